"""GPU suite (-m gpu): the CUDA kernels, called through the C ABI (robo_game_sim_b200 -> librgs_b200.so),
against the oracle on the same seeded inputs, against the committed golden fixtures, and -- at
BASELINE.json's full sizes -- through size-independent properties.

Bars: fp64 parity mode (K1) bit-exact on every state double, clock, damage and event record;
fast fp64 (K2 formulation, absolute clock) event-identical to the oracle except where rounding decides a
resting contact (measured rates in DESIGN.md section 4); fast fp32 (K2) statistically equal to the reference's
Monte Carlo score distribution (committed fixture tests/golden/mc_hist_2v2.npz).
"""
import ctypes as C
import os

import numpy as np
import pytest

import rgs_testlib as tl
from conftest import GOLDEN_CASES

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def rgs():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import robo_game_sim_b200

    return robo_game_sim_b200


def _gpu_rollout(rgs, mode, pd, x0, acts, steps, hold, log_stride=0):
    n = x0.shape[0]
    kw = {k: pd[k] for k in ("dt", "winning_score", "tau_puck", "tau_player", "player_mass", "puck_mass")}
    sim = rgs.BatchedGameSim(n, mode=mode, players_per_team=pd["players_per_team"], **kw)
    sim.reset(x0_ball=torch.from_numpy(x0).cuda())
    a = torch.from_numpy(np.ascontiguousarray(acts)).to(sim.action_dtype).cuda()
    traj = sim.run(a, n_steps=steps, hold=hold, log_stride=log_stride)
    out = dict(state=sim.state().cpu().numpy(), t=sim.time().cpu().numpy(), damage=sim.damage().cpu().numpy().reshape(n, -1),
               events=sim.events().cpu().numpy(), stats=sim.stats().cpu().numpy(),
               traj=None if traj is None else traj.cpu().numpy())
    return sim, out


def _assert_bit_exact(o, g):
    assert np.array_equal(o["state"], g["state"])
    assert np.array_equal(o["events"], g["events"])
    assert np.array_equal(o["damage"], g["damage"])
    assert np.array_equal(o["t"], g["t"])


# ------------------------------------------------------------------ K1: fp64 parity mode
def test_exact_bit_exact_s0_2v2(rgs):
    """North star: "fp64 mode matching the reference's goal/score events bit-exactly over 10^4 randomized games"
    -- here every state double, clock, damage entry and event record of 10^4 S0 games x 200 steps."""
    n, steps, hold = 10_000, 200, 10
    pd = tl.default_params(winning_score=3)
    x0, acts = tl.synth_inputs(42, 0, n, 2, steps // hold)
    o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), acts, steps, hold)
    _, g = _gpu_rollout(rgs, "exact", pd, x0, acts, steps, hold)
    _assert_bit_exact(o, g)
    assert o["events"][:, 2:8].sum() > 1e5 and o["events"][:, :2].sum() > 50   # collisions and goals are exercised


@pytest.mark.parametrize("n_team,n,steps", [(1, 256, 100), (4, 512, 100), (7, 256, 60)])
def test_exact_bit_exact_nvn(rgs, n_team, n, steps):
    """BASELINE configs[3]: 4v4 and 7v7 (O(n^2) pair stress); oracle = the n-v-n restatement."""
    pd = tl.default_params(n_team=n_team, winning_score=3)
    x0, acts = tl.synth_inputs(43, 0, n, n_team, steps // 10)
    o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), acts, steps, 10)
    _, g = _gpu_rollout(rgs, "exact", pd, x0, acts, steps, 10)
    _assert_bit_exact(o, g)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_exact_golden_cases(rgs, golden, name):
    pd = tl.default_params(winning_score=int(golden[name + "/winning_score"]), tau_puck=float(golden[name + "/tau_puck"]))
    acts = golden[name + "/actions"].reshape(1, -1, 4, 2)
    _, g = _gpu_rollout(rgs, "exact", pd, golden[name + "/x0"][None, :], acts, acts.shape[1], 1, log_stride=1)
    assert np.array_equal(g["traj"][0, :, 1:23], golden[name + "/states"])
    assert g["t"][0] == float(golden[name + "/t"])
    assert np.array_equal(g["damage"][0], golden[name + "/damage"])


def test_exact_golden_s0(rgs, golden):
    pd = tl.default_params(winning_score=3)
    _, g = _gpu_rollout(rgs, "exact", pd, golden["s0/x0"], golden["s0/actions"], 400, 10, log_stride=1)
    assert np.array_equal(g["state"], golden["s0/final"])
    assert np.array_equal(g["t"], golden["s0/t"])
    assert np.array_equal(g["damage"], golden["s0/damage"])
    assert np.array_equal(g["traj"][:3, :, 1:23], golden["s0/traj3"])


def test_collision_cases_batched_1e5(rgs, golden):
    """BASELINE configs[2]: the bounce-kick / wall-contact / KAT cases replicated 10^5 x with
    per-replica jitter ~U(-1e-3,1e-3) on the puck start, 400 steps, fp64 parity mode.  Full size on
    the GPU; the oracle checks a strided sample of 2000 replicas bit for bit, the rest through
    properties (finite states, event totals consistent with the statistics vector)."""
    names = ["kat1_goal", "kat2_wall_up", "kat4_head_on", "bounce_kick_up", "bounce_kick_down", "wall_contacts"]
    n, steps = 100_000, 400
    rng = np.random.default_rng(7)
    for tau in (0.1, 1.0):
        sel = [nm for nm in names if float(golden[nm + "/tau_puck"]) == tau]
        case = rng.integers(0, len(sel), n)
        x0 = np.stack([golden[sel[c] + "/x0"] for c in case]) + np.concatenate(
            [rng.uniform(-1e-3, 1e-3, (n, 2)), np.zeros((n, 2))], axis=1)
        uniq = np.stack([np.resize(golden[nm + "/actions"], (steps, 8)) for nm in sel])     # [cases, steps, 8]
        pd = tl.default_params(winning_score=3, tau_puck=tau)
        # per-game command tensors are case-indexed: build [n, steps, 4, 2] lazily on the device
        sim = rgs.BatchedGameSim(n, mode="exact", winning_score=3, tau_puck=tau)
        sim.reset(x0_ball=torch.from_numpy(x0).cuda())
        acts_dev = torch.from_numpy(uniq).cuda()[torch.from_numpy(case).cuda()].reshape(n, steps, 4, 2).contiguous()
        sim.run(acts_dev, n_steps=steps, hold=1)
        st, ev, dmg, t = (x.cpu().numpy() for x in (sim.state(), sim.events(), sim.damage(), sim.time()))
        assert np.isfinite(st).all()
        idx = np.arange(0, n, n // 2000)
        o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0[idx]), uniq[case[idx]].reshape(len(idx), steps, 4, 2), steps, 1)
        assert np.array_equal(o["state"], st[idx]) and np.array_equal(o["events"], ev[idx])
        assert np.array_equal(o["damage"], dmg.reshape(n, -1)[idx]) and np.array_equal(o["t"], t[idx])
        stats = sim.stats().cpu().numpy()
        assert stats[0] == n and np.array_equal(stats[8:22], ev[:, :14].sum(0))


def test_drop_in_gamesim_matches_reference_step_by_step(rgs, tmp_path):
    """robo_game_py.GameSim (reset / priming run / run loop, as py/MonteCarlo.py:28-45 drives it) against
    the compiled reference (or the oracle port) including the 29-double binary log."""
    from robo_game_py import GameSim

    steps = 120
    x0, acts = tl.synth_inputs(5, 3, 1, 2, steps // 10)
    a_steps = np.repeat(acts[0].reshape(-1, 8), 10, axis=0)
    sim = GameSim()
    assert repr(sim) == "GameSim Object"
    logname = str(tmp_path / "mine.log")
    s = sim.reset(0.05, 3, x0[0], 0.0, True, logname, 1, 0.1, 0.5, 1.0, 0.5)
    assert s.dtype == np.float64 and s.shape == (22,)
    mine = [sim.run(np.zeros(2), np.zeros(2), np.zeros(2), np.zeros(2))]
    for k in range(steps):
        mine.append(sim.run(a_steps[k, 0:2], a_steps[k, 2:4], list(a_steps[k, 4:6]), a_steps[k, 6:8].reshape(2, 1)))
        assert sim.undecided() == bool(mine[-1][0] < 3 and mine[-1][1] < 3)
    sim.reset(0.05, 3, x0[0], 0.0, False, "")        # 6-argument form of run_sim.py:43; closes the log
    mine = np.array(mine)
    full = np.concatenate([np.zeros((1, 8)), a_steps])
    lib = tl.ref_lib()
    if lib is not None:
        dp = C.POINTER(C.c_double)
        h = lib.ref_create()
        ref_log = str(tmp_path / "ref.log").encode()
        lib.ref_reset(h, 0.05, 3, x0[0].ctypes.data_as(dp), 0.0, 1, ref_log, 1, 0.1, 0.5, 1.0, 0.5, None)
        st = np.zeros((steps + 1, 22))
        lib.ref_rollout(h, np.ascontiguousarray(full).ctypes.data_as(dp), steps + 1, st.ctypes.data_as(dp), None)
        lib.ref_destroy(h)       # flushes and closes the reference log
        assert np.array_equal(mine, st)
        assert open(logname, "rb").read() == open(ref_log.decode(), "rb").read()
    else:
        pd = tl.default_params(winning_score=3)
        o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), full.reshape(1, -1, 4, 2), steps + 1, 1, traj=True)
        assert np.array_equal(mine, o["traj"][0])
    rec = np.fromfile(logname).reshape(-1, 29)
    assert rec.shape[0] == steps + 1 and np.allclose(rec[:, 0], 0.05 * np.arange(1, steps + 2))
    assert np.array_equal(rec[:, 3:5], mine[:, 2:4])
    with pytest.raises(TypeError):
        sim.run(np.zeros(3), np.zeros(2), np.zeros(2), np.zeros(2))
    with pytest.raises(TypeError):
        sim.reset(0.05, 3)


@pytest.mark.parametrize("binding", ["ctypes", "pybind11"])
def test_drop_in_run_sim_config0_2000_steps(rgs, binding, tmp_path):
    """BASELINE configs[0]: py/run_sim.py's game (T = 100 s = 2000 steps, dt 0.05, winning_score 3,
    x0_ball = (-1, 0, 0, 0), run_sim.py:15-28, reset called with SIX arguments, :43) through the drop-in class --
    the repository's ctypes module and the pybind11 translation unit of INTEGRATION.md section B -- bit-exact against
    the compiled reference on every one of the 2001 returned 22-vectors (commands: seeded piece-wise constant)."""
    GameSim = __import__("robo_game_py").GameSim if binding == "ctypes" else tl.pybind_module().GameSim
    steps = 2000
    rng = np.random.default_rng(7)
    u = np.repeat(rng.uniform(-10, 10, size=(steps // 10, 8)), 10, axis=0)
    x0 = np.array([-1.0, 0.0, 0.0, 0.0])
    sim = GameSim()
    assert repr(sim) == "GameSim Object"
    s0 = sim.reset(0.05, 3, x0, 0.0, False, "~/gamelog.log")
    assert s0.shape == (22,) and s0.dtype == np.float64
    mine = [sim.run(np.zeros(2), np.zeros(2), np.zeros(2), np.zeros(2))]
    for k in range(steps):
        mine.append(sim.run(u[k, 0:2], u[k, 2:4], u[k, 4:6], u[k, 6:8]))
    mine = np.array(mine)
    full = np.concatenate([np.zeros((1, 8)), u])
    lib = tl.ref_lib()
    if lib is not None:
        dp = C.POINTER(C.c_double)
        h = lib.ref_create()
        lib.ref_reset(h, 0.05, 3, x0.ctypes.data_as(dp), 0.0, 0, b"", 0, 0.1, 0.5, 1.0, 0.5, None)
        st = np.zeros((steps + 1, 22))
        lib.ref_rollout(h, np.ascontiguousarray(full).ctypes.data_as(dp), steps + 1, st.ctypes.data_as(dp), None)
        lib.ref_destroy(h)
    else:
        pd = tl.default_params(winning_score=3)
        st = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0[None, :]), full.reshape(1, -1, 4, 2), steps + 1, 1, traj=True)["traj"][0]
    assert np.array_equal(mine, st)
    assert sim.undecided() == bool(st[-1, 0] < 3 and st[-1, 1] < 3)
    with pytest.raises(TypeError):
        sim.run(np.zeros(3), np.zeros(2), np.zeros(2), np.zeros(2))


def test_resume_synth_stats_and_host_entry(rgs):
    """Rollout boundary = checkpoint; on-chip S0 commands == tensor commands; K4 == numpy; the
    host-buffer entry point == the device-tensor entry point."""
    n, steps, hold = 3000, 100, 10
    pd = tl.default_params(winning_score=1)
    x0, acts = tl.synth_inputs(9, 50, n, 2, steps // hold)
    sim, g = _gpu_rollout(rgs, "exact", pd, x0, acts, steps, hold)
    # device S0 generator == numpy generator
    dx0, dacts = rgs.synth_inputs(n, steps // hold, seed=9, first_game=50)
    assert np.array_equal(dx0.cpu().numpy(), x0) and np.array_equal(dacts.cpu().numpy(), acts)
    # 2 x 50 steps with on-chip commands
    sim2 = rgs.BatchedGameSim(n, mode="exact", winning_score=1)
    sim2.reset(x0_ball=dx0)
    sim2.run_synth(9, 50, hold=hold, first_game=50, step0=0)
    sim2.run_synth(9, 50, hold=hold, first_game=50, step0=50)
    assert np.array_equal(sim2.state().cpu().numpy(), g["state"])
    assert np.array_equal(sim2.events().cpu().numpy(), g["events"])
    # restore from a full-state checkpoint
    sim3 = rgs.BatchedGameSim(n, mode="exact", winning_score=1)
    sim3.reset(state0=torch.from_numpy(tl.oracle_reset(pd, x0)).cuda())
    sim3.run(dacts, n_steps=steps, hold=hold)
    assert np.array_equal(sim3.state().cpu().numpy(), g["state"])
    # statistics vector
    st, ev, s = g["state"], g["events"], g["stats"]
    assert s[0] == n and s[1] == (st[:, 0] >= 1).sum() and s[2] == ((st[:, 1] >= 1) & (st[:, 0] < 1)).sum()
    assert s[3] == ((st[:, 0] < 1) & (st[:, 1] < 1)).sum() and s[1] + s[2] + s[3] == n
    assert np.array_equal(s[8:22], ev[:, :14].sum(0))
    hist = np.zeros((5, 5), dtype=np.int64)
    np.add.at(hist, (np.minimum(st[:, 0], 4).astype(int), np.minimum(st[:, 1], 4).astype(int)), 1)
    assert np.array_equal(s[32:57].reshape(5, 5), hist)
    # host-buffer entry point
    hs, he, hst = sim.run_host(x0, acts, steps, hold=hold)
    assert np.array_equal(hs, g["state"]) and np.array_equal(he, g["events"]) and np.array_equal(hst, g["stats"])
    # frozen games do not advance their clock
    dec = ev[:, tl.EV["DECIDED_STEP"]] > 0
    assert dec.any() and np.allclose(g["t"][dec], 0.05 * ev[dec, tl.EV["DECIDED_STEP"]])


@pytest.mark.parametrize("mode", ["fast", "exact"])
def test_host_entry_chunked_equals_device_entry(rgs, mode):
    """rgs_batch_run_host splits batches of >= 65 536 games into game-range chunks over two private streams (copies
    overlapped with the rollouts); 70 001 games is ragged against every chunk bound.  Same bits as reset + run on
    device tensors, statistics included; then once more on the same batch object (staging buffers reused)."""
    n, steps, hold = 70_001, 40, 10
    dx0, dacts = rgs.synth_inputs(n, steps // hold, seed=17)
    sim = rgs.BatchedGameSim(n, mode=mode, winning_score=2)
    dacts = dacts.to(sim.action_dtype)
    sim.reset(x0_ball=dx0)
    sim.run(dacts, n_steps=steps, hold=hold)
    want = (sim.state().cpu().numpy(), sim.events().cpu().numpy(), sim.stats().cpu().numpy())
    host = rgs.BatchedGameSim(n, mode=mode, winning_score=2)
    hx0, hacts = dx0.cpu().pin_memory(), dacts.cpu().pin_memory()
    for _ in range(2):
        hs, he, hst = host.run_host(hx0, hacts, steps, hold=hold)
        assert np.array_equal(hs, want[0]) and np.array_equal(he, want[1]) and np.array_equal(hst, want[2])
    assert int(want[1][:, 2:8].sum()) > n // 10     # the games did collide


# ------------------------------------------------------------------ K2: throughput formulation
def test_fast64_tracks_oracle(rgs):
    """K2's formulation in fp64 against the UNMODIFIED oracle (absolute clock) over 200 steps.  Measured
    (tools/k2_fidelity.py): 99.8 % of the games event-hash-identical; the residue is rounding
    deciding a resting contact.  On the hash-identical games every integer outcome is equal."""
    n, steps, hold = 4096, 200, 10
    pd = tl.default_params(winning_score=3)
    x0, acts = tl.synth_inputs(42, 0, n, 2, steps // hold)
    o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), acts, steps, hold)
    _, g = _gpu_rollout(rgs, "fast64", pd, x0, acts, steps, hold)
    same = o["events"][:, 14] == g["events"][:, 14]
    assert same.mean() >= 0.99
    assert np.array_equal(o["events"][same][:, :14], g["events"][same][:, :14])
    # (the number of correctOverlap fixes is not part of the hash: an overlap of a few ulps is a fix on one side only)
    assert abs(int(o["events"][same][:, 15].sum()) - int(g["events"][same][:, 15].sum())) <= 0.01 * o["events"][same][:, 15].sum()
    assert np.array_equal(o["state"][same][:, :2], g["state"][same][:, :2])
    assert np.median(np.abs(o["state"] - g["state"])[same].max(1)) < 1e-8
    assert np.median(np.abs(o["damage"] - g["damage"])[same].max(1)) < 1e-8
    # and the emulator build of the same source gives the same bits as the GPU
    e = tl.emu_rollout(dict(pd, mode=tl.MODE_FAST_F64), tl.oracle_reset(pd, x0[:64]), acts[:64], steps, hold)
    assert np.array_equal(e["state"], g["state"][:64]) and np.array_equal(e["events"], g["events"][:64])


def _gpu_single_steps(rgs, mode, pd, before, acts1):
    """One game-step from every state record of `before` [m, 2+4E] with commands acts1 [m, 1, 2n, 2]."""
    m = before.shape[0]
    kw = {k: pd[k] for k in ("dt", "winning_score", "tau_puck", "tau_player", "player_mass", "puck_mass")}
    sim = rgs.BatchedGameSim(m, mode=mode, players_per_team=pd["players_per_team"], **kw)
    sim.reset(state0=torch.from_numpy(before).cuda())
    sim.run(torch.from_numpy(np.ascontiguousarray(acts1)).to(sim.action_dtype).cuda(), n_steps=1, hold=1)
    return dict(state=sim.state().cpu().numpy(), events=sim.events().cpu().numpy())


@pytest.mark.parametrize("n_team,dt,tau_puck,tau_player,pm,km", [
    (2, 0.02, 0.1, 0.5, 1.0, 0.5), (2, 0.05, 0.1, 0.5, 1.0, 0.5), (2, 0.05, 1.0, 0.5, 1.0, 0.5), (2, 0.1, 0.1, 0.5, 1.0, 0.5),
    (2, 0.2, 0.1, 0.5, 1.0, 0.5), (2, 0.2, 1.0, 0.2, 2.0, 0.1), (1, 0.1, 0.1, 0.5, 1.0, 0.5), (3, 0.1, 0.1, 0.5, 1.0, 0.5),
    (4, 0.1, 0.1, 0.5, 1.0, 0.5), (7, 0.1, 0.1, 0.5, 1.0, 0.5)])
def test_fast64_single_step_sweep(rgs, n_team, dt, tau_puck, tau_player, pm, km):
    """Every state of an oracle rollout stepped ONCE by the oracle and by K2 (fp64) on the GPU, clock 0 on both sides,
    swept over dt, tau, masses and team size (round-1 advisor finding: K2 departed on ordinary wall contacts as dt
    grew).  Bar: event-hash mismatch rate < 1e-3, no hash-identical step with a state error, identical event records
    (incl. the number of correctOverlap fixes) on the hash-identical steps."""
    n, steps, hold = (300, 40, 10) if n_team <= 2 else (100, 40, 10)
    pd = tl.default_params(n_team=n_team, winning_score=1 << 30, dt=dt, tau_puck=tau_puck, tau_player=tau_player,
                           player_mass=pm, puck_mass=km)
    x0, acts = tl.synth_inputs(42, 0, n, n_team, steps // hold)
    s0 = tl.oracle_reset(pd, x0)
    o = tl.oracle_rollout(pd, s0, acts, steps, hold, traj=True)
    before = np.ascontiguousarray(np.concatenate([s0[:, None], o["traj"][:, :-1]], axis=1).reshape(n * steps, -1))
    a1 = np.ascontiguousarray(np.repeat(acts, hold, axis=1)[:, :steps].reshape(n * steps, 1, 2 * n_team, 2))
    ro = tl.oracle_rollout(pd, before, a1, 1, 1)
    rg = _gpu_single_steps(rgs, "fast64", pd, before, a1)
    same = ro["events"][:, 14] == rg["events"][:, 14]
    err = np.abs(ro["state"] - rg["state"]).max(1)
    assert (~same).sum() <= 1e-3 * n * steps
    assert not (same & (err > 1e-6)).any()
    assert np.array_equal(ro["events"][same], rg["events"][same])
    # fp32 on the same steps: per-type event counts within 2 % (+ 3 sigma) of the oracle's
    rf = _gpu_single_steps(rgs, "fast", pd, before, a1)
    eo, ef = ro["events"].sum(0).astype(float), rf["events"].sum(0).astype(float)
    for k in ("WALL_UP", "WALL_DOWN", "WALL_RIGHT", "WALL_LEFT", "PLAYER_PLAYER", "PLAYER_PUCK"):
        assert abs(ef[tl.EV[k]] - eo[tl.EV[k]]) <= 0.02 * eo[tl.EV[k]] + 3 * np.sqrt(eo[tl.EV[k]] + 1), (k, eo[tl.EV[k]], ef[tl.EV[k]])


@pytest.mark.parametrize("n_team,n,steps", [(1, 4000, 100), (3, 2000, 100), (4, 2000, 100), (7, 2000, 100)])
def test_fast_nvn_tracks_oracle(rgs, n_team, n, steps):
    """K2 for team sizes other than the reference's 2 against the n-v-n oracle (the plain-C restatement; the
    reference itself is 2v2 only): fp64 -- share of event-hash-identical games, integer outcomes equal on those, per-type
    event rates on all; fp32 -- per-type event rates and goal totals."""
    hold = 10
    pd = tl.default_params(n_team=n_team, winning_score=3)
    x0, acts = tl.synth_inputs(7, 0, n, n_team, steps // hold)
    o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), acts, steps, hold)
    _, g = _gpu_rollout(rgs, "fast64", pd, x0, acts, steps, hold)
    same = o["events"][:, 14] == g["events"][:, 14]
    assert same.mean() >= 0.97, same.mean()
    assert np.array_equal(o["events"][same][:, :14], g["events"][same][:, :14])
    assert np.array_equal(o["state"][same][:, :2], g["state"][same][:, :2])
    # rounding differences are amplified by every impact: 1e-9 after 100 steps at 1v1, 1e-6 at 7v7 (measured)
    assert np.median(np.abs(o["state"] - g["state"])[same].max(1)) < 1e-5
    _, f = _gpu_rollout(rgs, "fast", pd, x0, acts.astype(np.float32), steps, hold)
    eo = o["events"].sum(0).astype(float)
    for name, r in (("fast64", g), ("fast", f)):
        er = r["events"].sum(0).astype(float)
        for k in ("WALL_UP", "WALL_DOWN", "WALL_RIGHT", "WALL_LEFT", "PLAYER_PLAYER", "PLAYER_PUCK"):
            tol = 0.02 if name == "fast64" else 0.03
            assert abs(er[tl.EV[k]] - eo[tl.EV[k]]) <= tol * eo[tl.EV[k]] + 3 * np.sqrt(eo[tl.EV[k]] + 1), (name, k, eo[tl.EV[k]], er[tl.EV[k]])
        goals_o, goals_r = eo[0] + eo[1], er[0] + er[1]
        assert abs(goals_r - goals_o) < 5 * np.sqrt(goals_o + 1)
        assert np.array_equal(o["events"][:, tl.EV["LIVE_STEPS"]] == steps, r["events"][:, tl.EV["LIVE_STEPS"]] == steps) or \
            abs((o["events"][:, 10] == steps).mean() - (r["events"][:, 10] == steps).mean()) < 0.01


def _chi2_two_sample(a, b):
    """Two-sample chi-square statistic of homogeneity and its degrees of freedom (cells empty in both dropped)."""
    a, b = np.asarray(a, float).ravel(), np.asarray(b, float).ravel()
    keep = (a + b) > 0
    a, b = a[keep], b[keep]
    ka, kb = np.sqrt(b.sum() / a.sum()), np.sqrt(a.sum() / b.sum())
    return float((((ka * a - kb * b) ** 2) / (a + b)).sum()), int(keep.sum() - 1)


def test_fast32_score_distribution_vs_reference(rgs):
    """North star: "an fp32 throughput mode is validated statistically against the reference's Monte Carlo score
    distributions".  Fixture tests/golden/mc_hist_2v2.npz = the UNMODIFIED reference's 5x5 score histogram, wins and
    event totals on 10^5 S0 games x 200 steps at winning_score 4 (tests/golden/make_mc_hist.py).
    (a) the SAME 10^5 games on the GPU in fp32: K4's histogram cell by cell within the paired-sampling noise of the
        games whose score differs, event totals per type within 1 %;
    (b) 10^6 OTHER games (different seed) on the GPU: two-sample chi-square of the 5x5 histogram and of wins
        A / B / undecided against the reference sample, p > 0.001."""
    from scipy.stats import chi2

    fx = np.load(os.path.join(os.path.dirname(__file__), "golden", "mc_hist_2v2.npz"))
    n, steps, hold, seed, ws = int(fx["n"]), int(fx["steps"]), int(fx["hold"]), int(fx["seed"]), int(fx["winning_score"])
    sim = rgs.BatchedGameSim(n, mode="fast", winning_score=ws)
    x0, acts = rgs.synth_inputs(n, steps // hold, seed=seed, dtype=torch.float32)
    sim.reset(x0_ball=x0)
    sim.run(acts, n_steps=steps, hold=hold)
    stats = sim.stats().cpu().numpy()
    hist_g = stats[32:57].reshape(5, 5)
    sc_g = sim.state()[:, :2].cpu().numpy().astype(np.int8)
    sc_o = fx["scores"]
    differ = int((sc_g != sc_o).any(1).sum())
    assert differ < 0.02 * n, differ                       # measured: see DESIGN.md section 4
    # a game whose score differs moves one count out of one cell and into another: |delta cell| <= differ, and for
    # unbiased differences ~ sqrt(differ)
    d = hist_g - fx["hist"]
    assert np.abs(d).max() <= 4 * np.sqrt(differ + 1), (d, differ)
    assert abs(int(stats[1]) - int(fx["wins"][0])) <= 4 * np.sqrt(differ + 1)
    assert abs(int(stats[2]) - int(fx["wins"][1])) <= 4 * np.sqrt(differ + 1)
    ev_g, ev_o = stats[8:22].astype(float), fx["events"][:14].astype(float)
    # Impact counts are dominated by RESTING contacts, whose event-loop length the reference decides by the rounding
    # of its absolute fp64 clock (base_idx, src/GameSim.cpp:237; run on a step-relative clock the reference itself
    # counts 5 % fewer impacts).  fp32 event times are step-relative, but resolve() takes that one decision in fp64 on
    # the reference's clock, so impacts track the reference too: 3 % here (chaotic games, different roundings).
    tol = {"PLAYER_PLAYER": 0.03, "PLAYER_PUCK": 0.03}
    for k in ("WALL_UP", "WALL_DOWN", "WALL_RIGHT", "WALL_LEFT", "PLAYER_PLAYER", "PLAYER_PUCK", "GOALS_A", "GOALS_B"):
        assert abs(ev_g[tl.EV[k]] - ev_o[tl.EV[k]]) <= tol.get(k, 0.01) * ev_o[tl.EV[k]] + 3 * np.sqrt(ev_o[tl.EV[k]]), (k, ev_o[tl.EV[k]], ev_g[tl.EV[k]])
    # (b) independent sample, winning_score 4 (the 5x5 histogram) and 1 (wins A / B / undecided, time to the first goal:
    # S0 games rarely see two goals in 200 steps, so the decided split only exists at winning_score 1)
    m = 1_000_000
    sim2 = rgs.BatchedGameSim(m, mode="fast", winning_score=ws)
    x0, _ = rgs.synth_inputs(m, 0, seed=seed + 77, dtype=torch.float32)
    sim2.reset(x0_ball=x0)
    sim2.run_synth(seed + 77, steps, hold=hold)
    st2 = sim2.stats().cpu().numpy()
    c2, dof = _chi2_two_sample(fx["hist"], st2[32:57].reshape(5, 5))
    assert dof >= 3 and chi2.sf(c2, dof) > 1e-3, (c2, dof)
    sim3 = rgs.BatchedGameSim(m, mode="fast", winning_score=1)
    sim3.reset(x0_ball=x0)
    sim3.run_synth(seed + 77, steps, hold=hold)
    st3 = sim3.stats().cpu().numpy()
    c2w, dofw = _chi2_two_sample(fx["wins_ws1"], st3[1:4])
    assert dofw == 2 and chi2.sf(c2w, dofw) > 1e-3, (c2w, dofw, fx["wins_ws1"], st3[1:4])
    ds = sim3.events()[:, tl.EV["DECIDED_STEP"]].cpu().numpy()
    dh = np.bincount(np.where(ds > 0, (ds - 1) // 20, steps // 20), minlength=steps // 20 + 1)
    c2d, dofd = _chi2_two_sample(fx["decided_hist_ws1"], dh)
    assert chi2.sf(c2d, dofd) > 1e-3, (c2d, dofd)


def test_fast32_full_size_properties(rgs):
    """BASELINE configs[1] size on one GPU (10^6 games x 200 steps): determinism, independence of the
    batch composition (a game's result does not depend on which other games share the launch),
    y-mirror symmetry of the physics, and event totals consistent with the statistics vector."""
    n, steps, hold = 1_000_000, 200, 10
    sim = rgs.BatchedGameSim(n, mode="fast", winning_score=4)
    x0, acts = rgs.synth_inputs(n, steps // hold, seed=11, dtype=torch.float32)
    sim.reset(x0_ball=x0)
    sim.run(acts, n_steps=steps, hold=hold)
    st, ev, stats = sim.state(), sim.events(), sim.stats()
    assert torch.isfinite(st).all()
    assert int(stats[0]) == n and torch.equal(stats[8:22], ev[:, :14].sum(0))
    assert int(ev[:, 10].min()) > 0 and int(ev[:, 10].max()) == steps
    sim.reset(x0_ball=x0)
    sim.run(acts, n_steps=steps, hold=hold)
    assert torch.equal(sim.state(), st) and torch.equal(sim.events(), ev)            # deterministic
    lo, cnt = 123_457, 10_000
    sub = rgs.BatchedGameSim(cnt, mode="fast", winning_score=4)
    sub.reset(x0_ball=x0[lo:lo + cnt].contiguous())
    sub.run(acts[lo:lo + cnt].contiguous(), n_steps=steps, hold=hold)
    assert torch.equal(sub.state(), st[lo:lo + cnt]) and torch.equal(sub.events(), ev[lo:lo + cnt])
    # mirror y -> -y of puck start and of every command: players start at +-1.25, so A1<->A2, B1<->B2 swap roles
    m = 100_000
    mx0 = x0[:m].clone()
    mx0[:, 1] *= -1
    mx0[:, 3] *= -1
    ma = acts[:m].clone()[:, :, [1, 0, 3, 2], :]
    ma[..., 1] *= -1
    mir = rgs.BatchedGameSim(m, mode="fast", winning_score=4)
    mir.reset(x0_ball=mx0)
    mir.run(ma.contiguous(), n_steps=steps, hold=hold)
    mev, mst = mir.events(), mir.state()
    same_goals = (mev[:, :2] == ev[:m, :2]).all(1).float().mean().item()
    assert same_goals > 0.98            # identical up to body-order tie-breaks in jams
    assert abs(int(mev[:, 2].sum()) - int(ev[:m, 3].sum())) / max(1, int(ev[:m, 3].sum())) < 0.02   # UP <-> DOWN


def test_monte_carlo_batched_harness(rgs, tmp_path):
    import types

    mc = types.SimpleNamespace(num_runs=5000, T=10.0, dt=0.05, winning_score=4, x0_ball=np.array([-1.0, 0, 0, 0]),
                               noise_stdev=0.0, tau_puck=0.1, tau_player=0.5, player_mass=1.0, puck_mass=0.5,
                               log_prefix=str(tmp_path / "mc"))
    out = rgs.monte_carlo_batched(mc, mode="exact", seed=3)
    assert out["games"] == 5000 and out["n_steps"] == 200
    assert out["live_steps"] <= 5000 * 201 and out["live_steps"] > 5000 * 150
    assert len(open(os.path.join(mc.log_prefix, "configuration.txt")).read().splitlines()) == 13
    # deterministic open-loop commands shared by all runs: every run is the same game (noise = 0)
    acts = np.zeros((20, 4, 2))
    acts[:, 0] = (5.0, -3.0)
    out2 = rgs.monte_carlo_batched(mc, actions=acts, mode="exact", hold=10)
    pd = tl.default_params(winning_score=4)
    full = np.concatenate([np.zeros((1, 4, 2)), np.repeat(acts, 10, axis=0)])
    o = tl.oracle_rollout(pd, tl.oracle_reset(pd, mc.x0_ball[None, :]), full, 201, 1)
    assert out2["player_puck"] == 5000 * o["events"][0, 7] and out2["wall_up"] == 5000 * o["events"][0, 2]
    # closed loop through a Python policy (host in the loop, like the reference harness)
    mc.num_runs, mc.T = 64, 1.0
    out3 = rgs.monte_carlo_batched(mc, policy=lambda s: torch.zeros((s.shape[0], 4, 2), device=s.device), mode="fast")
    assert out3["games"] == 64 and out3["live_steps"] == 64 * 21


def _reference_logs(tmp_path, x0, full_actions, ws=3):
    """mc_run-style binary logs written by the compiled reference itself, one game per file -> list of bytes."""
    lib = tl.ref_lib()
    dp = C.POINTER(C.c_double)
    out = []
    for g in range(x0.shape[0]):
        h = lib.ref_create()
        name = str(tmp_path / f"ref_{g}.log").encode()
        lib.ref_reset(h, 0.05, ws, np.ascontiguousarray(x0[g]).ctypes.data_as(dp), 0.0, 1, name, g + 1, 0.1, 0.5, 1.0, 0.5, None)
        a = np.ascontiguousarray(full_actions[g].reshape(-1, 8))
        lib.ref_rollout(h, a.ctypes.data_as(dp), a.shape[0], None, None)
        lib.ref_destroy(h)
        out.append(open(name.decode(), "rb").read())
    return out


def test_batched_log_writer_matches_reference_files(rgs, tmp_path):
    """BatchedGameSim.write_logs -> rgs_write_logs: 16 games logged every step by the batched path give the SAME
    BYTES as the compiled reference's own Logger for each game (29-double records, src/GameSim.cpp:161-170, incl. the
    damage.row(k) quirk of include/utils.h:75, and a decided game's log stopping at its last live step); the
    layout flag changes only the damage fields 14..29 (true rows)."""
    if tl.ref_lib() is None:
        pytest.skip("needs the compiled reference (oracle/_ref)")
    n, steps, hold, ws = 16, 120, 10, 1
    x0, acts = tl.synth_inputs(9, 0, n, 2, steps // hold)
    x0[0] = (4.0, 0.0, 5.0, 0.0)                      # KAT1: decided at step 5 -> a short log
    acts[1, :, 0] = (0.0, -10.0)                      # KAT4-like head-on impact: damage fields are non-zero
    acts[1, :, 1] = (0.0, 10.0)
    full = np.repeat(acts, hold, axis=1)             # [n, steps, 4, 2]
    ref = _reference_logs(tmp_path, x0, full, ws=ws)
    sim = rgs.BatchedGameSim(n, mode="exact", winning_score=ws)
    sim.reset(x0_ball=torch.from_numpy(x0).cuda())
    traj = sim.run(torch.from_numpy(acts).cuda(), n_steps=steps, hold=hold, log_stride=1)
    live = sim.events()[:, tl.EV["LIVE_STEPS"]]
    sim.write_logs(traj, str(tmp_path / "mine_{}.log"), first_index=1, layout="reference", n_records=live)
    for g in range(n):
        mine = open(tmp_path / f"mine_{g + 1}.log", "rb").read()
        assert mine == ref[g], g
    assert len(ref[0]) == 5 * 29 * 8 and len(ref[2]) == steps * 29 * 8
    sim.write_logs(traj, str(tmp_path / "rm_{}.log"), first_index=1, layout="row_major", n_records=live)
    a = np.fromfile(tmp_path / "mine_2.log").reshape(-1, 29)
    b = np.fromfile(tmp_path / "rm_2.log").reshape(-1, 29)
    assert np.array_equal(a[:, :13], b[:, :13]) and not np.array_equal(a[:, 13:], b[:, 13:])
    dm = sim.damage()[1].cpu().numpy()              # (aggressor, victim)
    assert dm[0, 1] > 0 and np.array_equal(b[-1, 13:].reshape(4, 4), dm)
    # append mode: two launches of 60 steps give the same files as one of 120
    sim.reset(x0_ball=torch.from_numpy(x0).cuda())
    t1 = sim.run(torch.from_numpy(acts[:, :6]).cuda(), n_steps=60, hold=hold, log_stride=1)
    l1 = sim.events()[:, tl.EV["LIVE_STEPS"]]
    sim.write_logs(t1, str(tmp_path / "two_{}.log"), n_records=l1)
    t2 = sim.run(torch.from_numpy(np.ascontiguousarray(acts[:, 6:])).cuda(), n_steps=60, hold=hold, log_stride=1)
    sim.write_logs(t2, str(tmp_path / "two_{}.log"), n_records=sim.events()[:, tl.EV["LIVE_STEPS"]] - l1, append=True)
    for g in range(n):
        assert open(tmp_path / f"two_{g + 1}.log", "rb").read() == ref[g]
    with pytest.raises(ValueError):
        sim.write_logs(traj, str(tmp_path / "no_placeholder.log"))


@pytest.mark.parametrize("mode", ["fast", "fast64", "exact"])
def test_trajectory_field_major_layout_equals_records(rgs, mode):
    """rgs_batch_set_traj_layout: the field-major (SoA) trajectory tensor [n_rec, 39, N] of a run holds exactly the
    records tensor [N, n_rec, 39] (ragged batch: 1 237 games is neither a multiple of 32 nor of 128; some games decide
    early, so the frozen-record path is covered; strides 1 and 4; also the on-chip command generator's entry point)."""
    n, steps, hold = 1237, 48, 10
    x0, acts = rgs.synth_inputs(n, 5, seed=21)
    for stride in (1, 4):
        sims = [rgs.BatchedGameSim(n, mode=mode, winning_score=1) for _ in range(2)]
        out = []
        for sim, layout in zip(sims, ("records", "field_major")):
            sim.reset(x0_ball=x0)
            out.append(sim.run(acts.to(sim.action_dtype), n_steps=steps, hold=hold, log_stride=stride, traj_layout=layout))
        assert out[0].shape == (n, steps // stride, 39) and out[1].shape == (steps // stride, 39, n)
        assert torch.equal(out[1].permute(2, 0, 1), out[0])
        assert torch.equal(sims[0].state(), sims[1].state())
        assert int((sims[0].state()[:, :2].sum(1) >= 1).sum()) > 0    # some games were decided and frozen
        # the layout is a property of the call: the same batch object switches back
        sims[1].reset(x0_ball=x0)
        back = sims[1].run_synth(3, steps, hold=hold, log_stride=stride)
        sims[0].reset(x0_ball=x0)
        fm = sims[0].run_synth(3, steps, hold=hold, log_stride=stride, traj_layout="field_major")
        assert torch.equal(fm.permute(2, 0, 1), back)
    with pytest.raises(ValueError):
        sims[0].run(acts.to(sims[0].action_dtype), n_steps=steps, hold=hold, log_stride=1, traj_layout="soa")


@pytest.mark.parametrize("mode", ["exact", "fast"])
def test_checkpoint_and_restore_resume_exactly(rgs, mode):
    """rgs_batch_get / rgs_batch_set: a batch restored from a checkpoint into ANOTHER batch object and carried on gives
    the bits of the uninterrupted run -- state, clock, damage and event records (hash included); reset(state0=...)
    alone would have zeroed the last three."""
    n, hold = 3000, 10
    x0, acts = rgs.synth_inputs(n, 10, seed=5)
    a = rgs.BatchedGameSim(n, mode=mode, winning_score=3)
    acts = acts.to(a.action_dtype)
    a.reset(x0_ball=x0)
    a.run(acts, n_steps=100, hold=hold)
    b = rgs.BatchedGameSim(n, mode=mode, winning_score=3)
    b.reset(x0_ball=x0)
    b.run(acts[:, :6], n_steps=60, hold=hold)
    ck = b.checkpoint()
    assert float(ck["time"].min()) > 0 and int(ck["events"][:, 2:8].sum()) > 0 and float(ck["damage"].sum()) > 0
    c = rgs.BatchedGameSim(n, mode=mode, winning_score=3)
    c.restore(**ck)
    c.run(acts[:, 6:].contiguous(), n_steps=40, hold=hold)
    for k, v in a.checkpoint().items():
        assert torch.equal(v, c.checkpoint()[k]), k
    with pytest.raises(ValueError):
        c.restore(state=ck["state"][:10])


def test_damage_statistics_on_device(rgs):
    """K4 sums the damage matrix over the batch in fixed point (one integer all-reduce vector): equal to the sum of
    the per-game matrices to the fixed-point resolution, in exact and fast mode."""
    n, steps, hold = 20000, 100, 10
    for mode in ("exact", "fast"):
        sim = rgs.BatchedGameSim(n, mode=mode, winning_score=4)
        x0, acts = rgs.synth_inputs(n, steps // hold, seed=13, dtype=sim.action_dtype)
        sim.reset(x0_ball=x0)
        sim.run(acts, n_steps=steps, hold=hold)
        d = rgs.stats_to_dict(sim.stats())
        dm = sim.damage().cpu().numpy()
        assert dm.sum() > 1.0
        assert np.abs(d["damage_sum"] - dm.sum(0)).max() <= n * 2.0 ** -24
        assert np.all(np.diag(d["damage_sum"]) == 0)
        team = np.array([[dm[:, :2, :2].sum(), dm[:, :2, 2:].sum()], [dm[:, 2:, :2].sum(), dm[:, 2:, 2:].sum()]])
        assert np.abs(d["damage_team_sum"] - team).max() <= 4 * n * 2.0 ** -24
        assert d["overlap_fixes"] == int(sim.events()[:, 15].sum())


def test_monte_carlo_reference_signature(rgs, tmp_path):
    """robo_game_sim_b200.monte_carlo(mc_params, team_A_run_method, team_B_run_method, cbf_run_method=None) -- the
    reference's signature (py/MonteCarlo.py:8) and artefacts.  Sequential (reference order, per-game SimState views),
    lock-step batched and on-device paths; the logs of the first two are the compiled reference's bytes."""
    import types

    mc = types.SimpleNamespace(num_runs=3, T=2.0, dt=0.05, winning_score=4, x0_ball=np.array([-1.0, 0.5, 2.0, 1.0]),
                               noise_stdev=0.0, tau_puck=0.1, tau_player=0.5, player_mass=1.0, puck_mass=0.5,
                               log_prefix=str(tmp_path / "seq"), input_limit=10.0, safety_radius=0.4, barrier_gain=30)

    def chase(team):        # a stateless controller: both players head for the puck
        def run(sim_state):
            assert isinstance(sim_state, rgs.SimState)
            pk = sim_state.get_puck_pos()
            return tuple(np.clip(4.0 * (pk - sim_state.get_player_pos(team, k)), -10, 10) for k in (1, 2))
        return run

    def damp(u_nominal, sim_state):   # a stand-in for CBF.get_centralized_safe_control...: scales the commands
        return [0.5 * u for u in u_nominal]

    out = rgs.monte_carlo(mc, chase("A"), chase("B"), damp, verbose=False)
    assert out["games"] == 3 and sorted(os.listdir(mc.log_prefix)) == ["configuration.txt", "mc_run_1.log", "mc_run_2.log", "mc_run_3.log"]
    assert len(open(os.path.join(mc.log_prefix, "configuration.txt")).read().splitlines()) == 13
    seq = [open(os.path.join(mc.log_prefix, f"mc_run_{i}.log"), "rb").read() for i in (1, 2, 3)]
    assert len(seq[0]) == 41 * 29 * 8 and seq[0] == seq[1] == seq[2]          # noise = 0: the runs are one game
    # the same loop against the compiled reference
    lib = tl.ref_lib()
    if lib is not None:
        dp = C.POINTER(C.c_double)
        h = lib.ref_create()
        name = str(tmp_path / "ref.log").encode()
        st = np.zeros(22)
        lib.ref_reset(h, mc.dt, 4, mc.x0_ball.ctypes.data_as(dp), 0.0, 1, name, 1, 0.1, 0.5, 1.0, 0.5, None)
        lib.ref_run(h, np.zeros(8).ctypes.data_as(dp), st.ctypes.data_as(dp))
        a, b = chase("A"), chase("B")
        for _ in range(40):
            ss = rgs.SimState(st)
            u = np.concatenate(damp([*a(ss), *b(ss)], ss))
            lib.ref_run(h, np.ascontiguousarray(u).ctypes.data_as(dp), st.ctypes.data_as(dp))
        lib.ref_destroy(h)
        assert open(name.decode(), "rb").read() == seq[0]
    # lock-step batched path: same controllers on [n, 22] views, exact mode -> the same files
    mc.log_prefix = str(tmp_path / "bat")
    out_b = rgs.monte_carlo(mc, chase("A"), chase("B"), damp, batched=True, verbose=False)
    assert [open(os.path.join(mc.log_prefix, f"mc_run_{i}.log"), "rb").read() for i in (1, 2, 3)] == seq
    for k in ("games", "wins_a", "wins_b", "lead_a", "lead_b"):
        assert out_b[k] == out[k]
    assert np.allclose(out_b["damage_sum"], out["damage_sum"], atol=1e-6)
    # on-device teams: one launch per Monte Carlo run set, logs of n_steps + 1 records
    mc.log_prefix, mc.num_runs = str(tmp_path / "dev"), 8
    out_d = rgs.monte_carlo(mc, "classical", "classical", "cbf", verbose=False)
    assert out_d["games"] == 8
    rec = np.fromfile(os.path.join(mc.log_prefix, "mc_run_8.log")).reshape(-1, 29)
    assert rec.shape[0] == 41 and np.allclose(rec[:, 0], 0.05 * np.arange(1, 42)) and np.isfinite(rec).all()
    with pytest.raises(ValueError):
        rgs.monte_carlo(mc, "classical", "dmpc", verbose=False)


@pytest.mark.parametrize("mode", ["fast64", "fast"])
def test_noise_mode_matches_reference_statistics(rgs, mode):
    """Process noise (reference src/GameSim.cpp:62-68; reset(noise, seed)) on the GPU: the same statistical checks as
    the CPU suite at the fixture's full sample sizes (tests/golden/noise_stats.npz = the reference run with noise > 0)."""
    fx = tl.noise_fixture()
    cols = list(range(6, 22))

    def run(n, x0, acts, steps, hold, noise, ws=1 << 30, log=1):
        sim = rgs.BatchedGameSim(n, mode=mode, winning_score=ws, noise=noise, seed=3)
        sim.reset(x0_ball=torch.as_tensor(x0).cuda())
        a = torch.as_tensor(np.ascontiguousarray(acts)).to(sim.action_dtype).cuda()
        traj = sim.run(a, n_steps=steps, hold=hold, log_stride=log)
        return sim, (None if traj is None else traj.cpu().numpy())

    n, steps, noise = int(fx["free_n"]), int(fx["free_steps"]), float(fx["free_noise"])
    _, tr = run(n, np.zeros(4), fx["free_actions"], steps, steps, noise)
    tr = tr[:, :, 1:23]
    tl.check_noise_moments(tr.mean(0), tr.var(0, ddof=1), n, fx["free_mean"], fx["free_var"], n, cols, 0.12, "free")
    vv, pp, _ = tl.analytic_noise_var(noise, steps)
    assert abs(tr[:, -1, 8].var(ddof=1) / vv - 1) < 0.1 and abs(tr[:, -1, 6].var(ddof=1) / pp - 1) < 0.1
    n, steps, noise, every = int(fx["press_n"]), int(fx["press_steps"]), float(fx["press_noise"]), int(fx["press_every"])
    sim, tr = run(n, np.zeros(4), fx["press_actions"], steps, steps, noise, log=every)
    tr = tr[:, :, 1:23]
    tl.check_noise_moments(tr.mean(0), tr.var(0, ddof=1), n, fx["press_mean"], fx["press_var"], n, cols[:12], 0.15, "press")
    assert int(sim.events()[:, tl.EV["WALL_UP"]].min()) >= 1
    n, steps, hold = int(fx["s0_n"]), int(fx["s0_steps"]), int(fx["s0_hold"])
    x0, a = tl.synth_inputs(int(fx["s0_seed"]), 0, n, 2, steps // hold)
    sim, _ = run(n, x0, a, steps, hold, float(fx["s0_noise"]), ws=1, log=0)
    st = sim.state().cpu().numpy()
    tl.check_noise_moments(st[:, 2:].mean(0), st[:, 2:].var(0, ddof=1), n, fx["s0_mean"], fx["s0_var"], n, list(range(4, 20)), 0.15, "s0")
    goals, ref_goals = st[:, :2].sum(), int(fx["s0_goals_a"]) + int(fx["s0_goals_b"])
    assert abs(goals - ref_goals) < 4.5 * np.sqrt(goals + ref_goals + 1)


def test_noise_through_the_drop_in_and_the_harness(rgs, tmp_path):
    """reset(noise > 0) works through robo_game_py.GameSim (the batch of one switches to the K2 fp64 noise kernel) and
    through the Monte Carlo harness; streams are keyed by (seed, global game index): reproducible, seed-dependent and
    independent of how a batch is sharded."""
    import types

    from robo_game_py import GameSim

    sim = GameSim()
    x0 = np.array([-1.0, 0.0, 0.0, 0.0])
    z = np.zeros(2)

    def play(seed, noise):
        sim.reset(0.05, 3, x0, noise, False, "", seed)
        return np.array([sim.run(np.array([3.0, -2.0]), z, z, z) for _ in range(30)])

    a, b, c, quiet = play(1, 0.5), play(1, 0.5), play(2, 0.5), play(1, 0.0)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert 1e-5 < np.abs(a - quiet).max() < 0.05 and np.allclose(a[:5, 2:6], quiet[:5, 2:6], atol=1e-9)   # puck: no noise of its own
    # sharding invariance: games 0..99 as one batch == two batches of 50 with game offsets 0 and 50
    n, steps, hold = 100, 60, 10
    xs, acts = tl.synth_inputs(5, 0, n, 2, steps // hold)
    kw = dict(mode="fast", winning_score=3, noise=8.0, seed=9)
    full = rgs.BatchedGameSim(n, **kw)
    full.reset(x0_ball=torch.from_numpy(xs).cuda())
    full.run(torch.from_numpy(acts).float().cuda(), n_steps=steps, hold=hold)
    for lo in (0, 50):
        part = rgs.BatchedGameSim(50, first_game=lo, **kw)
        part.reset(x0_ball=torch.from_numpy(xs[lo:lo + 50]).cuda())
        part.run(torch.from_numpy(acts[lo:lo + 50]).float().cuda(), n_steps=steps, hold=hold)
        assert torch.equal(part.state(), full.state()[lo:lo + 50]) and torch.equal(part.events(), full.events()[lo:lo + 50])
    with pytest.raises(NotImplementedError):
        rgs.BatchedGameSim(4, mode="exact", noise=0.1)
    mc = types.SimpleNamespace(num_runs=2000, T=2.0, dt=0.05, winning_score=4, x0_ball=np.array([-1.0, 0.5, 2.0, 1.0]),
                               noise_stdev=0.05, tau_puck=0.1, tau_player=0.5, player_mass=1.0, puck_mass=0.5,
                               log_prefix=str(tmp_path / "mc"))
    out = rgs.monte_carlo_batched(mc, policy="classical", mode="fast")
    assert out["games"] == 2000 and out["live_steps"] > 2000 * 20
    mc.num_runs = 2
    out2 = rgs.monte_carlo(mc, lambda s: (z, z), lambda s: (z, z), verbose=False)
    assert out2["games"] == 2 and os.path.exists(os.path.join(mc.log_prefix, "mc_run_2.log"))
    r1 = np.fromfile(os.path.join(mc.log_prefix, "mc_run_1.log")).reshape(-1, 29)
    r2 = np.fromfile(os.path.join(mc.log_prefix, "mc_run_2.log")).reshape(-1, 29)
    assert r1.shape == (41, 29) and not np.array_equal(r1, r2)        # seed = run index (py/MonteCarlo.py:28-30)


# ------------------------------------------------------------------ K3: shared-state rollout sampling
def test_rollout_sampler_matches_oracle(rgs):
    """BASELINE configs[4]: 4096 candidate command sequences x 50 steps from ONE state.  fp64 parity mode:
    every candidate's final state equals the oracle's bit for bit, so costs agree to rounding and the
    arg-min is the oracle's; fp32 mode: the device arg-min is the arg-min of the device costs and beats
    the do-nothing candidate."""
    from robo_game_sim_b200 import sampler as sp

    C, steps, hold = 4096, 50, 10
    pd = tl.default_params(winning_score=3)
    x0, warm = tl.synth_inputs(9, 0, 1, 2, 4)
    s0 = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), warm, 40, 10)["state"][0]     # a mid-game state
    acts = sp.candidate_actions(C, steps // hold, players=(0, 1), seed=1, dtype=np.float64)
    play = sp.dmpc_play(s0, field=-1)
    x_des = {0: sp.dmpc_target(s0, play, sp.ATTACKER, -1, [5.0, 0.0], [-5.0, 0.0]),
             1: sp.dmpc_target(s0, play, sp.DEFENDER, -1, [5.0, 0.0], [-5.0, 0.0])}
    o = tl.oracle_rollout(pd, np.repeat(s0[None], C, 0), acts, steps, hold)["state"]
    w = np.array(sp.OMEGA_N_MAX)
    ocost = sum(((o[:, 6 + 4 * p:10 + 4 * p] - x_des[p]) ** 2 * w).sum(1) for p in (0, 1))

    sim = rgs.BatchedGameSim(C, mode="exact", winning_score=3)
    best, first, cost = sp.sample_rollouts(sim, s0, acts, x_des, n_steps=steps, hold=hold)
    assert np.array_equal(sim.state().cpu().numpy(), o)
    assert np.allclose(cost.cpu().numpy(), ocost, rtol=1e-12, atol=0)
    assert best == int(np.argmin(ocost)) and np.array_equal(first, acts[best, 0, :2])

    fast = rgs.BatchedGameSim(C, mode="fast", winning_score=3)
    fb, _, fcost = sp.sample_rollouts(fast, s0, acts.astype(np.float32), x_des, n_steps=steps, hold=hold)
    assert fb == int(torch.argmin(fcost)) and float(fcost[fb]) <= float(fcost[0])
    assert np.abs(fcost.cpu().numpy() - ocost).max() / ocost.max() < 0.05 or fb == best


@pytest.mark.parametrize("n", [0, 1, 33, 18943, 18945])
def test_ragged_batch_sizes_and_cta_size_switch(rgs, n):
    """Empty, single-game and ragged batches, and both sides of the batch size at which the lane kernel switches
    from one-warp CTAs to 128-lane CTAs (148 SMs x 128): the first and last games equal the emulator build of the
    same source bit for bit (fp32), and the whole batch equals the oracle in fp64 parity mode (small n)."""
    steps, hold = 60, 10
    pd = tl.default_params(winning_score=2)
    if n == 0:
        sim = rgs.BatchedGameSim(0, mode="fast")
        sim.reset()
        sim.run(torch.zeros((0, 1, 4, 2), dtype=torch.float32, device="cuda"), n_steps=steps)
        assert sim.state().shape == (0, 22) and int(sim.stats()[0]) == 0
        return
    x0, acts = tl.synth_inputs(31, 0, n, 2, steps // hold)
    _, g = _gpu_rollout(rgs, "fast", pd, x0, acts.astype(np.float32), steps, hold)
    pick = np.unique(np.r_[np.arange(min(n, 40)), np.arange(max(0, n - 40), n)])
    e = tl.emu_rollout(dict(pd, mode=tl.MODE_FAST_F32), tl.oracle_reset(pd, x0[pick]), acts[pick], steps, hold)
    assert np.array_equal(e["state"], g["state"][pick]) and np.array_equal(e["events"], g["events"][pick])
    assert np.array_equal(e["damage"], g["damage"][pick]) and np.array_equal(e["t"], g["t"][pick])
    assert int(g["stats"][0]) == n and g["events"][:, tl.EV["LIVE_STEPS"]].max() == steps
    if n <= 33:
        o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), acts, steps, hold)
        _, x = _gpu_rollout(rgs, "exact", pd, x0, acts, steps, hold)
        _assert_bit_exact(o, x)


# ------------------------------------------------------------------ §8(f) rank 4: on-device surrogate policy
def test_on_device_policy_closed_loop(rgs, tmp_path):
    """The closed loop runs inside the K2 kernel.  Without the barrier filter the policy is add / mul / div / sqrt
    only: the GPU must equal the emulator build of the same source bit for bit (fp32).  With the filter (one expf per
    approaching pair, not bit-identical between CUDA and glibc) the first steps agree to rounding and the behaviour
    agrees: the filter removes the player-player impacts.  The Monte Carlo harness drives it with policy="classical"."""
    n, steps = 3000, 200
    pd = tl.default_params(winning_score=3, mode=tl.MODE_FAST_F32)
    x0, _ = tl.synth_inputs(31, 0, n, 2, 1)
    kw = {k: pd[k] for k in ("dt", "winning_score", "tau_puck", "tau_player", "player_mass", "puck_mass")}
    sim = rgs.BatchedGameSim(n, mode="fast", **kw)

    def gpu(n_steps, **pol):
        sim.reset(x0_ball=torch.from_numpy(x0).cuda())
        traj = sim.run_policy(n_steps, log_stride=n_steps, **pol)
        return dict(state=sim.state().cpu().numpy(), events=sim.events().cpu().numpy(), t=sim.time().cpu().numpy(),
                    damage=sim.damage().cpu().numpy().reshape(n, -1), traj=traj.cpu().numpy())

    s0 = tl.oracle_reset(tl.default_params(), x0)
    g = gpu(steps)
    sub = slice(0, 400)
    e = tl.emu_rollout(pd, s0[sub], None, steps, policy=tl.default_policy())
    assert np.array_equal(g["state"][sub], e["state"]) and np.array_equal(g["events"][sub], e["events"])
    assert np.array_equal(g["damage"][sub], e["damage"]) and np.array_equal(g["t"][sub], e["t"])
    assert np.array_equal(g["traj"][:, 0, 1:23], g["state"])          # the strided log holds the final state
    assert g["state"][:, :2].sum() > 100 and g["events"][:, 7].sum() > 10 * n   # games are played: goals, kicks
    gd = gpu(steps, kind=1)                                           # DMPC play logic and targets
    ed = tl.emu_rollout(pd, s0[sub], None, steps, policy=tl.default_policy(kind=1))
    assert np.array_equal(gd["state"][sub], ed["state"]) and np.array_equal(gd["events"][sub], ed["events"])
    assert not np.array_equal(gd["state"], g["state"])
    # barrier filter
    g3 = gpu(3, cbf=1)
    e3 = tl.emu_rollout(pd, s0[sub], None, 3, policy=tl.default_policy(cbf=1))
    same = (g3["events"][sub, 2:8] == e3["events"][:, 2:8]).all(axis=1)
    assert same.mean() > 0.99 and np.abs(g3["state"][sub][same] - e3["state"][same]).max() < 1e-3
    gs = gpu(steps, cbf=1)
    assert gs["events"][:, 6].sum() < 0.1 * g["events"][:, 6].sum()
    # error behaviour: exact mode and other team sizes are refused, loudly
    with pytest.raises(RuntimeError, match="on-device policy"):
        rgs.BatchedGameSim(8, mode="exact").reset(x0_ball=np.zeros(4)).run_policy(5)
    with pytest.raises(RuntimeError, match="on-device policy"):
        rgs.BatchedGameSim(8, mode="fast", players_per_team=3, noise=0.5).reset(x0_ball=np.zeros(4)).run_policy(5)
    # other team sizes (goalie + n-1 field players): GPU == emulator build, bit for bit, fp32 without the filter
    for nt in (1, 3, 4):
        xs, _ = tl.synth_inputs(33, 0, 256, nt, 1)
        s3 = rgs.BatchedGameSim(256, mode="fast", players_per_team=nt, winning_score=3)
        s3.reset(x0_ball=torch.from_numpy(xs).cuda())
        s3.run_policy(100)
        pdn = tl.default_params(n_team=nt, winning_score=3, mode=tl.MODE_FAST_F32)
        en = tl.emu_rollout(pdn, tl.oracle_reset(pdn, xs), None, 100, policy=tl.default_policy())
        assert np.array_equal(s3.state().cpu().numpy(), en["state"]) and np.array_equal(s3.events().cpu().numpy(), en["events"])
        assert en["events"][:, tl.EV["PLAYER_PUCK"]].sum() > 0
    with pytest.raises(TypeError):
        sim.run_policy(5, kick=1.0)
    # the Monte Carlo harness, both teams on the device-side baseline policy (py/mc_example_1.py)
    import types

    mc = types.SimpleNamespace(num_runs=20000, T=10.0, dt=0.05, winning_score=4, x0_ball=np.array([-1.0, 0, 0, 0]),
                               noise_stdev=0.0, tau_puck=0.1, tau_player=0.5, player_mass=1.0, puck_mass=0.5,
                               log_prefix=str(tmp_path / "mcp"))
    free = rgs.monte_carlo_batched(mc, policy="classical", randomize=True, seed=5)
    safe = rgs.monte_carlo_batched(mc, policy="classical+cbf", randomize=True, seed=5)
    assert free["games"] == 20000 and free["n_steps"] == 200 and free["goals_a"] + free["goals_b"] > 500
    assert 0.7 < free["goals_a"] / free["goals_b"] < 1.4            # mirror-symmetric teams, symmetric S0 pucks
    assert safe["player_player"] < 0.1 * free["player_player"]
