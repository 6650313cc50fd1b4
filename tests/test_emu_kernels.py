"""CPU suite, part 2: the PRODUCT's kernel bodies (robo_game_sim_b200/csrc/rgs_exact.cuh, rgs_fast.cuh)
compiled with g++ on the fiber warp emulator (tests/emu) and compared with the oracle.  This checks
the warp-level control flow (ballots, shuffles, lane roles) and the arithmetic of the very source
nvcc compiles for sm_100a; the -m gpu tests repeat the comparison on the real kernels."""
import numpy as np
import pytest

import rgs_testlib as tl
from conftest import GOLDEN_CASES


def _s0(n, steps, hold, n_team=2, seed=5, first=0, **kw):
    pd = tl.default_params(n_team=n_team, **kw)
    x0, acts = tl.synth_inputs(seed, first, n, n_team, steps // hold)
    return pd, tl.oracle_reset(pd, x0), acts


def _assert_same(o, e):
    assert np.array_equal(o["state"], e["state"])
    assert np.array_equal(o["events"], e["events"])
    assert np.array_equal(o["damage"], e["damage"])
    assert np.array_equal(o["t"], e["t"])


def test_exact_kernel_body_bit_exact_2v2():
    pd, s0, acts = _s0(24, 200, 10)
    o = tl.oracle_rollout(pd, s0, acts, 200, 10)
    e = tl.emu_rollout(pd, s0, acts, 200, 10)
    _assert_same(o, e)
    assert o["events"][:, 2:8].sum() > 500  # the sample does exercise collisions


@pytest.mark.parametrize("n_team", [1, 4, 7])
def test_exact_kernel_body_bit_exact_nvn(n_team):
    pd, s0, acts = _s0(4, 100, 10, n_team=n_team)
    o = tl.oracle_rollout(pd, s0, acts, 100, 10)
    e = tl.emu_rollout(pd, s0, acts, 100, 10)
    _assert_same(o, e)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_exact_kernel_body_golden(golden, name):
    pd = tl.default_params(winning_score=int(golden[name + "/winning_score"]), tau_puck=float(golden[name + "/tau_puck"]))
    s0 = tl.oracle_reset(pd, golden[name + "/x0"][None, :])
    acts = golden[name + "/actions"].reshape(-1, 4, 2)
    e = tl.emu_rollout(pd, s0, acts, acts.shape[0], 1, log_stride=1)
    assert np.array_equal(e["traj"][0, :, 1:23], golden[name + "/states"])
    assert e["t"][0] == float(golden[name + "/t"])
    assert np.array_equal(e["damage"][0], golden[name + "/damage"])
    assert np.array_equal(e["traj"][0, -1, 23:], golden[name + "/damage"])


def test_exact_resume_equals_single_launch():
    """A rollout boundary is a checkpoint: 2 x 50 steps == 100 steps (state, clock, damage, events)."""
    pd, s0, acts = _s0(6, 100, 10, seed=11)
    full = tl.emu_rollout(pd, s0, acts, 100, 10)
    a = tl.emu_rollout(pd, s0, acts[:, :5], 50, 10)
    b = tl.emu_rollout(pd, a["state"], acts[:, 5:], 50, 10, t=a["t"], damage=a["damage"], events=a["events"])
    _assert_same(full, b)


def test_on_chip_synth_commands_equal_tensor_commands():
    pd, s0, acts = _s0(4, 60, 10, seed=77, first=123)
    a = tl.emu_rollout(pd, s0, acts, 60, 10)
    b = tl.emu_rollout(pd, s0, None, 60, 10, synth=(77, 123, 0, 10.0))
    _assert_same(a, b)


def test_fast_formulation_fp64_tracks_oracle():
    """K2's closed-form formulation in fp64: event-for-event identical to the oracle except where the
    reference's own absolute-time rounding decides (jammed bodies, tau = 0 events)."""
    pd, s0, acts = _s0(32, 100, 10, seed=3)
    o = tl.oracle_rollout(pd, s0, acts, 100, 10)
    pf = dict(pd, mode=tl.MODE_FAST_F64)
    e = tl.emu_rollout(pf, s0, acts, 100, 10)
    same = o["events"][:, tl.EV["HASH"]] == e["events"][:, tl.EV["HASH"]]
    assert same.mean() >= 0.9
    assert np.array_equal(o["events"][same][:, :14], e["events"][same][:, :14])
    assert np.abs(o["state"] - e["state"])[same].max() < 1e-3   # rounding noise amplified by impacts
    assert np.median(np.abs(o["state"] - e["state"])[same].max(1)) < 1e-8


def test_fast_fp32_contact_free_and_first_contacts():
    pd, s0, acts = _s0(32, 20, 10, seed=3)
    o = tl.oracle_rollout(pd, s0, acts, 20, 10)
    pf = dict(pd, mode=tl.MODE_FAST_F32)
    e = tl.emu_rollout(pf, s0, acts, 20, 10)
    same = o["events"][:, tl.EV["HASH"]] == e["events"][:, tl.EV["HASH"]]
    assert same.mean() >= 0.8
    assert np.abs(o["state"] - e["state"])[same].max() < 5e-3
    assert np.array_equal(o["events"][:, tl.EV["LIVE_STEPS"]], e["events"][:, tl.EV["LIVE_STEPS"]])


def test_fast_kernel_goal_and_freeze(golden):
    """KAT1 in the fast formulation: same goal step, puck reset to exactly 0, frozen once decided."""
    for mode in (tl.MODE_FAST_F64, tl.MODE_FAST_F32):
        pd = tl.default_params(winning_score=1, mode=mode)
        s0 = tl.oracle_reset(pd, np.array([[4.0, 0.0, 5.0, 0.0]]))
        e = tl.emu_rollout(pd, s0, np.zeros((1, 4, 2)), 12, 12, log_stride=1)
        tr = e["traj"][0]
        assert tr[3, 1] == 0 and tr[4, 1] == 1.0 and np.all(tr[4, 3:7] == 0.0)
        assert e["events"][0, tl.EV["DECIDED_STEP"]] == 5 and e["events"][0, tl.EV["LIVE_STEPS"]] == 5
        assert e["t"][0] == pytest.approx(0.25)


# ---- K2 phased block kernel (rgs_phase.cuh): same bits as the one-lane-per-game body ----
MODE_PHASE_F32, MODE_PHASE_F64 = 34, 67   # emulator-only selectors (librgs_b200.so picks the kernel itself)


@pytest.mark.parametrize("geometry", [None, "64,4,8", "32,1,1", "7,3,2", "256,128,512"])
@pytest.mark.parametrize("n_team", [1, 2, 3])
def test_phase_kernel_equals_lane_kernel(monkeypatch, geometry, n_team):
    """Block-compacted schedule vs one lane per game: every state double, clock, damage, event record
    (incl. the order-sensitive event hash) and trajectory record identical, fp32 and fp64, also
    when the record pool runs dry (slots retry the step) and the item lists overflow (items searched
    on the spot), for any block size."""
    if geometry:
        monkeypatch.setenv("RGS_EMU_PHASE", geometry)
    n, steps, hold = (600 if n_team == 2 else 150), 120, 10
    x0, acts = tl.synth_inputs(5, 0, n, n_team, steps // hold)
    s0 = tl.oracle_reset(tl.default_params(n_team=n_team), x0)
    for lane_mode, phase_mode in ((tl.MODE_FAST_F32, MODE_PHASE_F32), (tl.MODE_FAST_F64, MODE_PHASE_F64)):
        a = tl.emu_rollout(tl.default_params(n_team=n_team, mode=lane_mode, winning_score=1), s0, acts, steps, hold, log_stride=30)
        b = tl.emu_rollout(tl.default_params(n_team=n_team, mode=phase_mode, winning_score=1), s0, acts, steps, hold, log_stride=30)
        _assert_same(a, b)
        assert np.array_equal(a["traj"], b["traj"])
        assert a["events"][:, 2:8].sum() > 500 and a["events"][:, tl.EV["DECIDED_STEP"]].any()


def test_phase_kernel_synth_and_resume():
    pd = tl.default_params(mode=MODE_PHASE_F32)
    x0, acts = tl.synth_inputs(77, 123, 40, 2, 6)
    s0 = tl.oracle_reset(tl.default_params(), x0)
    a = tl.emu_rollout(pd, s0, acts, 60, 10)
    b = tl.emu_rollout(pd, s0, None, 60, 10, synth=(77, 123, 0, 10.0))
    _assert_same(a, b)
    h = tl.emu_rollout(pd, s0, acts[:, :3], 30, 10)
    c = tl.emu_rollout(pd, h["state"], acts[:, 3:], 30, 10, t=h["t"], damage=h["damage"], events=h["events"])
    _assert_same(a, c)


@pytest.mark.parametrize("n_team,n,steps", [(2, 400, 120), (1, 100, 60), (4, 60, 40)])
def test_warp_scheduled_rollout_equals_lane_rollout(n_team, n, steps):
    """rollout_lane_warp (event loop cut at its iteration boundary, games loaded / stored inside the scheduling loop;
    A/B build -DRGS_LANE_DEB, emulator modes 35 / 68) executes the statements of rollout_lane_game per game: the
    results must be identical bit for bit, including the strided trajectory log and a resumed rollout."""
    x0, acts = tl.synth_inputs(5, 0, n, n_team, steps // 10)
    s0 = tl.oracle_reset(tl.default_params(n_team=n_team), x0)
    for lane_mode, warp_mode in ((tl.MODE_FAST_F32, 35), (tl.MODE_FAST_F64, 68)):
        a = tl.emu_rollout(tl.default_params(n_team=n_team, mode=lane_mode), s0, acts, steps, 10, log_stride=20)
        b = tl.emu_rollout(tl.default_params(n_team=n_team, mode=warp_mode), s0, acts, steps, 10, log_stride=20)
        _assert_same(a, b)
        assert np.array_equal(a["traj"], b["traj"])
        assert a["events"][:, 2:8].sum() > 50
