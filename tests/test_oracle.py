"""CPU suite, part 1: the oracle is pinned.

(a) oracle port (oracle/gamesim_oracle.c) == the UNMODIFIED reference compiled in place
    (oracle/_ref) bit for bit: 22 state doubles every step, clock and damage;
(b) oracle port == the committed golden fixtures (tests/golden, generated from (a)'s reference);
(c) the SURVEY 8(c) known-answer values and analytic closed forms.
"""
import ctypes as C

import numpy as np
import pytest

import rgs_testlib as tl
from conftest import GOLDEN_CASES


def _oracle_case(golden, name, traj=True):
    pd = tl.default_params(winning_score=int(golden[name + "/winning_score"]), tau_puck=float(golden[name + "/tau_puck"]))
    s0 = tl.oracle_reset(pd, golden[name + "/x0"][None, :])
    acts = golden[name + "/actions"].reshape(-1, 4, 2)
    return pd, s0, tl.oracle_rollout(pd, s0, acts, acts.shape[0], 1, traj=traj)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_oracle_matches_golden_cases(golden, name):
    pd, s0, o = _oracle_case(golden, name)
    assert np.array_equal(s0[0], golden[name + "/reset_state"])
    assert np.array_equal(o["traj"][0], golden[name + "/states"])          # every step, bit for bit
    assert o["t"][0] == float(golden[name + "/t"])
    assert np.array_equal(o["damage"][0], golden[name + "/damage"])


def test_oracle_matches_golden_s0(golden):
    pd = tl.default_params(winning_score=3)
    x0, acts = golden["s0/x0"], golden["s0/actions"]
    # the fixture inputs are the S0 generator's output for (seed, games 0..23)
    gx0, gacts = tl.synth_inputs(int(golden["s0/seed"]), 0, x0.shape[0], 2, acts.shape[1])
    assert np.array_equal(gx0, x0) and np.array_equal(gacts, acts)
    o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), acts, 400, 10, traj=True)
    assert np.array_equal(o["state"], golden["s0/final"])
    assert np.array_equal(o["t"], golden["s0/t"])
    assert np.array_equal(o["damage"], golden["s0/damage"])
    assert np.array_equal(o["traj"][:3], golden["s0/traj3"])


def test_known_answers(golden):
    """SURVEY 8(c) KAT1-KAT4 (values observed on the compiled reference)."""
    s = golden["kat1_goal/states"]
    assert s[0, 2] == 4.2379064549100676 and s[0, 4] == 4.5241870901798578
    assert s[3, 0] == 0 and s[4, 0] == 1.0 and np.all(s[4, 2:6] == 0.0)      # Team A scores during the 5th run
    s = golden["kat2_wall_up/states"]
    assert s[0, 3] == 2.2487541562707976 and s[0, 5] == 4.9502491687458408
    assert s[1, 5] == -4.900993366533779                                      # WALL_UP bounce
    s = golden["kat3_player_wall/states"]
    assert s[15, 9] == 3.6708038653407744 and s[16, 9] == -2.8437930934136637
    s = golden["kat4_head_on/states"]
    assert s[15, 9] == -3.6708038653407811 and s[16, 9] == 2.8437935470402627


def test_decided_game_is_frozen(golden):
    s = golden["decided_freeze/states"]
    k = int(np.argmax(s[:, 0] >= 1.0))
    assert np.all(s[k:] == s[k]) and float(golden["decided_freeze/t"]) == pytest.approx(0.05 * (k + 1))


def test_oracle_vs_compiled_reference_random_games():
    """Fresh randomized games (not the committed ones) against oracle/_ref, every step bit-exact."""
    lib = tl.ref_lib()
    if lib is None:
        pytest.skip("oracle/_ref not built and /root/reference absent")
    dp = C.POINTER(C.c_double)
    n, steps, hold = 40, 300, 10
    x0, acts = tl.synth_inputs(99, 1000, n, 2, steps // hold)
    for tau_puck, ws in ((0.1, 3), (1.0, 2)):
        pd = tl.default_params(winning_score=ws, tau_puck=tau_puck)
        o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), acts, steps, hold, traj=True)
        h = lib.ref_create()
        for g in range(n):
            s0 = np.zeros(22)
            lib.ref_reset(h, 0.05, ws, x0[g].ctypes.data_as(dp), 0.0, 0, b"", g, tau_puck, 0.5, 1.0, 0.5, s0.ctypes.data_as(dp))
            a = np.ascontiguousarray(np.repeat(acts[g].reshape(steps // hold, 8), hold, axis=0))
            st = np.zeros((steps, 22))
            lib.ref_rollout(h, a.ctypes.data_as(dp), steps, st.ctypes.data_as(dp), None)
            assert np.array_equal(st, o["traj"][g]), f"game {g}"
            assert lib.ref_time(h) == o["t"][g]
            d = np.zeros(16)
            lib.ref_damage(h, d.ctypes.data_as(dp))
            assert np.array_equal(d, o["damage"][g])
        lib.ref_destroy(h)


def test_analytic_free_motion():
    """Noise-free, contact-free motion: v(t) = u + (v0-u) exp(-t/tau_p); puck v0 exp(-0.2 t/tau_k)."""
    pd = tl.default_params(winning_score=3)
    s0 = tl.oracle_reset(pd, np.array([[0.0, 0.0, 1.0, -0.5]]))
    acts = np.zeros((1, 4, 2))
    acts[0, 0] = (1.0, 0.5)
    o = tl.oracle_rollout(pd, s0, acts, 10, 10)
    t = 0.5
    assert o["state"][0, 8] == pytest.approx(1.0 * (1 - np.exp(-t / 0.5)), abs=1e-12)
    assert o["state"][0, 4] == pytest.approx(1.0 * np.exp(-0.2 * t / 0.1), abs=1e-12)
    assert o["state"][0, 2] == pytest.approx((0.1 / 0.2) * (1 - np.exp(-0.2 * t / 0.1)), abs=1e-12)


def test_elastic_impulse_conserves_momentum_and_energy():
    """Head-on A1/A2 impact (KAT4 geometry): the impulse step conserves momentum and kinetic energy."""
    pd = tl.default_params(winning_score=3, tau_player=1e9)  # no drive: free flight between impacts
    s0 = tl.oracle_reset(pd, np.array([[4.0, 2.0, 0.0, 0.0]]))
    s0[0, 8:10] = (0.0, -3.0)
    s0[0, 12:14] = (0.0, 3.0)
    o = tl.oracle_rollout(pd, s0, np.zeros((1, 4, 2)), 20, 20, traj=True)
    assert o["events"][0, tl.EV["PLAYER_PLAYER"]] >= 1
    v = o["state"][0]
    assert v[9] + v[13] == pytest.approx(0.0, abs=1e-9)
    assert v[9] ** 2 + v[13] ** 2 == pytest.approx(18.0, rel=1e-6)


def test_log_record_layout():
    """29-double record (reference src/GameSim.cpp:161-170) incl. the contiguous-row quirk of utils.h:75."""
    lib = tl.oracle_lib()
    st = np.arange(22, dtype=np.float64)
    dmg = np.arange(16, dtype=np.float64) + 100  # row-major (aggressor, victim)
    out = np.zeros(29)
    dp = C.POINTER(C.c_double)
    lib.orc_log_record(st.ctypes.data_as(dp), 1.5, dmg.ctypes.data_as(dp), 0, out.ctypes.data_as(dp))
    assert out[0] == 1.5 and list(out[1:5]) == [0, 1, 2, 3] and list(out[5:7]) == [6, 7] and list(out[11:13]) == [18, 19]
    assert np.array_equal(out[13:], dmg)
    lib.orc_log_record(st.ctypes.data_as(dp), 1.5, dmg.ctypes.data_as(dp), 1, out.ctypes.data_as(dp))
    colmajor = dmg.reshape(4, 4).T.reshape(-1)  # the buffer Eigen holds
    for k in range(4):
        assert np.array_equal(out[13 + 4 * k:17 + 4 * k], colmajor[k:k + 4])
