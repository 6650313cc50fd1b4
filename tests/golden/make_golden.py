#!/usr/bin/env python
"""Generates tests/golden/gamesim_golden.npz from the UNMODIFIED reference GameSim
(/root/reference/src/GameSim.cpp compiled in place by oracle/Makefile -> oracle/_ref/libref_gamesim.so).

The reference ships no golden vectors of its own (its only "test", py/test_collisions.py, replays
command files that were never committed), so these fixtures ARE the pin: every case stores the
inputs and the reference's per-step 22-double state, clock and damage matrix.  Run in the build
container only (needs /root/reference); the .npz is committed and travels to the GPU box.

    python tests/golden/make_golden.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import rgs_testlib as tl  # noqa: E402


def run_reference(x0_ball, actions, tau_puck=0.1, winning_score=3, dt=0.05):
    """actions [steps, 8] -> (states [steps,22], t, damage[16], reset_state[22])"""
    lib = tl.ref_lib()
    assert lib is not None, "needs /root/reference (oracle/_ref)"
    dp = C.POINTER(C.c_double)
    h = lib.ref_create()
    x0 = np.ascontiguousarray(x0_ball, dtype=np.float64)
    s0 = np.zeros(22)
    lib.ref_reset(h, dt, winning_score, x0.ctypes.data_as(dp), 0.0, 0, b"", 0, tau_puck, 0.5, 1.0, 0.5, s0.ctypes.data_as(dp))
    acts = np.ascontiguousarray(actions, dtype=np.float64)
    states = np.zeros((acts.shape[0], 22))
    lib.ref_rollout(h, acts.ctypes.data_as(dp), acts.shape[0], states.ctypes.data_as(dp), None)
    t = lib.ref_time(h)
    dmg = np.zeros(16)
    lib.ref_damage(h, dmg.ctypes.data_as(dp))
    lib.ref_destroy(h)
    return states, t, dmg, s0


def const_actions(steps, a1=(0, 0), a2=(0, 0), b1=(0, 0), b2=(0, 0)):
    return np.tile(np.array([*a1, *a2, *b1, *b2], dtype=np.float64), (steps, 1))


def main():
    out = {}
    cases = {
        # SURVEY 8(c) known-answer cases
        "kat1_goal": dict(x0=(4, 0, 5, 0), tau_puck=0.1, acts=const_actions(10)),
        "kat2_wall_up": dict(x0=(0, 2, 1, 5), tau_puck=1.0, acts=const_actions(40)),
        "kat3_player_wall": dict(x0=(0, 0, 0, 0), tau_puck=0.1, acts=const_actions(40, a1=(0, 10))),
        "kat4_head_on": dict(x0=(0, 0, 0, 0), tau_puck=0.1, acts=const_actions(40, a1=(0, -10), a2=(0, 10))),
        # py/test_collisions.py scenarios (its command files are not in the repo): bounce-kick up / down
        # with tau_puck = 1.0 (py/run_sim.py:20), puck kicked by A1 / A2 to rebound off the top / bottom wall
        "bounce_kick_up": dict(x0=(-2.0, 1.6, 0, 0), tau_puck=1.0, acts=const_actions(400, a1=(10, 7), b1=(-3, 0))),
        "bounce_kick_down": dict(x0=(-2.0, -1.6, 0, 0), tau_puck=1.0, acts=const_actions(400, a2=(10, -7), b2=(-3, 0))),
        # puck shot straight at each wall, and a corner jam of three players
        "wall_contacts": dict(x0=(0, 0, -6, 6.5), tau_puck=1.0, acts=const_actions(400, a1=(-10, 10), a2=(-10, 10), b1=(10, -10), b2=(-10, 10))),
        # a full game to a decision: frozen state afterwards (GameSim::run no-op, :121-123)
        "decided_freeze": dict(x0=(4, 0, 5, 0), tau_puck=1.0, winning_score=1, acts=const_actions(30)),
    }
    for name, c in cases.items():
        st, t, dmg, s0 = run_reference(c["x0"], c["acts"], c.get("tau_puck", 0.1), c.get("winning_score", 3))
        out[name + "/x0"] = np.array(c["x0"], dtype=np.float64)
        out[name + "/tau_puck"] = np.float64(c.get("tau_puck", 0.1))
        out[name + "/winning_score"] = np.int64(c.get("winning_score", 3))
        out[name + "/actions"] = c["acts"]
        out[name + "/states"] = st
        out[name + "/t"] = np.float64(t)
        out[name + "/damage"] = dmg
        out[name + "/reset_state"] = s0
        print(f"{name}: final score {st[-1, 0]:.0f}-{st[-1, 1]:.0f}, t={t:.2f}")

    # randomized S0 games (the bench / parity workload): 24 games x 400 steps, commands held 10 steps
    n, steps, hold = 24, 400, 10
    x0, acts = tl.synth_inputs(2026, 0, n, 2, steps // hold)
    finals, ts, dmgs = np.zeros((n, 22)), np.zeros(n), np.zeros((n, 16))
    trajs = np.zeros((3, steps, 22))
    for g in range(n):
        a = np.repeat(acts[g].reshape(steps // hold, 8), hold, axis=0)
        st, t, dmg, _ = run_reference(x0[g], a, 0.1, 3)
        finals[g], ts[g], dmgs[g] = st[-1], t, dmg
        if g < 3:
            trajs[g] = st
    out["s0/seed"] = np.int64(2026)
    out["s0/x0"] = x0
    out["s0/actions"] = acts
    out["s0/final"] = finals
    out["s0/t"] = ts
    out["s0/damage"] = dmgs
    out["s0/traj3"] = trajs
    print("s0: goals", finals[:, :2].sum(0), "damage sum", dmgs.sum())
    np.savez_compressed(os.path.join(HERE, "gamesim_golden.npz"), **out)


if __name__ == "__main__":
    main()
