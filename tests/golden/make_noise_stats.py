#!/usr/bin/env python
"""Generates tests/golden/noise_stats.npz: statistics of the UNMODIFIED reference GameSim run WITH process noise
(reset(noise > 0, seed = run index), reference src/GameSim.cpp:30,41,62-68) -- what the K2 noise mode (Philox streams,
exact per-sub-step Gaussian increments) is validated against.  Bitwise parity is a non-goal (the reference's normal
stream is libstdc++'s std::normal_distribution over minstd_rand0), so the fixtures are sample moments and counts:

  free    : nobody touches anything (A1 commanded (3,-2), everything else at rest): per-step mean / variance of every
            player's state over 6000 runs x 20 steps, noise 2.0   -> the accumulated-noise covariance
  press   : A1 / A2 / B1 driven into the top / bottom / right wall, noise 5.0, 4000 runs x 40 steps: final-state
            moments and wall-event-free consistency (flagged steps: full noise paths, wall marches)
  s0      : contact-rich S0 games (random commands), noise 20.0, 6000 runs x 100 steps at winning_score 1:
            goals, decided fraction, final-state moments

Build container only (needs /root/reference -> oracle/_ref); ~3 min on 8 cores (one process per core).

    python tests/golden/make_noise_stats.py
"""
import ctypes as C
import os
import sys
from concurrent.futures import ProcessPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import rgs_testlib as tl  # noqa: E402


def _run(args):
    lo, hi, x0, acts, steps, hold, noise, ws, every = args
    lib = tl.ref_lib()
    dp = C.POINTER(C.c_double)
    h = lib.ref_create()
    n = hi - lo
    traj = np.zeros((n, steps, 22))
    for g in range(lo, hi):
        x = np.ascontiguousarray(x0 if x0.ndim == 1 else x0[g])
        a = acts if acts.ndim == 3 else acts[g]
        a8 = np.ascontiguousarray(np.repeat(a.reshape(a.shape[0], 8), hold, axis=0)[:steps])
        lib.ref_reset(h, 0.05, ws, x.ctypes.data_as(dp), noise, 0, b"", g + 1, 0.1, 0.5, 1.0, 0.5, None)
        lib.ref_rollout(h, a8.ctypes.data_as(dp), steps, traj[g - lo].ctypes.data_as(dp), None)
    lib.ref_destroy(h)
    return traj[:, every - 1::every]


def reference_runs(n, x0, acts, steps, hold, noise, ws=1 << 30, every=1, procs=8):
    cuts = [n * k // procs for k in range(procs + 1)]
    with ProcessPoolExecutor(procs) as ex:
        parts = list(ex.map(_run, [(cuts[k], cuts[k + 1], x0, acts, steps, hold, noise, ws, every) for k in range(procs)]))
    return np.concatenate(parts)


def moments(x):
    return x.mean(0), x.var(0, ddof=1)


def main():
    assert tl.ref_lib() is not None, "needs /root/reference (oracle/_ref)"
    out = {}
    # free flight
    n, steps, noise = 6000, 20, 2.0
    acts = np.zeros((1, 4, 2))
    acts[0, 0] = (3.0, -2.0)
    tr = reference_runs(n, np.zeros(4), acts, steps, steps, noise)
    m, v = moments(tr)
    out.update(free_n=n, free_steps=steps, free_noise=noise, free_actions=acts, free_mean=m, free_var=v)
    # pressed against the walls
    n, steps, noise = 4000, 40, 5.0
    acts = np.zeros((1, 4, 2))
    acts[0, 0], acts[0, 1], acts[0, 2] = (0.0, 10.0), (0.0, -10.0), (10.0, 0.0)
    tr = reference_runs(n, np.zeros(4), acts, steps, steps, noise, every=10)
    m, v = moments(tr)
    out.update(press_n=n, press_steps=steps, press_noise=noise, press_actions=acts, press_every=10, press_mean=m, press_var=v)
    # contact-rich S0 games
    n, steps, hold, noise, seed = 6000, 100, 10, 20.0, 31
    x0, a = tl.synth_inputs(seed, 0, n, 2, steps // hold)
    tr = reference_runs(n, x0, a, steps, hold, noise, ws=1, every=steps)[:, 0]
    tr0 = reference_runs(n, x0, a, steps, hold, 0.0, ws=1, every=steps)[:, 0]
    out.update(s0_n=n, s0_steps=steps, s0_hold=hold, s0_noise=noise, s0_seed=seed, s0_goals_a=int(tr[:, 0].sum()),
               s0_goals_b=int(tr[:, 1].sum()), s0_mean=tr[:, 2:].mean(0), s0_var=tr[:, 2:].var(0, ddof=1),
               s0_noise_free_goals=int(tr0[:, :2].sum()),
               s0_same_as_noise_free=float((np.abs(tr - tr0).max(1) < 1e-9).mean()),
               s0_rms_shift=float(np.sqrt(((tr[:, 2:] - tr0[:, 2:]) ** 2).mean())))
    np.savez_compressed(os.path.join(HERE, "noise_stats.npz"), **out)
    print("free sd(vx A1) per step:", np.sqrt(out["free_var"][:, 8]).round(5))
    print("press mean A1 py:", out["press_mean"][:, 7].round(4), "var", out["press_var"][:, 7].round(6))
    print("s0 goals", out["s0_goals_a"], out["s0_goals_b"], "noise-free", out["s0_noise_free_goals"], "rms shift", out["s0_rms_shift"])


if __name__ == "__main__":
    main()
