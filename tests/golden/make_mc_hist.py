#!/usr/bin/env python
"""Generates tests/golden/mc_hist_2v2.npz: the Monte Carlo score distribution of the UNMODIFIED reference GameSim on
BASELINE configs[1]'s workload shape (S0 randomized 2v2 games, 200 steps, winning_score = 4), 10^5 games.

The fp32 throughput kernel (K2) is validated against THIS distribution (north star: "validated statistically against
the reference's Monte Carlo score distributions"); tests do not recompute it.  The games are stepped by the plain-C
restatement (oracle/gamesim_oracle.c, 8 threads; it also records the event counters) and a 4 000-game sub-sample is
stepped by the compiled reference itself (oracle/_ref, /root/reference/src/GameSim.cpp in place): final states must
agree bit for bit.  Build container only (needs /root/reference); ~4 min on 8 cores.

    python tests/golden/make_mc_hist.py
"""
import ctypes as C
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import rgs_testlib as tl  # noqa: E402

N, STEPS, HOLD, SEED, WS = 100_000, 200, 10, 1, 4
REF_SAMPLE = 4000


def main():
    pd = tl.default_params(winning_score=WS)
    ev = np.zeros(tl.N_EVENTS, dtype=np.int64)
    scores = np.zeros((N, 2), dtype=np.int8)
    chunk = 10_000
    t0 = time.time()
    for g0 in range(0, N, chunk):
        x0, acts = tl.synth_inputs(SEED, g0, chunk, 2, STEPS // HOLD)
        o = tl.oracle_rollout(pd, tl.oracle_reset(pd, x0), acts, STEPS, HOLD, threads=8)
        scores[g0:g0 + chunk] = o["state"][:, :2].astype(np.int8)
        ev += o["events"].sum(0)
        if g0 == 0:
            lib = tl.ref_lib()
            assert lib is not None, "needs /root/reference (oracle/_ref)"
            dp = C.POINTER(C.c_double)
            out = np.zeros((REF_SAMPLE, 22))
            xs, as_ = np.ascontiguousarray(x0[:REF_SAMPLE]), np.ascontiguousarray(acts[:REF_SAMPLE])
            lib.ref_rollout_batch(REF_SAMPLE, STEPS, HOLD, pd["dt"], WS, xs.ctypes.data_as(dp), as_.ctypes.data_as(dp),
                                  STEPS // HOLD, pd["tau_puck"], pd["tau_player"], pd["player_mass"], pd["puck_mass"],
                                  out.ctypes.data_as(dp))
            assert np.array_equal(out, o["state"][:REF_SAMPLE]), "restatement != compiled reference"
        print(f"{g0 + chunk} games, {time.time() - t0:.0f} s", flush=True)
    # second pass, winning_score = 1: S0 games rarely see two goals in 200 steps, so the win / undecided split and the
    # time to the first goal are the informative part of the score distribution
    pd1 = tl.default_params(winning_score=1)
    wins1 = np.zeros(3, dtype=np.int64)
    decided_hist = np.zeros(STEPS // 20 + 1, dtype=np.int64)   # [k] = games decided in steps (20 k, 20 k + 20]; last = never
    for g0 in range(0, N, chunk):
        x0, acts = tl.synth_inputs(SEED, g0, chunk, 2, STEPS // HOLD)
        o = tl.oracle_rollout(pd1, tl.oracle_reset(pd1, x0), acts, STEPS, HOLD, threads=8)
        a, b = o["state"][:, 0] >= 1, o["state"][:, 1] >= 1
        wins1 += np.array([a.sum(), (b & ~a).sum(), (~a & ~b).sum()])
        ds = o["events"][:, tl.EV["DECIDED_STEP"]]
        np.add.at(decided_hist, np.where(ds > 0, (ds - 1) // 20, STEPS // 20), 1)
        print(f"ws=1: {g0 + chunk} games, {time.time() - t0:.0f} s", flush=True)
    hist = np.zeros((5, 5), dtype=np.int64)
    np.add.at(hist, (np.minimum(scores[:, 0], 4), np.minimum(scores[:, 1], 4)), 1)
    wins_a = int((scores[:, 0] >= WS).sum())
    wins_b = int(((scores[:, 1] >= WS) & (scores[:, 0] < WS)).sum())
    ev_names = np.array(list(tl.EV))
    ev[tl.EV["HASH"]] = 0   # a sum of hashes means nothing
    np.savez_compressed(os.path.join(HERE, "mc_hist_2v2.npz"), n=N, steps=STEPS, hold=HOLD, seed=SEED, winning_score=WS,
                        hist=hist, wins=np.array([wins_a, wins_b, N - wins_a - wins_b]), events=ev, event_names=ev_names,
                        scores=scores, a_leads=int((scores[:, 0] > scores[:, 1]).sum()),
                        b_leads=int((scores[:, 1] > scores[:, 0]).sum()), ref_sample_checked=REF_SAMPLE,
                        wins_ws1=wins1, decided_hist_ws1=decided_hist)
    print("wins ws=1", wins1, "decided hist", decided_hist)
    print("hist\n", hist, "\nwins", wins_a, wins_b, "events", dict(zip(ev_names.tolist(), ev.tolist())))


if __name__ == "__main__":
    main()
