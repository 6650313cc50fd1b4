#!/usr/bin/env python
"""tools/dump_fast.py OUT.npz [n_team] : state / clock / damage / events (+ a strided trajectory) of a fast-f32 S0 rollout
through whichever library RGS_B200_LIB names; tools/dump_fast.py --cmp A.npz B.npz compares two dumps bit for bit.
A/B builds of the K2 kernel must agree exactly (games are independent of how they are scheduled)."""
import sys

import numpy as np

if sys.argv[1] == "--cmp":
    a, b = np.load(sys.argv[2]), np.load(sys.argv[3])
    bad = [k for k in a.files if not np.array_equal(a[k], b[k], equal_nan=True)]
    print("dump_fast cmp:", "IDENTICAL" if not bad else f"DIFFER in {bad}", {k: a[k].shape for k in a.files})
    sys.exit(1 if bad else 0)

import torch  # noqa: E402

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from robo_game_sim_b200 import BatchedGameSim, synth_inputs  # noqa: E402

nt = int(sys.argv[2]) if len(sys.argv) > 2 else 2
out = {}
for tag, n, steps, stride in (("big", 150_001, 200, 0), ("log", 20_000, 100, 10), ("small", 4096, 50, 0)):
    sim = BatchedGameSim(n, mode="fast", players_per_team=nt, winning_score=3)
    x0, acts = synth_inputs(n, steps // 10, seed=7, players_per_team=nt, dtype=torch.float32)
    sim.reset(x0_ball=x0)
    traj = sim.run(acts, n_steps=steps, hold=10, log_stride=stride)
    out[tag + "_state"] = sim.state().cpu().numpy()
    out[tag + "_time"] = sim.time().cpu().numpy()
    out[tag + "_damage"] = sim.damage().cpu().numpy()
    out[tag + "_events"] = sim.events().cpu().numpy()
    if traj is not None:
        out[tag + "_traj"] = traj.cpu().numpy()
np.savez(sys.argv[1], **out)
print("dumped", sys.argv[1])
