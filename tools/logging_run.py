#!/usr/bin/env python
"""Throughput of a LOGGING run (north star: "achieved HBM GB/s on logging runs"): fp32 rollout with a trajectory
record (39 doubles = 312 B per game-step at 2v2: t, 22 state doubles, 16 damage entries) written every
`log_stride` steps.  Prints one JSON line.  GPU box only."""
import json
import sys

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from robo_game_sim_b200 import BatchedGameSim, synth_inputs  # noqa: E402

n, steps, hold = 200_000, 100, 10
out = {}
for stride in (0, 10, 1):
    sim = BatchedGameSim(n, mode="fast", winning_score=4)
    x0, acts = synth_inputs(n, steps // hold, seed=3, dtype=torch.float32)
    ms = []
    for rep in range(4):
        sim.reset(x0_ball=x0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        traj = sim.run(acts, n_steps=steps, hold=hold, log_stride=stride)
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    t = min(ms[1:]) * 1e-3
    nbytes = 0 if traj is None else traj.numel() * 8
    out[f"log_stride_{stride}"] = {"game_steps_per_s": n * steps / t, "ms": t * 1e3, "traj_bytes": nbytes,
                                   "traj_write_gbs": nbytes / t / 1e9}
    del traj
print(json.dumps(out))
