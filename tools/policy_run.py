#!/usr/bin/env python
"""Closed-loop Monte Carlo on the on-device surrogate policy (10^6 2v2 games x 200 steps, S0 pucks), with and without
the barrier filter, against the same loop stepped from the host with a torch policy of one kernel launch per step
(the reference's structure, py/MonteCarlo.py:35-46).  Prints one JSON line.  GPU box only."""
import json
import sys

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from robo_game_sim_b200 import BatchedGameSim, synth_inputs  # noqa: E402
from robo_game_sim_b200 import _capi  # noqa: E402

n, steps = 1_000_000, 200
sim = BatchedGameSim(n, mode="fast", winning_score=4)
x0, _ = synth_inputs(n, 1, seed=9, dtype=torch.float32)
out = {"games": n, "steps": steps}
for tag, pol in (("policy", {}), ("policy_cbf", {"cbf": 1})):
    ms = []
    for rep in range(3):
        sim.reset(x0_ball=x0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        sim.run_policy(steps, **pol)
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    st = sim.stats().cpu().numpy()
    live = int(st[_capi.ST_EVENTS0 + 10])
    out[tag] = {"ms": min(ms), "game_steps_per_s": live / (min(ms) * 1e-3), "live_steps": live,
                "goals": [int(st[_capi.ST_EVENTS0]), int(st[_capi.ST_EVENTS0 + 1])],
                "player_player": int(st[_capi.ST_EVENTS0 + 6]), "player_puck": int(st[_capi.ST_EVENTS0 + 7]),
                "loop_iters_per_step": float(st[_capi.ST_EVENTS0 + 9]) / max(1, live)}
# host in the loop: zero-command torch policy, one launch per game-step (lower bound on the reference structure's cost)
m = 100_000
sim2 = BatchedGameSim(m, mode="fast", winning_score=4)
sim2.reset(x0_ball=x0[:m])
zero = torch.zeros((m, 1, 4, 2), dtype=torch.float32, device="cuda")
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(50):
    s = sim2.state()
    sim2.run(zero + 0.0 * s[:, :1, None, None].to(torch.float32), n_steps=1)
b.record()
torch.cuda.synchronize()
out["host_stepped_zero_policy"] = {"games": m, "steps": 50, "ms": a.elapsed_time(b),
                                   "game_steps_per_s": m * 50 / (a.elapsed_time(b) * 1e-3)}
print(json.dumps(out))
