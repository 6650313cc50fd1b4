#!/usr/bin/env python
"""Device-timed throughput of the BASELINE.json configurations that are NOT the bench.py headline line:
  configs[3]  n-v-n team games (1v1 .. 7v7), fp32 throughput kernel and fp64 exact kernel;
  configs[4]  one controller tick of the rollout sampler: 4096 candidates x 50 steps from a shared state
              (reset_shared + rollout + terminal cost/arg-min), latency per tick;
  configs[2]  the fp64 exact kernel on a 10^5-game batch (the size of the batched collision cases).
Prints one JSON line.  GPU box only; inputs are the K5 synthetic S0 generator's, resident in HBM."""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from robo_game_sim_b200 import BatchedGameSim, synth_inputs, sampler as sp  # noqa: E402
from robo_game_sim_b200 import _capi  # noqa: E402

HOLD = 10


def timed(fn, reps=4):
    ms = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return float(np.median(ms[1:])) * 1e-3


def rollout_rate(n_team, mode, n_games, steps):
    sim = BatchedGameSim(n_games, mode=mode, players_per_team=n_team, winning_score=4)
    x0, acts = synth_inputs(n_games, steps // HOLD, seed=11, players_per_team=n_team, dtype=sim.action_dtype)

    def go():
        sim.reset(x0_ball=x0)
        sim.run(acts, n_steps=steps, hold=HOLD)

    sec = timed(go)
    live = int(sim.stats().cpu().numpy()[_capi.ST_EVENTS0 + 10])
    return {"n_team": n_team, "mode": mode, "games": n_games, "steps": steps, "ms": sec * 1e3,
            "live_game_steps": live, "game_steps_per_s": live / sec}


out = {"nvn_fast_f32": [], "nvn_exact_f64": []}
for nt in (1, 2, 3, 4, 5, 6, 7):
    out["nvn_fast_f32"].append(rollout_rate(nt, "fast", 400_000, 200))
for nt in (1, 2, 4, 7):
    out["nvn_exact_f64"].append(rollout_rate(nt, "exact", 100_000, 100))

# configs[4]: the sampler tick
C, H = 4096, 50
sim = BatchedGameSim(C, mode="fast")
state = np.zeros(22)
state[0:4] = [-3.0, 1.0, 0.5, 0.0]
state[4:8] = [-2.0, -1.5, 0.0, 0.3]
state[8:12] = [3.0, 1.0, -0.5, 0.0]
state[12:16] = [2.0, -1.0, 0.0, 0.0]
state[16:20] = [0.3, 0.2, 1.0, -0.5]
play = sp.dmpc_play(state, field=-1)
x_des = {0: sp.dmpc_target(state, play, sp.ATTACKER, -1, goal_pos=[5, 0], home_pos=[-5, 0]),
         1: sp.dmpc_target(state, play, sp.DEFENDER, -1, goal_pos=[5, 0], home_pos=[-5, 0])}
acts = torch.as_tensor(sp.candidate_actions(C, H // HOLD, seed=5)).cuda()
xd = np.zeros((4, 4))
w = np.zeros((4, 4))
for p, t in x_des.items():
    xd[p], w[p] = t, sp.OMEGA_N_MAX
xd_t, w_t = torch.as_tensor(xd).cuda(), torch.as_tensor(w).cuda()


def tick_device():
    sim.reset_shared(state)
    sim.run(acts, n_steps=H, hold=HOLD)
    sim.terminal_cost(xd_t, w_t)     # ends with the 16-byte D2H of (best index, best cost)


sec = timed(tick_device, reps=12)
out["sampler_tick"] = {"candidates": C, "horizon": H, "device_ms": sec * 1e3, "candidate_steps_per_s": C * H / sec}
import time  # noqa: E402

best = None
for _ in range(3):
    best, first, cost = sp.sample_rollouts(sim, state, acts, x_des, n_steps=H, hold=HOLD)
torch.cuda.synchronize()
t0 = time.perf_counter()
reps = 20
for _ in range(reps):
    best, first, cost = sp.sample_rollouts(sim, state, acts, x_des, n_steps=H, hold=HOLD)
torch.cuda.synchronize()
wall = (time.perf_counter() - t0) / reps
out["sampler_tick"].update({"public_call_wall_ms": wall * 1e3, "best": int(best),
                            "first_command": np.asarray(first).round(4).tolist()})
print(json.dumps(out))
