#!/usr/bin/env python
"""Three sampler ticks (4096 candidates x 50 steps from a shared state) for a launch list:
   ncu --metrics gpu__time_duration.sum --clock-control none --csv python tools/tick_once.py"""
import sys

import numpy as np
import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from robo_game_sim_b200 import BatchedGameSim, sampler as sp  # noqa: E402

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
acts = torch.as_tensor(sp.candidate_actions(C, H // 10, seed=5)).cuda()
for _ in range(3):
    best, first, cost = sp.sample_rollouts(sim, state, acts, x_des, n_steps=H, hold=10)
torch.cuda.synchronize()
ev = sim.events().cpu().numpy()
print("best", best, "loop iters mean/max", ev[:, 9].mean(), ev[:, 9].max())
