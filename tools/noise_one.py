#!/usr/bin/env python
"""tools/noise_one.py : one fp32 K2 rollout with process noise (for ncu captures; GPU box only)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from robo_game_sim_b200 import BatchedGameSim, synth_inputs
n = 200_000
sim = BatchedGameSim(n, mode="fast", winning_score=4, noise=0.05, seed=1)
x0, acts = synth_inputs(n, 20, seed=11, dtype=sim.action_dtype)
for _ in range(2):
    sim.reset(x0_ball=x0)
    sim.run(acts, n_steps=200, hold=10)
torch.cuda.synchronize()
