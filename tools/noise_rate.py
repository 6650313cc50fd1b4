#!/usr/bin/env python
"""tools/noise_rate.py : fp32 / fp64 K2 throughput with process noise on (reset(noise > 0), a16), 2v2 S0 games (GPU box only)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
from robo_game_sim_b200 import BatchedGameSim, _capi, synth_inputs
from bench_extras import timed
for mode, n in (("fast", 400_000), ("fast64", 200_000)):
    for noise in (0.0, 0.05, 2.0):
        sim = BatchedGameSim(n, mode=mode, winning_score=4, noise=noise, seed=1)
        x0, acts = synth_inputs(n, 20, seed=11, dtype=sim.action_dtype)
        sim.reset(x0_ball=x0)
        sec = timed(lambda: (sim.reset(x0_ball=x0), sim.run(acts, n_steps=200, hold=10)))
        live = int(sim.stats().cpu().numpy()[_capi.ST_EVENTS0 + 10])
        print(mode, "noise", noise, "%.4g game-steps/s  %.1f ms" % (live / sec, sec * 1e3))
