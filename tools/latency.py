#!/usr/bin/env python
"""tools/latency.py : BASELINE configs[0] alone -- microseconds per GameSim.run() call of the drop-in class, and of the
bare C-ABI call underneath it (GPU box only)."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from bench_extras import single_step_latency
from robo_game_sim_b200 import GameSim
for _ in range(3):
    print(single_step_latency())
sim = GameSim()
sim.reset(0.05, 1 << 30, np.array([-1.0, 0.0, 0.0, 0.0]), 0.0, False, "")
rng = np.random.default_rng(0)
u = rng.uniform(-10, 10, size=(201, 4, 2))      # the command sequence of single_step_latency()
acts = [(C.c_double * 8)(*u[k].ravel()) for k in range(201)]
o = (C.c_double * 22)()
f, h = sim._lib.rgs_single_run, sim._h
for k in range(50):
    f(h, acts[0], o)
sim.reset(0.05, 1 << 30, np.array([-1.0, 0.0, 0.0, 0.0]), 0.0, False, "")
t0 = time.perf_counter()
for k in range(2000):
    f(h, acts[k // 10], o)
print("bare rgs_single_run through ctypes, same game: %.2f us per call" % ((time.perf_counter() - t0) / 2000 * 1e6))
