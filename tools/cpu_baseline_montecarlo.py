#!/usr/bin/env python
"""tools/cpu_baseline_montecarlo.py -- the two CPU-container-only baseline figures BASELINE.md section 3 promises:

  driver (i): the reference's own py/MonteCarlo.py::monte_carlo, UNMODIFIED and read in place from /root/reference, driven
      with open-loop replay callables as the two team methods.  Its `from robo_game_py import GameSim` resolves to a
      ctypes class over the compiled reference (oracle/_ref: the unmodified src/GameSim.cpp; the reference's pybind
      module itself cannot be built here, Eigen is absent), `progressbar` to a 3-line stand-in.  log=True is hard-wired
      in MonteCarlo.py:29, so the figure includes the binary log.
  no-O build: the same reference sources compiled with the reference's own flags (CMakeLists.txt:4: -std=c++11, no
      optimisation) and stepped through the C wrapper.

Needs /root/reference: build container only (the GPU box has no reference tree).  Prints one JSON line.
"""
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import rgs_testlib as tl  # noqa: E402

REF = "/root/reference"
dp = C.POINTER(C.c_double)


def make_modules(lib):
    class GameSim:   # robo_game_py.GameSim over the compiled reference (reference src/RoboGamePy.cpp:15-24)
        def __init__(self):
            self.h = lib.ref_create()

        def __del__(self):
            lib.ref_destroy(self.h)

        def reset(self, dt, ws, x0, noise, log, logname, seed, tau_puck, tau_player, pm, km):
            out = np.zeros(22)
            x0 = np.ascontiguousarray(x0, dtype=np.float64)
            lib.ref_reset(self.h, dt, ws, x0.ctypes.data_as(dp), noise, 1 if log else 0, logname.encode(), seed, tau_puck,
                          tau_player, pm, km, out.ctypes.data_as(dp))
            return out

        def run(self, a1, a2, b1, b2):
            u = np.ascontiguousarray(np.concatenate([a1, a2, b1, b2]), dtype=np.float64)
            out = np.zeros(22)
            lib.ref_run(self.h, u.ctypes.data_as(dp), out.ctypes.data_as(dp))
            return out

    m = types.ModuleType("robo_game_py")
    m.GameSim = GameSim
    sys.modules["robo_game_py"] = m
    pb = types.ModuleType("progressbar")
    pb.progressbar = lambda it: it
    sys.modules["progressbar"] = pb


def driver_i(runs=8, T=10.0):
    lib = tl.ref_lib()
    assert lib is not None and os.path.exists(REF), "needs /root/reference"
    make_modules(lib)
    sys.path.insert(0, os.path.join(REF, "py"))
    from MonteCarlo import monte_carlo   # the reference's file, unmodified

    steps = int(np.ceil(T / 0.05))
    _, acts = tl.synth_inputs(20261104, 0, 1, 2, steps // 10 + 1)
    k = {"A": 0, "B": 0}

    def team(name, lo):
        def run(sim_state):
            a = acts[0, min(k[name] // 10, acts.shape[1] - 1)]
            k[name] += 1
            return a[lo], a[lo + 1]
        return run

    with tempfile.TemporaryDirectory() as td:
        p = types.SimpleNamespace(num_runs=runs, T=T, dt=0.05, winning_score=4, x0_ball=np.array([-1.0, 0.0, 0.0, 0.0]),
                                  noise_stdev=0.0, tau_puck=0.1, tau_player=0.5, player_mass=1.0, puck_mass=0.5,
                                  log_prefix=os.path.join(td, "mc"))
        out = open(os.devnull, "w")
        saved, sys.stdout = sys.stdout, out
        try:
            t0 = time.perf_counter()
            monte_carlo(p, team("A", 0), team("B", 2))
            sec = time.perf_counter() - t0
        finally:
            sys.stdout = saved
    return {"runs": runs, "steps_per_run": steps + 1, "seconds": sec, "game_steps_per_s_1core": runs * (steps + 1) / sec}


def no_opt_build(games=8, steps=200):
    with tempfile.TemporaryDirectory() as td:
        so = os.path.join(td, "libref_noopt.so")
        subprocess.run(["g++", "-std=c++11", "-w", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "oracle", "eigen_shim"),
                        "-I" + os.path.join(REF, "include"), os.path.join(ROOT, "oracle", "ref_wrapper.cpp"),
                        os.path.join(REF, "src", "GameSim.cpp"), "-o", so], check=True)
        lib = C.CDLL(so)
        lib.ref_rollout_batch.argtypes = [C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_int, dp, dp, C.c_int64, C.c_double,
                                          C.c_double, C.c_double, C.c_double, dp]
        x0, acts = tl.synth_inputs(20261104, 0, games, 2, steps // 10)
        out = np.zeros((games, 22))
        t0 = time.perf_counter()
        lib.ref_rollout_batch(games, steps, 10, 0.05, 4, x0.ctypes.data_as(dp), np.ascontiguousarray(acts).ctypes.data_as(dp),
                              steps // 10, 0.1, 0.5, 1.0, 0.5, out.ctypes.data_as(dp))
        sec = time.perf_counter() - t0
        ref = tl.ref_lib()
        out2 = np.zeros((games, 22))
        t0 = time.perf_counter()
        ref.ref_rollout_batch(games, steps, 10, 0.05, 4, x0.ctypes.data_as(dp), np.ascontiguousarray(acts).ctypes.data_as(dp),
                              steps // 10, 0.1, 0.5, 1.0, 0.5, out2.ctypes.data_as(dp))
        sec2 = time.perf_counter() - t0
        assert np.array_equal(out, out2), "-O0 and -O2 builds of the reference disagree"
    return {"games": games, "steps": steps, "no_O_game_steps_per_s_1core": games * steps / sec,
            "O2_game_steps_per_s_1core": games * steps / sec2, "bit_identical": True}


if __name__ == "__main__":
    print(json.dumps({"driver_i_MonteCarlo_py": driver_i(), "reference_default_flags": no_opt_build()}))
