"""Device-timed figures for the BASELINE.json configurations and modes that are NOT the bench.py headline line.
bench.py calls measure_extras() on rank 0 at N = 1 and carries the result in the same JSON line:

  modes.exact            K1 (fp64 event-exact parity mode), 2x10^5 S0 games x 200 steps, own FP64-issue roofline
  configs.nvn            BASELINE configs[3]: 4v4 and 7v7 team games, fp32 throughput kernel and fp64 exact kernel
  configs.collision_batch  BASELINE configs[2] size: the exact kernel on 10^5 games x 400 steps
  configs.noise          the FAST kernels with process noise on (reset(noise = 0.05), reference src/GameSim.cpp:62-68), fp32 and fp64
  configs.sampler_tick_ms  BASELINE configs[4]: 4096 candidates x 50 steps from one shared state (reset_shared +
                         rollout + terminal cost / arg-min + 16-byte D2H), latency per controller tick
  configs.logging        fp32 rollout with a trajectory record written every 1 / 10 steps: game-steps/s and achieved
                         HBM GB/s of the trajectory writes (north star: "achieved HBM GB/s on logging runs"); every
                         step also in the field-major (SoA) trajectory layout
  latency.single_step_us BASELINE configs[0]: the drop-in GameSim (batch of one through rgs_single_run) over a
                         2000-step game (py/run_sim.py: T = 100 s), microseconds per run() call

GPU box only; inputs are the K5 synthetic S0 generator's, resident in HBM."""
import ctypes as C
import time

import numpy as np
import torch

from robo_game_sim_b200 import BatchedGameSim, GameSim, _capi, synth_inputs
from robo_game_sim_b200 import sampler as sp

HOLD = 10
W_REF = {1: 50 * (2 * 71 + 79 + 8 * 3 + 7 * 3), 2: 23650, 4: 48550, 7: 96400}     # RK4 formulation
W_AFF = {1: 50 * (2 * 14 + 6 + 8 * 3 + 6 * 3), 2: 8100, 4: 20300, 7: 47600}       # affine formulation


def timed(fn, reps=4):
    """median of reps-1 device-timed calls after one warm call, seconds"""
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


def rollout_rate(n_team, mode, n_games, steps, peak_tflops=None, reps=4, noise=0.0):
    sim = BatchedGameSim(n_games, mode=mode, players_per_team=n_team, winning_score=4, noise=noise, seed=1)
    x0, acts = synth_inputs(n_games, steps // HOLD, seed=11, players_per_team=n_team, dtype=sim.action_dtype)
    sim.reset(x0_ball=x0)
    sec = timed(lambda: (sim.reset(x0_ball=x0), sim.run(acts, n_steps=steps, hold=HOLD)), reps=reps)
    live = int(sim.stats().cpu().numpy()[_capi.ST_EVENTS0 + 10])
    out = {"players_per_team": n_team, "mode": mode, "games": n_games, "steps": steps, "ms": sec * 1e3,
           "live_game_steps": live, "game_steps_per_s": live / sec}
    if noise:
        out["noise"] = noise
    if peak_tflops:
        w = (W_REF if mode == "exact" else W_AFF)[n_team]
        out["roofline"] = {"bound": "fp64-issue" if mode != "fast" else "fp32-issue", "flop_per_game_step": w,
                           "achieved": w * live / sec / 1e12, "peak": peak_tflops, "unit": "TFLOP/s",
                           "frac": w * live / sec / 1e12 / peak_tflops}
    return out


def sampler_tick():
    Cn, H = 4096, 50
    sim = BatchedGameSim(Cn, mode="fast")
    state = np.zeros(22)
    state[0:4] = [-3.0, 1.0, 0.5, 0.0]
    state[4:8] = [-2.0, -1.5, 0.0, 0.3]
    state[8:12] = [3.0, 1.0, -0.5, 0.0]
    state[12:16] = [2.0, -1.0, 0.0, 0.0]
    state[16:20] = [0.3, 0.2, 1.0, -0.5]
    play = sp.dmpc_play(state, field=-1)
    x_des = {0: sp.dmpc_target(state, play, sp.ATTACKER, -1, goal_pos=[5, 0], home_pos=[-5, 0]),
             1: sp.dmpc_target(state, play, sp.DEFENDER, -1, goal_pos=[5, 0], home_pos=[-5, 0])}
    acts = torch.as_tensor(sp.candidate_actions(Cn, H // HOLD, seed=5)).cuda()
    xd, w = np.zeros((4, 4)), np.zeros((4, 4))
    for p, t in x_des.items():
        xd[p], w[p] = t, sp.OMEGA_N_MAX
    xd_t, w_t = torch.as_tensor(xd).cuda(), torch.as_tensor(w).cuda()

    def tick():
        sim.reset_shared(state)
        sim.run(acts, n_steps=H, hold=HOLD)
        sim.terminal_cost(xd_t, w_t)     # ends with the 16-byte D2H of (best index, best cost)

    sec = timed(tick, reps=12)
    for _ in range(3):
        sp.sample_rollouts(sim, state, acts, x_des, n_steps=H, hold=HOLD)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        sp.sample_rollouts(sim, state, acts, x_des, n_steps=H, hold=HOLD)
    torch.cuda.synchronize()
    return {"candidates": Cn, "horizon": H, "device_ms": sec * 1e3, "public_call_wall_ms": (time.perf_counter() - t0) / 20 * 1e3}


def logging_runs(hbm_peak_gbs):
    n, steps = 200_000, 100
    out = {}
    sim = BatchedGameSim(n, mode="fast", winning_score=4)
    x0, acts = synth_inputs(n, steps // HOLD, seed=3, dtype=torch.float32)
    for stride, layout in ((0, "records"), (10, "records"), (1, "records"), (1, "field_major")):
        traj = [None]

        def go():
            sim.reset(x0_ball=x0)
            traj[0] = sim.run(acts, n_steps=steps, hold=HOLD, log_stride=stride, traj_layout=layout)

        # reset is inside the timed call on purpose: the same region for every stride
        sec = timed(go)
        nbytes = 0 if traj[0] is None else traj[0].numel() * 8
        out[f"log_stride_{stride}" + ("_field_major" if layout != "records" else "")] = {"game_steps_per_s": n * steps / sec, "ms": sec * 1e3, "traj_bytes": nbytes,
                                       "traj_write_gbs": nbytes / sec / 1e9, "hbm_frac": nbytes / sec / 1e9 / hbm_peak_gbs}
        traj[0] = None
    base = out["log_stride_0"]["game_steps_per_s"]
    for k in ("log_stride_10", "log_stride_1", "log_stride_1_field_major"):
        out[k]["loss_vs_no_log"] = 1.0 - out[k]["game_steps_per_s"] / base
    out["games"], out["steps"] = n, steps
    return out


def single_step_latency(steps=2000):
    """BASELINE configs[0]: py/run_sim.py's loop (T = 100 s, dt = 0.05, x0_ball = (-1, 0, 0, 0), run_sim.py:28)
    against the drop-in class, wall time per run() call (H2D 64 B, one 1-step launch, D2H, stream sync)."""
    sim = GameSim()
    rng = np.random.default_rng(0)
    u = rng.uniform(-10, 10, size=(steps // HOLD + 1, 4, 2))
    sim.reset(0.05, 1 << 30, np.array([-1.0, 0.0, 0.0, 0.0]), 0.0, False, "")
    for k in range(50):
        sim.run(*u[0])
    sim.reset(0.05, 1 << 30, np.array([-1.0, 0.0, 0.0, 0.0]), 0.0, False, "")
    t0 = time.perf_counter()
    for k in range(steps):
        a = u[k // HOLD]
        sim.run(a[0], a[1], a[2], a[3])
    sec = time.perf_counter() - t0
    return {"steps": steps, "single_step_us": sec / steps * 1e6, "game_wall_ms": sec * 1e3,
            "api": "robo_game_py.GameSim.run -> rgs_single_run (K1 fp64, batch of one)"}


def measure_extras(device, hbm_peak_gbs):
    lib = _capi.load()
    pk32, pk64 = C.c_double(0.0), C.c_double(0.0)
    _capi.check(lib.rgs_measure_peak(device, 0, C.byref(pk32)))
    _capi.check(lib.rgs_measure_peak(device, 1, C.byref(pk64)))
    out = {"peaks": {"fp32_fma_tflops": pk32.value, "fp64_fma_tflops": pk64.value,
                     "fp32_non_fma_lane_ops_tflops": pk32.value / 2, "fp64_non_fma_lane_ops_tflops": pk64.value / 2,
                     "fp32_nominal_tflops": 148 * 128 * 2 * 1.965e9 / 1e12,
                     "source": "rgs_measure_peak FMA-chain micro-kernel on this GPU; nominal = 148 SMs x 128 lanes x 2 x 1.965 GHz"}}
    out["modes"] = {"exact": rollout_rate(2, "exact", 200_000, 200, pk64.value),
                    "fast64": rollout_rate(2, "fast64", 200_000, 200, pk64.value)}
    out["configs"] = {
        "nvn": [rollout_rate(4, "fast", 400_000, 200, pk32.value), rollout_rate(7, "fast", 400_000, 200, pk32.value),
                rollout_rate(4, "exact", 100_000, 100, pk64.value), rollout_rate(7, "exact", 100_000, 100, pk64.value)],
        "collision_batch": rollout_rate(2, "exact", 100_000, 400, pk64.value, reps=3),
        "sampler_tick": sampler_tick(),
        "noise": [rollout_rate(2, "fast", 200_000, 200, reps=3, noise=0.05), rollout_rate(2, "fast64", 100_000, 200, reps=3, noise=0.05)],
        "logging": logging_runs(hbm_peak_gbs),
    }
    out["configs"]["sampler_tick_ms"] = out["configs"]["sampler_tick"]["device_ms"]
    out["latency"] = single_step_latency()
    return out
