#!/usr/bin/env python
"""bench.py -- game-steps/s of the GameSim hot path (BASELINE.json metric) on N B200s.

One bench "step" = one Monte Carlo pass over the per-GPU batch: batched GameSim::reset, then
`horizon` fused game-steps of every game from device-resident fp32 command tensors (K2), then the K4
statistics reduction (+ one NCCL all-reduce of int64[96] when N > 1).  Workload = BASELINE.json
configs[1] (MonteCarlo 2v2, 10^6 randomized games x 200 steps, S0 inputs); per-GPU work is fixed as
N grows (weak scaling).  `value` counts LIVE game-steps (GameSim::run is a no-op once decided).

  python bench.py [--gpus N] [--steps K] [--warmup W]            # our arm (torchrun for N > 1)
  python bench.py --impl reference ...                            # reference C++ GameSim on host cores
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

W_AFF_2V2 = 8100      # algorithmic flop / game-step, affine fp32 formulation
W_REF_2V2 = 23650     # algorithmic flop / game-step, RK4 formulation (K1)
HOLD = 10
SEED = 20261104
WINNING_SCORE = 4     # py/mc_example_1.py:31
METRIC = "game-steps/sec (2v2, whole box, device-timed)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--games", type=int, default=1_000_000, help="games per GPU")
    ap.add_argument("--horizon", type=int, default=200, help="game-steps per game per pass (T=10 s at dt=0.05)")
    ap.add_argument("--mode", default="fast", choices=["fast", "exact", "fast64"])
    ap.add_argument("--cpu-games-per-core", type=int, default=192)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip modes / configs / latency (tools/bench_extras.py)")
    ap.add_argument("--_worker", nargs=4, default=None, help=argparse.SUPPRESS)
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# CPU side: the reference's own C++ GameSim (oracle/_ref, compiled from /root/reference) or, when
# that is absent, the plain-C oracle port -- one PROCESS per host core (GameSim keeps function-local
# static scratch, reference src/GameSim.cpp:72,100, so threads are not safe).
# ------------------------------------------------------------------------------------------------
def _cpu_kind():
    ref = os.path.join(ROOT, "oracle", "_ref", "libref_gamesim.so")
    if os.path.exists(ref):
        return "reference", ref
    port = os.path.join(ROOT, "oracle", "liboracle_gamesim.so")
    if not os.path.exists(port):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, stdout=subprocess.DEVNULL)
    return "port", port


def _worker(npz, lo, hi, horizon):
    """Runs in a child process: games [lo, hi) of the sample, timing only the C call."""
    import ctypes as C

    import numpy as np

    lo, hi, horizon = int(lo), int(hi), int(horizon)
    d = np.load(npz)
    x0 = np.ascontiguousarray(d["x0"][lo:hi])
    acts = np.ascontiguousarray(d["acts"][lo:hi])
    n, n_act = x0.shape[0], acts.shape[1]
    kind, path = _cpu_kind()
    lib = C.CDLL(path)
    dp = C.POINTER(C.c_double)
    final = np.zeros((n, 22))
    if kind == "reference":
        lib.ref_rollout_batch_live.restype = C.c_int64
        lib.ref_rollout_batch_live.argtypes = [C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_int, dp, dp, C.c_int64,
                                               C.c_double, C.c_double, C.c_double, C.c_double, dp]
        t0 = time.perf_counter()
        live = lib.ref_rollout_batch_live(n, horizon, HOLD, 0.05, WINNING_SCORE, x0.ctypes.data_as(dp),
                                          acts.ctypes.data_as(dp), n_act, 0.1, 0.5, 1.0, 0.5, final.ctypes.data_as(dp))
        sec = time.perf_counter() - t0
    else:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import rgs_testlib as tl

        pd = tl.default_params(winning_score=WINNING_SCORE)
        s0 = tl.oracle_reset(pd, x0)
        t0 = time.perf_counter()
        r = tl.oracle_rollout(pd, s0, acts, horizon, HOLD, threads=1)
        sec = time.perf_counter() - t0
        final, live = r["state"], int(r["events"][:, tl.EV["LIVE_STEPS"]].sum())
    decided = int(((final[:, 0] >= WINNING_SCORE) | (final[:, 1] >= WINNING_SCORE)).sum())
    # LIVE game-steps (run() is a no-op once a game is decided), the same basis as the GPU arm's value
    print(json.dumps({"steps": int(live), "seconds": sec, "decided": decided}))


def cpu_pass(x0, acts, horizon, cores):
    """One bounded-sample pass on `cores` processes; returns (game-steps, seconds = max over workers)."""
    import numpy as np

    n = x0.shape[0]
    with tempfile.TemporaryDirectory() as td:
        npz = os.path.join(td, "sample.npz")
        np.savez(npz, x0=x0, acts=acts)
        procs = []
        for c in range(cores):
            lo, hi = n * c // cores, n * (c + 1) // cores
            if hi > lo:
                procs.append(subprocess.Popen([sys.executable, os.path.abspath(__file__), "--_worker", npz, str(lo), str(hi),
                                               str(horizon)], stdout=subprocess.PIPE, text=True))
        outs = [json.loads(p.communicate()[0].strip().splitlines()[-1]) for p in procs]
    return sum(o["steps"] for o in outs), max(o["seconds"] for o in outs), sum(o["decided"] for o in outs)


def host_cores():
    return len(os.sched_getaffinity(0))


def cpu_sample(cores, games_per_core, horizon):
    from robo_game_sim_b200.synth_np import synth_inputs

    n = cores * games_per_core
    x0, acts = synth_inputs(SEED, 0, n, 2, horizon // HOLD)   # the first n games of the GPU workload
    return x0, acts.reshape(n, horizon // HOLD, 8)


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    kind, _ = _cpu_kind()
    x0, acts = cpu_sample(cores, args.cpu_games_per_core, args.horizon)
    for _ in range(max(0, min(args.warmup, 1))):   # one warm pass is enough for a CPU loop
        cpu_pass(x0, acts, args.horizon, cores)
    steps, sec = 0, 0.0
    for _ in range(args.steps):
        s, t, _ = cpu_pass(x0, acts, args.horizon, cores)
        steps += s
        sec += t
    value = steps / sec
    games_timed = int(x0.shape[0]) * args.steps
    sample = f"{x0.shape[0]} games x {args.horizon} steps per pass ({cores} processes x {args.cpu_games_per_core} games), first games of the S0 workload"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "game-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sec / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": value, "unit": "game-steps/s", "cores": cores, "kind": kind, "sample": sample},
        "games_timed": games_timed, "live_game_steps_timed": int(steps),
        "e2e": {"value": value, "unit": "game-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def workload_config(args, n_gpus):
    return {"workload": f"MonteCarlo 2v2 (BASELINE configs[1]): {args.games} randomized games per GPU x {args.horizon} "
                        f"game-steps, S0 inputs (puck ~U, commands ~U(-10,10) redrawn every {HOLD} steps), dt=0.05, "
                        f"winning_score={WINNING_SCORE}",
            "games_per_gpu": args.games, "horizon": args.horizon, "hold": HOLD, "players_per_team": 2,
            "mode": args.mode, "parallelism": f"games sharded over {n_gpus} GPU(s), one all-reduce of int64[96] stats",
            "l2": "inputs larger than L2 (fp32 command tensor 32 B x games x horizon/hold; state 176 B x games)"}


# ------------------------------------------------------------------------------------------------
def clocks_sampler_start(gpu_index):
    try:
        f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index),
                              "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
                              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
                              "clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-lms", "100"],
                             stdout=f, stderr=subprocess.DEVNULL)
        return p, f
    except OSError:
        return None, None


def clocks_sampler_stop(p, f):
    if p is None:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    p.terminate()
    try:
        p.wait(timeout=5)
    except subprocess.TimeoutExpired:
        p.kill()
    f.flush()
    f.seek(0)
    rows = [r.strip().split(",") for r in f.read().strip().splitlines() if r.strip()]
    f.close()
    os.unlink(f.name)
    sm, smax, reasons, power = [], [], set(), []
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for r in rows:
        try:
            sm.append(float(r[0]))
            smax.append(float(r[1]))
            power.append(float(r[2]))
            for k, nm in enumerate(names):
                if r[3 + k].strip().lower().startswith("active"):
                    reasons.add(nm)
        except (ValueError, IndexError):
            continue
    if not sm:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
    sm.sort()
    return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(smax), "power_w_max": max(power), "samples": len(sm),
            "reasons": sorted(reasons)}


def bind_to_gpu_numa_node(local_rank):
    """Pin this rank (and therefore the first-touch placement of its pinned host buffers) to the CPUs of the NUMA node
    its GPU hangs off: the end-to-end path stages ~1 GB per pass per GPU through host memory.  Returns what it did."""
    try:
        import torch

        pr = torch.cuda.get_device_properties(local_rank)
        if not all(hasattr(pr, k) for k in ("pci_domain_id", "pci_bus_id", "pci_device_id")):
            return {"numa_node": None, "note": "no PCI ids in the device properties"}
        path = f"/sys/bus/pci/devices/{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return {"numa_node": node, "note": "platform reports no NUMA affinity for the GPU"}
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "cpus": len(cpus)}
    except Exception as exc:  # binding is an optimisation, never a failure
        return {"numa_node": None, "note": f"{type(exc).__name__}: {exc}"[:120]}


def main():
    args = parse()
    if args._worker:
        _worker(*args._worker)
        return
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    from robo_game_sim_b200 import BatchedGameSim, _capi, synth_inputs

    # stdout carries exactly one JSON line: anything native libraries print there (NCCL's version banner) is sent
    # to stderr by pointing fd 1 at fd 2 for the duration of the run; the line itself goes to the saved descriptor
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; robo_game_sim_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else {"numa_node": None, "note": "single rank: not bound"}
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    lib = _capi.load()
    G, H = args.games, args.horizon
    n_act = (H + HOLD - 1) // HOLD
    first_game = rank * G                      # global game index keys every random stream
    sim = BatchedGameSim(G, device=local_rank, mode=args.mode, winning_score=WINNING_SCORE)
    x0, acts = synth_inputs(G, n_act, seed=SEED, first_game=first_game, dtype=sim.action_dtype, device=local_rank)
    torch.cuda.synchronize()

    stats = None

    def one_pass(ev_pair=None):
        nonlocal stats
        sim.reset(x0_ball=x0)
        if ev_pair:
            ev_pair[0].record()
        sim.run(acts, n_steps=H, hold=HOLD)
        if ev_pair:
            ev_pair[1].record()
        stats = sim.stats()
        if world > 1:
            dist.all_reduce(stats, op=dist.ReduceOp.SUM)   # NCCL over NVLink: the path's only collective

    for _ in range(args.warmup):
        one_pass()
    torch.cuda.synchronize()
    barrier()

    sampler = clocks_sampler_start(local_rank) if rank == 0 else (None, None)
    launches0 = _capi.launch_count()
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_beg, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    t_beg.record()
    for k in range(args.steps):
        one_pass(kev[k])
    t_end.record()
    torch.cuda.synchronize()
    barrier()
    launches = _capi.launch_count() - launches0
    clocks = clocks_sampler_stop(*sampler) if rank == 0 else None

    ms = torch.tensor([t_beg.elapsed_time(t_end)], dtype=torch.float64, device=dev)
    kms = torch.tensor([sum(a.elapsed_time(b) for a, b in kev) / max(1, args.steps)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(kms, op=dist.ReduceOp.MAX)
    total_ms, kernel_ms = float(ms.item()), float(kms.item())
    st = stats.cpu().numpy()
    live_per_pass = int(st[_capi.ST_EVENTS0 + 10])          # LIVE_STEPS summed over all games of all ranks
    value = live_per_pass * args.steps / (total_ms * 1e-3)

    # ---- end to end: host buffers in, host results out, through the public host-buffer call ----
    e2e = None
    if not args.no_e2e:
        hx0 = torch.empty((G, 4), dtype=torch.float64).pin_memory()
        hact = torch.empty((G, n_act, 4, 2), dtype=sim.action_dtype).pin_memory()
        hx0.copy_(x0)
        hact.copy_(acts)
        out = {"state": torch.empty((G, 22), dtype=torch.float64).pin_memory(),
               "events": torch.empty((G, 16), dtype=torch.int64).pin_memory()}
        sim.run_host(hx0, hact, H, hold=HOLD, out=out)     # warm-up (allocates staging)
        reps = max(1, min(args.steps, 3))
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            _, _, hstats = sim.run_host(hx0, hact, H, hold=HOLD, out=out)
        my_sec = time.perf_counter() - t0
        sec = torch.tensor([my_sec], dtype=torch.float64, device=dev)
        live = torch.tensor([int(hstats[_capi.ST_EVENTS0 + 10])], dtype=torch.int64, device=dev)
        rank_secs = [sec.clone() for _ in range(world)]
        if world > 1:
            dist.all_gather(rank_secs, sec)
            dist.all_reduce(sec, op=dist.ReduceOp.MAX)
            dist.all_reduce(live, op=dist.ReduceOp.SUM)
        esz = 4 if sim.action_dtype == torch.float32 else 8
        e2e = {"value": float(live.item()) * reps / float(sec.item()), "unit": "game-steps/s",
               "h2d_bytes_per_step": world * (G * 4 * 8 + G * n_act * 8 * esz),
               "d2h_bytes_per_step": world * (G * 22 * 8 + G * 16 * 8 + 96 * 8), "passes": reps,
               # per-rank staging rates: (H2D + D2H bytes of one rank) / its own wall time per pass
               "per_rank_host_gbs": [round((G * 4 * 8 + G * n_act * 8 * esz + G * 22 * 8 + G * 16 * 8) * reps / float(t.item()) / 1e9, 2)
                                     for t in rank_secs],
               "per_rank_pass_ms": [round(float(t.item()) / reps * 1e3, 2) for t in rank_secs], "numa": numa,
               "api": "BatchedGameSim.run_host -> rgs_batch_run_host (pinned host buffers, 6 game-range chunks of 2, 6, 8, 8, 6, "
                      "2 thirty-seconds of the batch double-buffered over 2 streams)"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (the fused rollout) ----
    peak = C_double = None
    import ctypes as C
    pk = C.c_double(0.0)
    fp64_mode = args.mode != "fast"
    _capi.check(lib.rgs_measure_peak(local_rank, 1 if fp64_mode else 0, C.byref(pk)))
    W = W_REF_2V2 if args.mode == "exact" else W_AFF_2V2
    live_per_gpu_launch = live_per_pass / world
    achieved = W * live_per_gpu_launch / (kernel_ms * 1e-3) / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    esz = 4 if args.mode == "fast" else 8
    alg_bytes = G * (2 * 22 * 8 + 2 * 16 * 8 + 8) + G * n_act * 8 * esz   # state+events in/out, clock, commands
    # DRAM traffic is measured nowhere but in ncu (--set full capture of the rollout kernel); the file names the
    # commit the capture was taken on, so a stale figure is visible in the line
    traffic, traffic_src = None, None
    tfile = os.path.join(ROOT, "profiles", "traffic_latest.json")
    if os.path.exists(tfile) and args.mode == "fast":
        try:
            tj = json.load(open(tfile))
            traffic = tj.get("dram_bytes_per_launch")
            traffic_src = {k: tj.get(k) for k in ("source", "commit", "kernel")}
        except (OSError, ValueError):
            pass
    roofline = {"bound": "fp64-issue" if fp64_mode else "fp32-issue", "achieved": achieved, "peak": pk.value,
                "unit": "TFLOP/s", "frac": achieved / pk.value if pk.value else None, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": "rgs_measure_peak FMA micro-kernel on this GPU (MEASURED_PEAKS.json holds only HBM and bf16 tensor peaks; "
                               "this path uses neither)",
                "flop_per_game_step": W, "kernel_ms": kernel_ms, "game_steps_per_launch": live_per_gpu_launch,
                "hbm": {"achieved_gbs": alg_bytes / (kernel_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                        "frac": alg_bytes / (kernel_ms * 1e-3) / 1e9 / hbm_peak, "algorithmic_bytes_per_launch": alg_bytes,
                        "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"}}

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        cores = host_cores()
        kind, _ = _cpu_kind()
        cx0, cacts = cpu_sample(cores, args.cpu_games_per_core, H)
        # the device-generated inputs and the host generator must be the same numbers
        assert np.array_equal(cx0, x0[:cx0.shape[0]].cpu().numpy()), "S0 generator mismatch host vs device"
        s, t, _ = cpu_pass(cx0, cacts, H, cores)
        cpu_baseline = {"value": s / t, "unit": "game-steps/s", "cores": cores, "kind": kind,
                        "sample": f"{cx0.shape[0]} games x {H} steps ({cores} processes x {args.cpu_games_per_core} games), "
                                  f"first games of the same S0 workload, fp64 C++ GameSim at -O2"}

    extras = {}
    if world == 1 and not args.no_extras:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from bench_extras import measure_extras

        del sim, x0, acts
        torch.cuda.empty_cache()
        extras = measure_extras(local_rank, hbm_peak)

    line = {"metric": METRIC, "value": value, "unit": "game-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / max(1, args.steps), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": {"fast": "f32", "exact": "f64", "fast64": "f64"}[args.mode],
            "data": "synthetic", "config": workload_config(args, world), "clocks": clocks, "e2e": e2e,
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu_baseline,
            "stats": {"live_game_steps_per_pass": live_per_pass, "games": int(st[0]), "goals": int(st[8] + st[9]),
                      "events_per_step": float(st[10:16].sum()) / max(1, live_per_pass)}}
    if cpu_baseline:
        line["vs_cpu_baseline"] = {"device_timed": value / cpu_baseline["value"],
                                   "e2e": (e2e["value"] / cpu_baseline["value"]) if e2e else None,
                                   "cores": cpu_baseline["cores"], "basis": "live game-steps on both arms"}
    line.update(extras)
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
