#!/usr/bin/env python
"""tools/e2e_probe.py : what the end-to-end pass of each rank is made of (run under torchrun on N GPUs): device-only
pass, host-buffer pass with and without the per-game event records, and the bare H2D / D2H copies of one pass timed
alone and with all ranks copying at once.  Prints one line per rank-0 with the per-rank figures (ms)."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch, torch.distributed as dist
from robo_game_sim_b200 import BatchedGameSim, synth_inputs

world, rank, lr = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
G, H, HOLD = 1_000_000, 200, 10
sim = BatchedGameSim(G, mode="fast", winning_score=4)
x0, acts = synth_inputs(G, H // HOLD, seed=1, first_game=rank * G, dtype=torch.float32)
hx0 = torch.empty((G, 4), dtype=torch.float64).pin_memory(); hx0.copy_(x0)
hact = torch.empty(tuple(acts.shape), dtype=torch.float32).pin_memory(); hact.copy_(acts)
out = {"state": torch.empty((G, 22), dtype=torch.float64).pin_memory(), "events": torch.empty((G, 16), dtype=torch.int64).pin_memory()}
dst = torch.empty_like(acts)
dstate = torch.empty((G, 38), dtype=torch.float64, device=dev)
hstate = torch.empty((G, 38), dtype=torch.float64).pin_memory()

def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()

def wall(fn, reps=3):
    fn(); barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3

def device_pass():
    sim.reset(x0_ball=x0); sim.run(acts, n_steps=H, hold=HOLD); sim.stats()

res = {
    "device": wall(device_pass),
    "e2e": wall(lambda: sim.run_host(hx0, hact, H, hold=HOLD, out=out)),
    "e2e_no_events": wall(lambda: sim.run_host(hx0, hact, H, hold=HOLD, want_events=False, out=out)),
    "h2d_640MB": wall(lambda: dst.copy_(hact, non_blocking=True)),
    "d2h_304MB": wall(lambda: hstate.copy_(dstate, non_blocking=True)),
}
keys = list(res)
t = torch.tensor([res[k] for k in keys], dtype=torch.float64, device=dev)
allr = [t.clone() for _ in range(world)]
if world > 1:
    dist.all_gather(allr, t)
if rank == 0:
    for i, k in enumerate(keys):
        print(k, [round(float(a[i]), 2) for a in allr])
if world > 1:
    dist.destroy_process_group()
