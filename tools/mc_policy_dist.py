#!/usr/bin/env python
"""Sharding invariance of the closed-loop Monte Carlo: `torchrun --nproc-per-node N tools/mc_policy_dist.py` runs
monte_carlo_batched(policy="classical+cbf") with the runs sharded over N ranks (statistics summed by one NCCL
all-reduce) and, on rank 0, the same runs on one GPU; the two statistics vectors must be equal (run index = global
game index keys every random stream).  Prints one JSON line."""
import json
import os
import sys
import types

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from robo_game_sim_b200 import monte_carlo_batched  # noqa: E402

world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
solo = dist.new_group([0]) if world > 1 else None     # rank 0 alone: the single-GPU run reduces over nobody
mc = types.SimpleNamespace(num_runs=400_001, T=10.0, dt=0.05, winning_score=4, x0_ball=np.array([-1.0, 0, 0, 0]),
                           noise_stdev=0.0, tau_puck=0.1, tau_player=0.5, player_mass=1.0, puck_mass=0.5, log_prefix=None)
keys = ("games", "wins_a", "wins_b", "undecided", "goals_a", "goals_b", "player_player", "player_puck", "wall_up",
        "loop_iters", "live_steps")
sharded = monte_carlo_batched(mc, policy="classical+cbf", randomize=True, seed=11, device=local)
if int(os.environ.get("RANK", "0")) == 0:
    single = monte_carlo_batched(mc, policy="classical+cbf", randomize=True, seed=11, device=local, rank=0, world_size=1,
                                 group=solo)
    same = all(sharded[k] == single[k] for k in keys) and np.array_equal(sharded["score_hist"], single["score_hist"])
    print(json.dumps({"world": world, "equal": bool(same), "sharded": {k: sharded[k] for k in keys},
                      "single": {k: single[k] for k in keys}}))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
