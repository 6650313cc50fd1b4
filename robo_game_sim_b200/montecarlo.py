"""Batched Monte Carlo harness: the role of py/MonteCarlo.py::monte_carlo (reference
py/MonteCarlo.py:8-49) with its inner `for run: for step: sim.run(...)` loop replaced by fused
rollout launches over a batch of games, and the offline MATLAB statistics
(matlab/interpret_mc_results.m:29-50) replaced by the on-device K4 reduction.

Multi-GPU: one process per GPU; the runs are split into contiguous slices (`shard_range`), every
per-run random stream is keyed by the GLOBAL run index so results do not depend on the number of
ranks, and the only collective is one all-reduce(sum) of the int64[64] statistics vector.
"""
import os
from math import ceil

import numpy as np

from . import _capi


def shard_range(num_runs, rank, world_size):
    """Contiguous slice [first, first+count) of 0-based run indices owned by `rank`."""
    per = -(-num_runs // world_size)
    first = min(rank * per, num_runs)
    return first, max(0, min(per, num_runs - first))


def write_configuration(mc_params, directory):
    """configuration.txt exactly as the reference writes it (py/MonteCarlo.py:14-21): 13 lines."""
    os.makedirs(directory, exist_ok=True)
    vals = [mc_params.num_runs, mc_params.T, mc_params.dt, mc_params.winning_score, mc_params.x0_ball[0],
            mc_params.x0_ball[1], mc_params.x0_ball[2], mc_params.x0_ball[3], mc_params.noise_stdev,
            mc_params.tau_puck, mc_params.tau_player, mc_params.player_mass, mc_params.puck_mass]
    with open(os.path.join(directory, "configuration.txt"), "w") as f:
        for v in vals:
            f.write(str(v) + "\n")


def reduce_stats(stats, group=None):
    """Sum the int64[64] statistics vector over all ranks (no-op without torch.distributed).
    NCCL reduces the device tensor in place over NVLink; gloo (CPU tests) reduces a host copy."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return stats
    if dist.get_backend(group) == "gloo" and stats.is_cuda:
        host = stats.cpu()
        dist.all_reduce(host, op=dist.ReduceOp.SUM, group=group)
        return host.to(stats.device)
    dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def stats_to_dict(stats):
    """Name the fields of the K4 statistics vector (include/rgs_capi.h RGS_ST_*)."""
    s = np.asarray(stats.cpu() if hasattr(stats, "cpu") else stats, dtype=np.int64)
    d = dict(games=int(s[_capi.ST_GAMES]), wins_a=int(s[_capi.ST_WINS_A]), wins_b=int(s[_capi.ST_WINS_B]),
             undecided=int(s[_capi.ST_UNDECIDED]), lead_a=int(s[_capi.ST_LEAD_A]), lead_b=int(s[_capi.ST_LEAD_B]),
             tied=int(s[_capi.ST_TIED]))
    for k in range(14):
        d[_capi.EV_NAMES[k]] = int(s[_capi.ST_EVENTS0 + k])
    d["score_hist"] = s[_capi.ST_SCORE_HIST:_capi.ST_SCORE_HIST + 25].reshape(5, 5).copy()
    return d


def monte_carlo_batched(mc_params, actions=None, policy=None, *, mode="fast", randomize=None, seed=1, hold=10,
                        u_max=10.0, device=None, rank=None, world_size=None, group=None, return_sim=False):
    """Run `mc_params.num_runs` games of `ceil(T/dt)` steps (after the reference's priming zero-action
    step, py/MonteCarlo.py:33) and return the reduced statistics as a dict.

    mc_params: the reference's parameter object (num_runs, T, dt, winning_score, x0_ball,
      noise_stdev, tau_puck, tau_player, player_mass, puck_mass, log_prefix).
    actions:  None, or an open-loop command tensor [n_act, 4, 2] shared by all runs / [runs, n_act, 4, 2].
    policy:   None; a callable  state[n,22] (device tensor) -> commands[n,4,2] (the batch is then stepped one
              tick per call: closed loop with the host in the loop, like the reference); or "classical" /
              "classical+cbf" / "dmpc" / "dmpc+cbf" / a dict of rgs_policy fields: the ON-DEVICE surrogate of the reference's baseline
              team for both sides (py/mc_example_1.py: ClassicalTeam vs ClassicalTeam, optionally behind the CBF
              filter), the whole closed loop in one kernel launch (BatchedGameSim.run_policy).
    randomize: True draws the synthetic S0 puck state (and, with no actions/policy, S0 commands
              redrawn every `hold` steps) per GLOBAL run index; default: True iff neither actions
              nor policy is given (the reference MC with noise=0 is otherwise N copies of one game).
    """
    import torch
    import torch.distributed as dist

    from .gamesim import BatchedGameSim

    if rank is None:
        rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
    if world_size is None:
        world_size = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    if randomize is None:
        randomize = actions is None and policy is None
    n_steps = int(ceil(mc_params.T / mc_params.dt))
    first, count = shard_range(int(mc_params.num_runs), rank, world_size)
    if rank == 0 and getattr(mc_params, "log_prefix", None):
        write_configuration(mc_params, mc_params.log_prefix)

    sim = BatchedGameSim(count, device=device, mode=mode, dt=mc_params.dt, winning_score=int(mc_params.winning_score),
                         noise=mc_params.noise_stdev, tau_puck=mc_params.tau_puck, tau_player=mc_params.tau_player,
                         player_mass=mc_params.player_mass, puck_mass=mc_params.puck_mass)
    lib = _capi.load()
    if randomize:
        x0 = torch.empty((count, 4), dtype=torch.float64, device=sim.device)
        # run index i (1-based, the reference's seed, py/MonteCarlo.py:23-30) = global game index
        _capi.check(lib.rgs_synth_inputs(sim.device.index or 0, int(seed), first + 1, count, 2, 0, float(u_max),
                                         x0.data_ptr() if count else None, None, 64, sim._stream()))
        sim.reset(x0_ball=x0)
    else:
        sim.reset(x0_ball=np.asarray(mc_params.x0_ball, dtype=np.float64))
    zero = torch.zeros((1, 4, 2), dtype=sim.action_dtype, device=sim.device)
    sim.run(zero, n_steps=1)  # priming step (py/MonteCarlo.py:33)
    if isinstance(policy, (str, dict)):
        named = {"classical": {}, "classical+cbf": {"cbf": 1}, "dmpc": {"kind": 1}, "dmpc+cbf": {"kind": 1, "cbf": 1}}
        fields = dict(named[policy]) if isinstance(policy, str) else dict(policy)
        fields.setdefault("input_limit", float(getattr(mc_params, "input_limit", 10.0)))
        if fields.get("cbf"):
            fields.setdefault("safety_radius", float(getattr(mc_params, "safety_radius", 0.4)))
            fields.setdefault("barrier_gain", float(getattr(mc_params, "barrier_gain", 30.0)))
        sim.run_policy(n_steps, **fields)
    elif policy is not None:
        for _ in range(n_steps):
            cmd = policy(sim.state())
            sim.run(torch.as_tensor(cmd).reshape(count, 1, 4, 2), n_steps=1)
    elif actions is not None:
        a = torch.as_tensor(actions)
        if a.dim() == 4:
            a = a[first:first + count]
        sim.run(a, n_steps=n_steps, hold=hold)
    else:
        sim.run_synth(seed, n_steps, hold=hold, u_max=u_max, first_game=first + 1, step0=0)
    stats = reduce_stats(sim.stats(), group)
    out = stats_to_dict(stats)
    out["n_steps"] = n_steps
    out["rank_runs"] = (first, count)
    if return_sim:
        out["sim"] = sim
    return out
