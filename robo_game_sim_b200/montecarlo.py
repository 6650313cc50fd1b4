"""Batched Monte Carlo harness: the role of py/MonteCarlo.py::monte_carlo (reference
py/MonteCarlo.py:8-49) with its inner `for run: for step: sim.run(...)` loop replaced by fused
rollout launches over a batch of games, and the offline MATLAB statistics
(matlab/interpret_mc_results.m:29-50) replaced by the on-device K4 reduction.

Multi-GPU: one process per GPU; the runs are split into contiguous slices (`shard_range`), every
per-run random stream is keyed by the GLOBAL run index so results do not depend on the number of
ranks, and the only collective is one all-reduce(sum) of the int64[96] statistics vector.
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
    """Sum the int64[96] statistics vector over all ranks (no-op without torch.distributed).
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
    # damage(aggressor, victim) summed over the runs (reference src/GameSim.cpp:277-282; the quantity
    # matlab/interpret_mc_results.m:37-48 histograms per run); 4x4 for 2v2 batches, team-level 2x2 always
    d["damage_sum"] = s[_capi.ST_DAMAGE:_capi.ST_DAMAGE + 16].reshape(4, 4) / _capi.DAMAGE_FIXED_SCALE
    d["damage_team_sum"] = s[_capi.ST_DAMAGE_TEAM:_capi.ST_DAMAGE_TEAM + 4].reshape(2, 2) / _capi.DAMAGE_FIXED_SCALE
    d["overlap_fixes"] = int(s[_capi.ST_OVERLAP_FIXES])
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

    if mc_params.noise_stdev and mode == "exact":
        mode = "fast64"   # process noise runs in the K2 formulation (statistical parity; see rgs_params.noise)
    sim = BatchedGameSim(count, device=device, mode=mode, first_game=first + 1, dt=mc_params.dt,
                         winning_score=int(mc_params.winning_score), noise=mc_params.noise_stdev, tau_puck=mc_params.tau_puck,
                         tau_player=mc_params.tau_player, player_mass=mc_params.player_mass, puck_mass=mc_params.puck_mass)
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


_ON_DEVICE = {"classical": 0, "dmpc": 1}


def _write_run_logs(sim, traj, live_before, log_dir, first_index, layout, append):
    """mc_run_{i}.log for every game of `sim` from the trajectory of ONE launch; a decided game stops logging
    (GameSim::run is a no-op then, src/GameSim.cpp:121-123), so only its live steps of this launch are written."""
    live_now = sim.events()[:, _capi.EV_NAMES.index("live_steps")]
    sim.write_logs(traj, os.path.join(log_dir, "mc_run_{}.log"), first_index=first_index, layout=layout,
                   n_records=(live_now - live_before), append=append)
    return live_now


def monte_carlo(mc_params, team_A_run_method, team_B_run_method, cbf_run_method=None, *, batched=False, mode=None,
                layout="reference", device=None, verbose=True, rank=None, world_size=None, group=None):
    """Drop-in for the reference's `monte_carlo(mc_params, team_A_run_method, team_B_run_method, cbf_run_method=None)`
    (py/MonteCarlo.py:8-49): same positional signature, same artefacts -- `<log_prefix>/configuration.txt` (13 lines,
    :14-21) and one binary `<log_prefix>/mc_run_{i}.log` per run i = 1..num_runs (29 doubles per record, the priming
    step included, :26-33) that matlab/interpret_mc_results.m and visualize_mc_game.m read -- and the per-run `reset`
    arguments of :28-30 (log on, seed = i).  Returns the statistics dict of `stats_to_dict` (the reference returns
    None and leaves the statistics to MATLAB).

    Three ways to supply the teams:
      * two CALLABLES (+ optional CBF callable), `batched=False` (default) -- exactly the reference's loop: the runs
        are played one after the other, every tick calls `team_A_run_method(SimState)`, `team_B_run_method(SimState)`
        on a per-game 22-vector view (and `cbf_run_method([uA1, uA2, uB1, uB2], SimState)`), so stateful controllers
        (py/src/ClassicalTeam.py replays a stored trajectory) behave as they do against the reference.  The step is
        the drop-in `GameSim` (batch of one, K1 fp64), which writes the log itself, byte for byte.
      * the same callables with `batched=True` -- all runs advance in lock step: the callables get ONE `SimState`
        over the [num_runs, 22] array and return ([num_runs, 2], [num_runs, 2]); one launch per tick.  Only for
        controllers that are functions of the state.
      * the strings "classical" / "dmpc" for both teams (+ `cbf_run_method="cbf"`): the ON-DEVICE surrogate of those
        teams, the whole closed loop of every run in one kernel launch (BatchedGameSim.run_policy).
    mode: "exact" | "fast" | "fast64" for the batched paths (default "exact" for callables, "fast" on device).
    layout: damage fields of the batched paths' logs, "reference" (byte-compatible) or "row_major".
    Multi-GPU (batched paths): runs are sharded by `shard_range`, statistics all-reduced; every rank writes its logs.
    """
    import torch
    import torch.distributed as dist

    from .gamesim import BatchedGameSim, GameSim
    from .simstate import SimState

    num_runs = int(mc_params.num_runs)
    n_steps = int(ceil(mc_params.T / mc_params.dt))
    log_dir = mc_params.log_prefix
    if rank is None:
        rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
    if world_size is None:
        world_size = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    if rank == 0:
        write_configuration(mc_params, log_dir)
    else:
        os.makedirs(log_dir, exist_ok=True)
    on_device = isinstance(team_A_run_method, str) or isinstance(team_B_run_method, str)
    common = dict(dt=mc_params.dt, winning_score=int(mc_params.winning_score), noise=mc_params.noise_stdev,
                  tau_puck=mc_params.tau_puck, tau_player=mc_params.tau_player, player_mass=mc_params.player_mass,
                  puck_mass=mc_params.puck_mass)

    if not on_device and not batched:
        # ---- the reference's loop, verbatim in structure (py/MonteCarlo.py:23-46) ----
        sim = GameSim() if device is None else GameSim(device)
        first, count = shard_range(num_runs, rank, world_size)
        finals = np.zeros((count, 22))
        dmg = np.zeros((count, 4, 4))
        for k, i in enumerate(range(first + 1, first + count + 1)):
            if verbose:
                print("MC RUN # {} / {}:".format(i, num_runs))
            logname = os.path.join(log_dir, "mc_run_{}.log".format(i))
            sim.reset(mc_params.dt, mc_params.winning_score, mc_params.x0_ball, mc_params.noise_stdev, True, logname, i,
                      mc_params.tau_puck, mc_params.tau_player, mc_params.player_mass, mc_params.puck_mass)
            z = np.zeros(2)
            sim_state = SimState(sim.run(z, z, z, z))
            for _ in range(n_steps):
                velA1, velA2 = team_A_run_method(sim_state)
                velB1, velB2 = team_B_run_method(sim_state)
                if cbf_run_method is not None:
                    velA1, velA2, velB1, velB2 = cbf_run_method([velA1, velA2, velB1, velB2], sim_state)
                sim_state = SimState(sim.run(velA1, velA2, velB1, velB2))
            finals[k] = sim_state.state
            dmg[k] = sim.damage()
        sim.reset(mc_params.dt, mc_params.winning_score, mc_params.x0_ball, mc_params.noise_stdev, False, "")  # closes the last log
        if verbose:
            print("DONE!")
        return _host_stats(finals, dmg, int(mc_params.winning_score), group)

    first, count = shard_range(num_runs, rank, world_size)
    mode = mode or ("fast" if on_device else "exact")
    if mc_params.noise_stdev and mode == "exact":
        mode = "fast64"   # process noise runs in the K2 formulation (statistical parity; see rgs_params.noise)
    sim = BatchedGameSim(count, device=device, mode=mode, first_game=first + 1, **common)   # run i <-> game index i
    sim.reset(x0_ball=np.asarray(mc_params.x0_ball, dtype=np.float64))
    zero = torch.zeros((1, 4, 2), dtype=sim.action_dtype, device=sim.device)
    live = torch.zeros(count, dtype=torch.int64, device=sim.device)
    traj = sim.run(zero, n_steps=1, log_stride=1)            # priming step (py/MonteCarlo.py:33), logged like any other
    live = _write_run_logs(sim, traj, live, log_dir, first + 1, layout, append=False)
    if on_device:
        if team_A_run_method != team_B_run_method or team_A_run_method not in _ON_DEVICE:
            raise ValueError("on-device teams: both team methods must be the same one of 'classical', 'dmpc'")
        if cbf_run_method not in (None, "cbf"):
            raise ValueError("on-device CBF filter: cbf_run_method must be None or 'cbf'")
        fields = {"kind": _ON_DEVICE[team_A_run_method], "input_limit": float(getattr(mc_params, "input_limit", 10.0))}
        if cbf_run_method:
            fields.update(cbf=1, safety_radius=float(getattr(mc_params, "safety_radius", 0.4)),
                          barrier_gain=float(getattr(mc_params, "barrier_gain", 30.0)))
        traj = sim.run_policy(n_steps, log_stride=1, **fields)
        _write_run_logs(sim, traj, live, log_dir, first + 1, layout, append=True)
    else:
        chunks = []
        for _ in range(n_steps):
            sim_state = SimState(sim.state().cpu().numpy())
            velA1, velA2 = team_A_run_method(sim_state)
            velB1, velB2 = team_B_run_method(sim_state)
            if cbf_run_method is not None:
                velA1, velA2, velB1, velB2 = cbf_run_method([velA1, velA2, velB1, velB2], sim_state)
            cmd = np.stack([np.broadcast_to(np.asarray(v, dtype=np.float64), (count, 2)) for v in (velA1, velA2, velB1, velB2)], axis=1)
            chunks.append(sim.run(torch.as_tensor(cmd).reshape(count, 1, 4, 2), n_steps=1, log_stride=1))
        _write_run_logs(sim, torch.cat(chunks, dim=1), live, log_dir, first + 1, layout, append=True)
    out = stats_to_dict(reduce_stats(sim.stats(), group))
    out["n_steps"] = n_steps
    out["rank_runs"] = (first, count)
    return out


def _host_stats(finals, damage, winning_score, group=None):
    """The K4 statistics vector for runs that were played one by one through the drop-in GameSim."""
    import torch

    s = np.zeros(_capi.N_STATS, dtype=np.int64)
    sa, sb = finals[:, 0], finals[:, 1]
    s[_capi.ST_GAMES] = len(finals)
    s[_capi.ST_WINS_A] = int((sa >= winning_score).sum())
    s[_capi.ST_WINS_B] = int(((sb >= winning_score) & (sa < winning_score)).sum())
    s[_capi.ST_UNDECIDED] = len(finals) - s[_capi.ST_WINS_A] - s[_capi.ST_WINS_B]
    s[_capi.ST_LEAD_A], s[_capi.ST_LEAD_B], s[_capi.ST_TIED] = int((sa > sb).sum()), int((sb > sa).sum()), int((sa == sb).sum())
    np.add.at(s, _capi.ST_SCORE_HIST + 5 * np.minimum(sa, 4).astype(int) + np.minimum(sb, 4).astype(int), 1)
    q = np.rint(damage * _capi.DAMAGE_FIXED_SCALE).astype(np.int64)
    s[_capi.ST_DAMAGE:_capi.ST_DAMAGE + 16] = q.sum(0).reshape(-1)
    for a in range(4):
        for v in range(4):
            s[_capi.ST_DAMAGE_TEAM + 2 * (a >= 2) + (v >= 2)] += q[:, a, v].sum()
    return stats_to_dict(reduce_stats(torch.from_numpy(s), group))
