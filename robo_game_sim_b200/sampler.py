"""Shared-state rollout sampling for MPC (BASELINE.json configs[4]).

The reference's DMPC / CMPC controllers solve a direct-collocation programme over an internal LINEAR model
with SNOPT at every tick (reference py/src/DMPC.py:16-54, py/src/CentralizedMPC.py:17-72) and never roll out
the simulator.  Here the same decision -- "which command should my players apply now to approach x_des?" --
is taken by rolling C candidate command sequences through the real GameSim step (collisions, walls, goals)
on the GPU from the current state, scoring the simulated terminal state with the reference's terminal
weight Omega_N_max = diag(10, 10, 20, 20) (py/src/DMPCTeam.py:56-59) and taking the arg-min.

Host side only builds targets and candidate tensors; rollouts, costs and the arg-min run in
librgs_b200.so (rgs_batch_reset_shared / rgs_batch_run / rgs_batch_terminal_cost).  No CPU fallback.
"""
import numpy as np

OMEGA_N_MAX = (10.0, 10.0, 20.0, 20.0)      # reference py/src/DMPCTeam.py:56-59
V_HIT = 4.0                                  # reference py/src/DMPCTeam.py:73
OFFENSE, DEFENSE = 0, 1
ATTACKER, DEFENDER = 1, 2                    # player ids, reference py/src/DMPCPlayer.py


def dmpc_play(state22, field, arena_x=10.0):
    """OFFENSE / DEFENSE decision of DMPCTeam.execute (reference py/src/DMPCTeam.py:78-88).
    field = +1 for the team defending the +x goal side as the reference passes it, -1 otherwise."""
    px, vx = state22[2], state22[4]
    v_thresh = arena_x / 10.0                # :69
    d_off, d_def = arena_x / 2.0, arena_x / 4.0   # :70-71
    if field * px > arena_x - d_def:
        return DEFENSE
    if field * px > arena_x - d_off and field * vx > v_thresh:
        return DEFENSE
    return OFFENSE


def dmpc_target(state22, play, player_id, field, goal_pos, home_pos, v_hit=V_HIT):
    """x_des of DMPCPlayer.get_action (reference py/src/DMPCPlayer.py:27-55): [px, py, vx, vy]."""
    puck = np.asarray(state22[2:4], dtype=np.float64)
    goal_pos = np.asarray(goal_pos, dtype=np.float64)
    home_pos = np.asarray(home_pos, dtype=np.float64)

    def hit(scale):
        d = goal_pos - puck
        d = scale * v_hit * d / np.linalg.norm(d)
        return np.array([puck[0], puck[1], d[0], d[1]])

    def defend():
        p = home_pos + (puck - home_pos) / 2.0
        return np.array([p[0], p[1], 0.0, 0.0])

    if play == OFFENSE:
        if player_id == ATTACKER:
            return hit(1.0)
        return defend() if field * puck[0] > 0 else hit(1.0)
    if player_id == ATTACKER:
        return hit(2.0)
    return defend()


def candidate_actions(n_candidates, n_act, players=(0, 1), n_players=4, u_max=10.0, seed=0, others=None, dtype=np.float32):
    """[C, n_act, n_players, 2] command tensor: the sampled players' commands ~ U(-u_max, u_max)^2 (the
    reference's input box, py/run_sim.py:23), every other player follows `others` [n_act, n_players, 2] (zeros
    by default).  Candidate 0 is all-zero for the sampled players (the do-nothing baseline)."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    a = np.zeros((n_candidates, n_act, n_players, 2), dtype=np.float64)
    if others is not None:
        a[:] = np.asarray(others, dtype=np.float64)[None]
    for p in players:
        a[:, :, p, :] = rng.uniform(-u_max, u_max, size=(n_candidates, n_act, 2))
        a[0, :, p, :] = 0.0
    return a.astype(dtype)


def sample_rollouts(sim, state22, actions, x_des, weights=None, n_steps=50, hold=10, players=(0, 1)):
    """Roll every candidate of `actions` [C, n_act, 2n, 2] out for n_steps from `state22` on `sim` (a
    BatchedGameSim with C games) and return (best_index, first_command_of_best [len(players), 2], cost tensor).
    x_des: {player index: [4]} targets; weights default to OMEGA_N_MAX for the players in x_des."""
    import torch

    npl = 2 * sim.n_team
    xd = np.zeros((npl, 4))
    w = np.zeros((npl, 4))
    for p, t in x_des.items():
        xd[p] = np.asarray(t, dtype=np.float64)
        w[p] = OMEGA_N_MAX if weights is None else np.asarray(weights[p], dtype=np.float64)
    acts = torch.as_tensor(actions)
    if acts.shape[0] != sim.n_games:
        raise ValueError(f"actions holds {acts.shape[0]} candidates, the batch {sim.n_games} games")
    acts = acts.to(device=sim.device, dtype=sim.action_dtype)
    sim.reset_shared(np.asarray(state22, dtype=np.float64))
    sim.run(acts, n_steps=n_steps, hold=hold)
    cost, best, _ = sim.terminal_cost(xd, w)
    first = acts[best, 0, list(players), :].to(torch.float64).cpu().numpy()
    return best, first, cost
