"""On-device surrogate policy (robo_game_sim_b200/csrc/rgs_policy.cuh), CPU side.

The policy body is the kernel's own source compiled by the emulator build (tests/emu).  It is checked against
  * a numpy restatement of the reference's play logic and target states, written from the reference files
    (py/src/ClassicalTeam.py:32-75, py/src/ClassicalPlayer.py:190-244) with the surrogate tracking law on top;
  * the reference's control-barrier constraints in the form CBF.py writes them (hij_dot >= -alpha(hij),
    py/src/CBF.py:46-76) and, for the minimal-modification QP (:78-85), scipy's SLSQP solution;
  * itself: a closed-loop rollout inside the kernel body == stepping the game from the host with the commands
    the policy returns for every state (fp64 formulation, bit for bit).
pydrake (the reference's trajectory optimiser and QP solver) is absent: parity of the optimiser is out of scope."""
import math

import numpy as np
import pytest

import rgs_testlib as tl

ARENA_X, R_PLAYER, R_PUCK, TAU = 10.0, 0.2, 0.175, 0.5


def _unit(v):
    n = math.sqrt(v[0] * v[0] + v[1] * v[1])
    return n


def ref_policy(state, pol):
    """state record [TAS, TBS, puck(4), A1, A2, B1, B2] -> nominal commands [4, 2]."""
    puck = state[2:6]
    pl = state[6:22].reshape(4, 4)
    u = np.zeros((4, 2))
    reach = R_PLAYER + R_PUCK
    for team, field in ((0, -1.0), (1, 1.0)):
        # ClassicalTeam.evaluateGame :32-37
        defense = field * puck[0] >= (ARENA_X / 2.0) * 0.5 and field * puck[2] >= 0
        goal = np.array([-field * (ARENA_X / 2.0), 0.0])   # get_adversary_goal_pos :218-223
        home = np.array([field * (ARENA_X / 2.0), 0.0])    # get_home_goal_pos :225-230
        shoot = goal - puck[:2]                            # get_shoot_direction :239-244
        n = _unit(shoot)
        if n > 1e-4:
            shoot = shoot / n
        opp = pl[(1 - team) * 2:(1 - team) * 2 + 2]
        clear = np.array([-field, np.sign(opp[0][1] + opp[1][1])])   # defend_kick :198-200
        clear = clear / _unit(clear)
        for q in range(2):
            me = pl[team * 2 + q]
            if not defense:   # ClassicalTeam.execute :44-53
                task = "defend" if (q == 0 and field * puck[0] >= 0) else "kick"
            else:             # :55-75 (the goalie's defend_kick branch is dead in the reference)
                task = "defend" if q == 0 else "clear"
            if task == "defend":   # ClassicalPlayer.defend :209-222
                pf, vf = home + 0.5 * (puck[:2] - home), np.zeros(2)
            elif task == "kick":   # simple_kick :95-110 + get_final_state_for_kick :232-237
                pf, vf = puck[:2] - shoot * reach, pol["kick_velocity"] * shoot
            else:                  # defend_kick :190-207 with 1.5 x kick velocity (ClassicalTeam.py:58)
                pf, vf = puck[:2] - clear * reach, clear * (1.5 * pol["kick_velocity"])
            u[team * 2 + q] = np.clip(vf + pol["gain"] * (pf - me[:2]), -pol["input_limit"], pol["input_limit"])
    return u


def cbf_margin(state, u, pol):
    """min over approaching ordered pairs of hij_dot + alpha(hij), exactly as CBF.py:46-76 writes it."""
    pl = state[6:22].reshape(4, 4)
    a_rel_max, Ds = 2 * pol["input_limit"], pol["safety_radius"]
    worst = np.inf
    for i in range(4):
        for j in range(4):
            if i == j:
                continue
            pij, vij = pl[i][:2] - pl[j][:2], pl[i][2:] - pl[j][2:]
            norm = np.linalg.norm(pij)
            unit = pij / norm
            vcol = unit.dot(vij)
            if vcol < 0:
                uij = u[i] - u[j]
                ef = np.exp((norm + TAU * vcol - Ds) / (TAU * a_rel_max))
                w = vij.dot(vij) - vcol ** 2
                hdot = ef * (-w / norm + (a_rel_max - vcol) / (TAU * a_rel_max) * (vcol + TAU / norm * w)
                             + vcol / (TAU * a_rel_max) * (vcol - unit.dot(uij)))
                h = (a_rel_max - vcol) * ef - a_rel_max
                worst = min(worst, hdot + pol["barrier_gain"] * h ** 3)
    return worst


def random_states(n, seed, close=False):
    rng = np.random.default_rng(seed)
    s = np.zeros((n, 22))
    s[:, 2:4] = rng.uniform([-4.5, -2.2], [4.5, 2.2], size=(n, 2))
    s[:, 4:6] = rng.uniform(-3, 3, size=(n, 2))
    for p in range(4):
        spread = 0.8 if close else 1.0
        s[:, 6 + 4 * p:8 + 4 * p] = rng.uniform([-4.5 * spread, -2.2 * spread], [4.5 * spread, 2.2 * spread], size=(n, 2))
        s[:, 8 + 4 * p:10 + 4 * p] = rng.uniform(-4, 4, size=(n, 2))
    if close:   # a second player of every game sits near the first one and runs into it
        s[:, 10:12] = s[:, 6:8] + rng.uniform(0.5, 1.2, size=(n, 2)) * rng.choice([-1, 1], size=(n, 2))
        s[:, 12:14] = s[:, 8:10] + 2.0 * (s[:, 6:8] - s[:, 10:12])
    return s


def test_nominal_commands_follow_the_reference_play_logic():
    pd, pol = tl.default_params(mode=tl.MODE_FAST_F64), tl.default_policy()
    states = random_states(400, 1)
    states[:50, 4] = 0.0            # puck at rest in x: both '>= 0' tests of evaluateGame on their boundary
    states[50:60, 2:4] = [5.0, 0.0]  # puck on the goal centre: zero shooting direction for team A (norm <= 1e-4)
    got = tl.emu_policy_commands(pd, pol, states)
    want = np.stack([ref_policy(s, pol) for s in states])
    assert np.abs(got - want).max() <= 1e-12
    plays = [(s[2] >= 2.5 and s[4] >= 0, -s[2] >= 2.5 and -s[4] >= 0) for s in states]
    assert any(a for _, a in plays) and any(b for b, _ in plays) and any(not a and not b for b, a in plays)
    assert (np.abs(got) == pol["input_limit"]).any() and (np.abs(got) < pol["input_limit"]).any()


def test_barrier_filter_satisfies_the_reference_constraints_and_its_qp():
    from scipy.optimize import minimize

    pd = tl.default_params(mode=tl.MODE_FAST_F64)
    nominal, filt = tl.default_policy(), tl.default_policy(cbf=1, sweeps=400)
    states = random_states(400, 2, close=True)
    u0 = tl.emu_policy_commands(pd, nominal, states)
    u1 = tl.emu_policy_commands(pd, filt, states)
    changed = 0
    for s, a, b in zip(states, u0, u1):
        m0, m1 = cbf_margin(s, a, filt), cbf_margin(s, b, filt)
        inside = np.abs(b).max() < filt["input_limit"] - 1e-9
        if m0 >= 0:                       # nominal commands already safe: untouched (CBF.py:78-80, zero cost)
            assert np.abs(a - b).max() <= 1e-9
            continue
        changed += 1
        if inside:                        # input box inactive: the filtered commands satisfy every barrier row ...
            assert m1 >= -1e-6 * max(1.0, abs(m0))
            # ... and are the QP's minimiser: compare with SLSQP on the same constraints
            cons = [{"type": "ineq", "fun": lambda x, s=s: cbf_margin(s, x.reshape(4, 2), filt)}]
            res = minimize(lambda x: ((x - a.ravel()) ** 2).sum(), b.ravel(), constraints=cons, method="SLSQP",
                           options={"ftol": 1e-12, "maxiter": 200})
            cost_ours, cost_qp = ((b - a) ** 2).sum(), res.fun
            assert cost_ours <= cost_qp * (1 + 1e-3) + 1e-9
    assert changed >= 20
    # the shipped sweep count is close to the converged filter
    u8 = tl.emu_policy_commands(pd, tl.default_policy(cbf=1, sweeps=8), states)
    assert np.abs(u8 - u1).max() <= 0.05


@pytest.mark.parametrize("cbf", [0, 1])
def test_closed_loop_in_the_kernel_equals_stepping_from_the_host(cbf):
    pd = tl.default_params(mode=tl.MODE_FAST_F64, winning_score=2)
    pol = tl.default_policy(cbf=cbf)
    x0, _ = tl.synth_inputs(21, 0, 24, 2, 1)
    s0 = tl.oracle_reset(tl.default_params(), x0)
    n_steps = 60
    a = tl.emu_rollout(pd, s0, None, n_steps, policy=pol)
    st, t, dmg, ev = s0, None, None, None
    for _ in range(n_steps):
        cmd = tl.emu_policy_commands(pd, pol, st)[:, None]          # [n, 1, 4, 2]
        r = tl.emu_rollout(pd, st, cmd, 1, 1, t=t, damage=dmg, events=ev)
        st, t, dmg, ev = r["state"], r["t"], r["damage"], r["events"]
    assert np.array_equal(a["state"], st) and np.array_equal(a["events"], ev) and np.array_equal(a["damage"], dmg)
    assert a["events"][:, tl.EV["PLAYER_PUCK"]].sum() > 0     # the teams do reach the puck
    assert a["state"][:, :2].sum() > 0                        # and score


def test_policy_games_are_played_out():
    """Behaviour: both surrogate teams chase the puck and score; the barrier filter removes most player-player impacts."""
    pd = tl.default_params(mode=tl.MODE_FAST_F32, winning_score=3)
    x0, _ = tl.synth_inputs(22, 0, 200, 2, 1)
    s0 = tl.oracle_reset(tl.default_params(), x0)
    free = tl.emu_rollout(pd, s0, None, 200, policy=tl.default_policy())
    safe = tl.emu_rollout(pd, s0, None, 200, policy=tl.default_policy(cbf=1))
    goals = free["state"][:, :2].sum(axis=0)
    assert goals[0] >= 5 and goals[1] >= 5 and goals.sum() >= 20                 # both sides score
    pp_free, pp_safe = free["events"][:, tl.EV["PLAYER_PLAYER"]].sum(), safe["events"][:, tl.EV["PLAYER_PLAYER"]].sum()
    assert pp_free > 1000 and pp_safe < 0.1 * pp_free


def test_dmpc_kind_follows_the_dmpc_play_logic_and_targets():
    """kind = 1: play and x_des of DMPCTeam.execute / DMPCPlayer.get_action -- restated on the host in
    robo_game_sim_b200/sampler.py (dmpc_play, dmpc_target), which test_capi_and_host.py checks against the reference
    text -- with the surrogate tracking law on top."""
    from robo_game_sim_b200 import sampler as sp

    pd, pol = tl.default_params(mode=tl.MODE_FAST_F64), tl.default_policy(kind=1, gain=3.0)
    states = random_states(300, 4)
    states[:40, 2] = np.linspace(-4.9, 4.9, 40)
    got = tl.emu_policy_commands(pd, pol, states)
    for s, u in zip(states, got):
        for team, field in ((0, -1), (1, 1)):
            play = sp.dmpc_play(s, field)
            assert play == sp.OFFENSE      # the reference's thresholds are out of the arena's reach (DMPCTeam.py:81-84)
            goal, home = [-field * 5.0, 0.0], [field * 5.0, 0.0]
            for q, pid in ((0, sp.ATTACKER), (1, sp.DEFENDER)):
                xd = sp.dmpc_target(s, play, pid, field, goal, home, v_hit=pol["kick_velocity"])
                me = s[6 + 4 * (2 * team + q):8 + 4 * (2 * team + q)]
                want = np.clip(xd[2:] + pol["gain"] * (xd[:2] - me), -pol["input_limit"], pol["input_limit"])
                assert np.abs(u[2 * team + q] - want).max() <= 1e-12
    # and it plays: closed loop with the barrier filter on, fp32
    x0, _ = tl.synth_inputs(23, 0, 100, 2, 1)
    r = tl.emu_rollout(dict(pd, mode=tl.MODE_FAST_F32, winning_score=3), tl.oracle_reset(tl.default_params(), x0), None, 200,
                       policy=tl.default_policy(kind=1, cbf=1))
    assert r["events"][:, tl.EV["PLAYER_PUCK"]].sum() > 100 and r["state"][:, :2].sum() >= 5


@pytest.mark.parametrize("n_team", [1, 3, 4])
def test_policy_for_other_team_sizes(n_team):
    """The surrogate teams for n != 2 (roles of ClassicalTeam.py:44-75 generalised: player 1 of a team is the goalie,
    the other n-1 are field players): the in-kernel closed loop equals stepping from the host with the policy's
    commands (fp64, bit for bit, with and without the barrier filter), the field players follow the 2v2 field
    player's law, and games are played out."""
    pd = tl.default_params(n_team=n_team, mode=tl.MODE_FAST_F64, winning_score=2)
    x0, _ = tl.synth_inputs(23, 0, 12, n_team, 1)
    s0 = tl.oracle_reset(tl.default_params(n_team=n_team), x0)
    for cbf in (0, 1):
        pol = tl.default_policy(cbf=cbf)
        a = tl.emu_rollout(pd, s0, None, 40, policy=pol)
        st, t, dmg, ev = s0, None, None, None
        for _ in range(40):
            cmd = tl.emu_policy_commands(pd, pol, st)[:, None]
            r = tl.emu_rollout(pd, st, cmd, 1, 1, t=t, damage=dmg, events=ev)
            st, t, dmg, ev = r["state"], r["t"], r["damage"], r["events"]
        assert np.array_equal(a["state"], st) and np.array_equal(a["events"], ev)
    # every field player (q >= 1) of a team gets the command the 2v2 field player would get at its position
    pol = tl.default_policy()
    st = s0.copy()
    st[:, 2:6] = np.random.default_rng(1).uniform([-4, -2, -2, -2], [4, 2, 2, 2], size=(12, 4))
    u = tl.emu_policy_commands(pd, pol, st)
    if n_team >= 3:
        pd2 = tl.default_params(mode=tl.MODE_FAST_F64)
        for team in (0, 1):
            for q in range(1, n_team):
                s2 = np.zeros((12, 22))
                s2[:, :6] = st[:, :6]
                for tt in (0, 1):       # goalie and the probed field player of each team
                    s2[:, 6 + 8 * tt:10 + 8 * tt] = st[:, 6 + 4 * (tt * n_team):10 + 4 * (tt * n_team)]
                    s2[:, 10 + 8 * tt:14 + 8 * tt] = st[:, 6 + 4 * (tt * n_team + q):10 + 4 * (tt * n_team + q)]
                # the clearing direction looks at the adversary's first two players: same in both layouts
                s2[:, 10 + 8 * (1 - team):14 + 8 * (1 - team)] = st[:, 6 + 4 * ((1 - team) * n_team + 1):10 + 4 * ((1 - team) * n_team + 1)]
                u2 = tl.emu_policy_commands(pd2, pol, s2)
                assert np.allclose(u[:, team * n_team + q], u2[:, team * 2 + 1], atol=1e-12)
    played = tl.emu_rollout(dict(pd, mode=tl.MODE_FAST_F32, winning_score=3), s0, None, 200, policy=pol)
    assert played["events"][:, tl.EV["PLAYER_PUCK"]].sum() > 0 and np.isfinite(played["state"]).all()
