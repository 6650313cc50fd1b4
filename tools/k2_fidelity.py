#!/usr/bin/env python
"""tools/k2_fidelity.py -- how far is the K2 formulation (fp64 instantiation, CPU emulator build of the kernel
body) from the oracle?  TEST-SIDE tool: uses tests/rgs_testlib (oracle + emulator), never the product package.

  single-step mode: the oracle is rolled out with trajectory recording; every recorded state is then stepped ONCE by
  both the emulator (fast64) and the oracle with the clock at 0, so clock rounding is excluded and errors do not
  accumulate.  Reports the share of steps whose event-hash differs and the state error on hash-identical steps.

  rollout mode: whole rollouts against oracle_rollout (absolute clock) and oracle_rollout_reltime.

usage: k2_fidelity.py [--n 300] [--steps 40] [--nt 2] [--dt 0.05] [--tau-puck 0.1] [--umax 10] [--rollout]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import rgs_testlib as tl  # noqa: E402


def single_step(pd, n, steps, hold, seed, umax, verbose=False, abs_clock=False):
    nt = pd["players_per_team"]
    x0, acts = tl.synth_inputs(seed, 0, n, nt, (steps + hold - 1) // hold, u_max=umax)
    st0 = tl.oracle_reset(pd, x0)
    o = tl.oracle_rollout(pd, st0, acts, steps, hold, traj=True)
    states = np.concatenate([st0[:, None], o["traj"][:, :-1]], axis=1)  # state BEFORE step k
    mism = same_err = total = 0
    worst = 0.0
    bad = []
    ev_o_tot = np.zeros(tl.N_EVENTS)
    ev_e_tot = np.zeros(tl.N_EVENTS)
    tk = 0.0
    for k in range(steps):
        t_now, tk = tk, tk + pd["dt"]   # the reference clock: t_ += dt_ per step (src/GameSim.cpp:158)
        a = acts[:, k // hold:k // hold + 1]
        sk = np.ascontiguousarray(states[:, k])
        live = (sk[:, 0] < pd["winning_score"]) & (sk[:, 1] < pd["winning_score"])
        sk, a = sk[live], a[live]
        if len(sk) == 0:
            continue
        tt = np.full(len(sk), t_now if abs_clock else 0.0)
        ro = tl.oracle_rollout(pd, sk, a, 1, 1, t=tt)
        re = tl.emu_rollout(dict(pd, mode=tl.MODE_FAST_F64), sk, a, 1, 1, t=tt)
        same = ro["events"][:, 14] == re["events"][:, 14]
        err = np.abs(ro["state"] - re["state"]).max(1)
        total += len(sk)
        mism += int((~same).sum())
        same_err += int((same & (err > 1e-6)).sum())
        worst = max(worst, float(err[same].max()) if same.any() else 0.0)
        ev_o_tot += ro["events"].sum(0)
        ev_e_tot += re["events"].sum(0)
        if verbose:
            for g in np.nonzero(~same)[0][:3]:
                bad.append((k, int(np.nonzero(live)[0][g]), ro["events"][g, :10].tolist(), re["events"][g, :10].tolist(),
                            float(err[g])))
    return dict(total=total, mismatch=mism, same_hash_state_err=same_err, worst_same_err=worst, bad=bad,
                ev_o=ev_o_tot, ev_e=ev_e_tot)


def rollout(pd, n, steps, hold, seed, umax):
    nt = pd["players_per_team"]
    x0, acts = tl.synth_inputs(seed, 0, n, nt, (steps + hold - 1) // hold, u_max=umax)
    st0 = tl.oracle_reset(pd, x0)
    o = tl.oracle_rollout(pd, st0, acts, steps, hold)
    r = tl.oracle_rollout_reltime(pd, st0, acts, steps, hold)
    e = tl.emu_rollout(dict(pd, mode=tl.MODE_FAST_F64), st0, acts, steps, hold)
    out = {}
    for name, ref in (("abs", o), ("rel", r)):
        same = ref["events"][:, 14] == e["events"][:, 14]
        out[name] = dict(hash_same=float(same.mean()),
                         ints_equal_on_same=bool(np.array_equal(ref["events"][same][:, :14], e["events"][same][:, :14])),
                         median_err_same=float(np.median(np.abs(ref["state"] - e["state"])[same].max(1))) if same.any() else None,
                         ev_ref=ref["events"].sum(0)[:14].tolist(), ev_emu=e["events"].sum(0)[:14].tolist())
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=300)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--hold", type=int, default=10)
    ap.add_argument("--nt", type=int, default=2)
    ap.add_argument("--dt", type=float, default=0.05)
    ap.add_argument("--tau-puck", type=float, default=0.1)
    ap.add_argument("--tau-player", type=float, default=0.5)
    ap.add_argument("--player-mass", type=float, default=1.0)
    ap.add_argument("--puck-mass", type=float, default=0.5)
    ap.add_argument("--umax", type=float, default=10.0)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--ws", type=int, default=1 << 30)
    ap.add_argument("--rollout", action="store_true")
    ap.add_argument("--abs-clock", action="store_true", help="single-step mode: both sides start from the reference clock k*dt")
    ap.add_argument("-v", action="store_true")
    a = ap.parse_args()
    pd = tl.default_params(n_team=a.nt, winning_score=a.ws, dt=a.dt, tau_puck=a.tau_puck, tau_player=a.tau_player,
                           player_mass=a.player_mass, puck_mass=a.puck_mass)
    if a.rollout:
        r = rollout(pd, a.n, a.steps, a.hold, a.seed, a.umax)
        for k, v in r.items():
            print(k, v)
        return
    r = single_step(pd, a.n, a.steps, a.hold, a.seed, a.umax, a.v, a.abs_clock)
    print(f"nt={a.nt} dt={a.dt} tau_puck={a.tau_puck}: steps {r['total']}  hash mismatch {r['mismatch']} "
          f"({r['mismatch'] / max(r['total'], 1):.2e})  same-hash state err>1e-6: {r['same_hash_state_err']}  "
          f"worst same-hash err {r['worst_same_err']:.2e}")
    names = list(tl.EV)
    print("  oracle events:", {names[i]: int(r["ev_o"][i]) for i in (2, 3, 4, 5, 6, 7, 8, 9, 12, 15)})
    print("  emu    events:", {names[i]: int(r["ev_e"][i]) for i in (2, 3, 4, 5, 6, 7, 8, 9, 12, 15)})
    for b in r["bad"][:20]:
        print("  bad step", b)


if __name__ == "__main__":
    main()
