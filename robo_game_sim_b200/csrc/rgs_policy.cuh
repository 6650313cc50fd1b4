// rgs_policy.cuh -- on-device SURROGATE policies for closed-loop Monte Carlo: the commands
// of all 2n players are computed from the game state inside the rollout kernel, every game-step, so a closed-loop
// batch never leaves the GPU.
//
// What is the reference's and what is not.  The reference's baseline team (py/src/ClassicalTeam.py) picks a play
// from the state and then asks pydrake optimisers for open-loop minimum-time trajectories towards a target state
// (py/src/ClassicalPlayer.py:95-110, 190-222).  pydrake is a controller dependency outside this path (and absent
// here), so the trajectory optimiser is replaced by a closed-form tracking law; the play logic and the target
// states are the reference's:
//   * play selection: defence iff the puck is in the own quarter and not moving away from the own goal
//     (ClassicalTeam.evaluateGame, py/src/ClassicalTeam.py:32-37);
//   * offence (:44-53): the field player kicks towards the adversary goal (simple_kick), the goalie defends while
//     the puck is in the own half and kicks otherwise;
//   * defence (:55-75): the field player clears the puck with 1.5 x the kick speed towards the side the opponents
//     are not on (defend_kick); the goalie defends -- the reference compares the distance of the ADVERSARY's first
//     player with itself (:65-72, `goalie_pos` is read from the adversary team), so its defend_kick branch is dead;
//   * targets: kick = touch the puck from behind along the shooting direction with the kick velocity
//     (get_final_state_for_kick, py/src/ClassicalPlayer.py:232-237; get_shoot_direction :239-244; defend_kick
//     :190-207), defend = half way between the own goal and the puck at rest (defend :209-222);
//   * kind = 1: the DMPC teams' play selection and target states x_des instead (DMPCTeam.execute,
//     py/src/DMPCTeam.py:78-93; DMPCPlayer.get_action, py/src/DMPCPlayer.py:27-57); the SNOPT collocation
//     (py/src/DMPC.py:16-54) that steers towards x_des is replaced by the same tracking law;
//   * surrogate tracking law (ours): u = clamp(v_f + gain (p_f - p), +-input_limit) on the first-order velocity plant.
// Optional safety filter: the reference's centralised control-barrier constraints for the damped double
// integrator (py/src/CBF.py:36-76: one half-space in u_i - u_j per approaching player pair, closed form) with
// the same class-K function alpha(h) = barrier_gain h^3 (:170-172); the QP that minimally modifies the nominal
// commands (:78-85, solved by pydrake there) is solved here by a fixed number of Hildreth sweeps over the pair
// constraints (exact for the half-space QP as the sweeps grow), followed by the input box.
//
// Plain scalar C++ (one IEEE operation per statement, no contraction), shared with the CPU emulator build.
#pragma once
#include "rgs_device.cuh"

// pair loops of the barrier filter: rolled, with the rows in local memory.  Measured on B200 (profiles/r01_ncu_policy_full.md):
// the policy kernel is instruction-fetch-bound, and the rolled form (4312 instead of 5360 SASS instructions for the
// 2v2 fp32 kernel) runs the filtered closed loop 9 % faster than the unrolled one (-DRGS_POLICY_UNROLLED, rows in
// registers), with identical results.
#ifdef RGS_POLICY_UNROLLED
#define RGS_POL_PAIRS _Pragma("unroll")
#else
#define RGS_POL_PAIRS _Pragma("unroll 1")
#endif

namespace rgs {

RGS_DEV float pol_sqrt(float x) { return sqrtf(x); }
RGS_DEV double pol_sqrt(double x) { return sqrt(x); }
RGS_DEV float pol_exp(float x) { return expf(x); }
RGS_DEV double pol_exp(double x) { return exp(x); }

template <typename T>
RGS_DEV T pol_clamp(T x, T lim) { return x > lim ? lim : (x < -lim ? -lim : x); }

// X(e, c): component c (px, py, vx, vy) of body e at the start of the step (players 0..2n-1, team A first; puck 2n)
// u[2*player + axis]: the commands (output)
template <typename T, int NT, typename StateFn>
RGS_DEV void surrogate_policy(const DevParams<T> &d, const PolicyArgs &pol, StateFn X, T *u) {
    constexpr int NP = 2 * NT, PK = 2 * NT;
    const T kv = T(pol.kick_velocity), gain = T(pol.gain), lim = T(pol.input_limit);
    const T px = X(PK, 0), py = X(PK, 1), pvx = X(PK, 2);
    const T reach = d.rad[0] + d.rad[1];
#pragma unroll
    for (int team = 0; team < 2; ++team) {
        const T f = team == 0 ? T(-1) : T(1);                // team A defends the left goal (mc_example_1.py:41-42)
        const bool defense = (f * px >= d.half_x * T(0.5)) && (f * pvx >= T(0));
        const T goal_x = -f * d.half_x, home_x = f * d.half_x;
        T sx = goal_x - px, sy = T(0) - py;                  // shooting direction
        const T sn = pol_sqrt(sx * sx + sy * sy);
        if (sn > T(1e-4)) { sx = sx / sn; sy = sy / sn; }
        const int o1 = (1 - team) * NT, o2 = o1 + (NT > 1 ? 1 : 0);
        const T ysum = X(o1, 1) + X(o2, 1);
        T kx = -f, ky = ysum > T(0) ? T(1) : (ysum < T(0) ? T(-1) : T(0));   // clearing direction
        const T kn = pol_sqrt(kx * kx + ky * ky);
        kx = kx / kn; ky = ky / kn;
#pragma unroll
        for (int q = 0; q < NT; ++q) {
            const int me = team * NT + q;
            int task;  // 0 defend, 1 kick at the goal, 2 clear, 3 / 4 DMPC hit with 1 x / 2 x the hit speed
            if (pol.kind == 1) {
                // DMPCTeam.execute :78-88 (thresholds arena_x - arena_x/4 and arena_x - arena_x/2 on |x| <= arena_x/2:
                // as written in the reference both are out of reach, the team always plays OFFENSE)
                const T ax = d.half_x + d.half_x;
                const bool dmpc_def = (f * px > ax - ax * T(0.25)) ||
                                      ((f * px > ax - ax * T(0.5)) && (f * pvx > ax / T(10)));
                const bool attacker = q != 1;     // A1 / B1 (get_action :27-57); the second player is the defender
                if (!dmpc_def) task = attacker ? 3 : (f * px > T(0) ? 0 : 3);
                else task = attacker ? 4 : 0;
            } else if (!defense) task = (q == 0 && f * px >= T(0)) ? 0 : 1;
            else task = q == 0 ? 0 : 2;
            T pfx, pfy, vfx, vfy;
            if (task >= 3) {   // x_des = [puck, scale v_hit (goal - puck) / |goal - puck|]
                T hx = goal_x - px, hy = T(0) - py;
                if (sn > T(0)) { hx = hx / sn; hy = hy / sn; }
                const T sc = task == 4 ? T(2) * kv : kv;
                pfx = px; pfy = py;
                vfx = sc * hx; vfy = sc * hy;
            } else if (task == 0) {
                pfx = home_x + T(0.5) * (px - home_x); pfy = T(0.5) * py;
                vfx = T(0); vfy = T(0);
            } else if (task == 1) {
                pfx = px - sx * reach; pfy = py - sy * reach;
                vfx = kv * sx; vfy = kv * sy;
            } else {
                const T kk = T(1.5) * kv;
                pfx = px - kx * reach; pfy = py - ky * reach;
                vfx = kx * kk; vfy = ky * kk;
            }
            u[2 * me] = pol_clamp(vfx + gain * (pfx - X(me, 0)), lim);
            u[2 * me + 1] = pol_clamp(vfy + gain * (pfy - X(me, 1)), lim);
        }
    }
    if (!pol.cbf) return;
    // ---- control-barrier filter: a_k . (u_i - u_j) >= b_k for every approaching pair (CBF.py:46-76) ----
    constexpr int NC = NP * (NP - 1) / 2;
    T ax[NC], ay[NC], b[NC], lam[NC];
    bool on[NC];
    const T tau = d.tau[0], amax = T(2) * lim, Ds = T(pol.safety_radius), gam = T(pol.barrier_gain);
    {
        int k = 0;
RGS_POL_PAIRS
        for (int i = 0; i < NP - 1; ++i)
RGS_POL_PAIRS
            for (int j = i + 1; j < NP; ++j, ++k) {
                const T dx = X(i, 0) - X(j, 0), dy = X(i, 1) - X(j, 1);
                const T dvx = X(i, 2) - X(j, 2), dvy = X(i, 3) - X(j, 3);
                const T n = pol_sqrt(dx * dx + dy * dy);
                const T ex = n > T(0) ? dx / n : dx, ey = n > T(0) ? dy / n : dy;
                const T vcol = ex * dvx + ey * dvy;
                on[k] = vcol < T(0) && n > T(0);
                lam[k] = T(0);
                ax[k] = ex; ay[k] = ey; b[k] = T(0);
                if (on[k]) {
                    const T ef = pol_exp((n + tau * vcol - Ds) / (tau * amax));
                    const T w = (dvx * dvx + dvy * dvy) - vcol * vcol;
                    const T h = (amax - vcol) * ef - amax;
                    // hij_dot = c0 + c1 (e . u_ij)  with  c1 = -ef vcol / (tau amax) > 0
                    const T c0 = ef * (-w / n + (amax - vcol) / (tau * amax) * (vcol + tau / n * w) + vcol / (tau * amax) * vcol);
                    const T c1 = -(ef * vcol) / (tau * amax);
                    b[k] = (-(gam * h * h * h) - c0) / c1;
                }
            }
    }
#pragma unroll 1
    for (int sweep = 0; sweep < pol.sweeps; ++sweep) {
        int k = 0;
RGS_POL_PAIRS
        for (int i = 0; i < NP - 1; ++i)
RGS_POL_PAIRS
            for (int j = i + 1; j < NP; ++j, ++k) {
                if (!on[k]) continue;
                const T au = ax[k] * (u[2 * i] - u[2 * j]) + ay[k] * (u[2 * i + 1] - u[2 * j + 1]);
                T nl = lam[k] + (b[k] - au) * T(0.5);   // |(e, -e)|^2 = 2
                if (nl < T(0)) nl = T(0);
                const T dl = nl - lam[k];
                lam[k] = nl;
                u[2 * i] = u[2 * i] + dl * ax[k]; u[2 * i + 1] = u[2 * i + 1] + dl * ay[k];
                u[2 * j] = u[2 * j] - dl * ax[k]; u[2 * j + 1] = u[2 * j + 1] - dl * ay[k];
            }
    }
#pragma unroll
    for (int c = 0; c < 2 * NP; ++c) u[c] = pol_clamp(u[c], lim);
}

}  // namespace rgs
