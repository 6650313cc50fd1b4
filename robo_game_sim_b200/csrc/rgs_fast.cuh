// rgs_fast.cuh -- K2 (warp body): fused multi-step rollout in the THROUGHPUT formulation.
//
// Same event semantics as the reference step (GameSim::updateSim, reference src/GameSim.cpp:126-171:
// overlap detected at the 50 grid columns, linear back-tracking from column i-1, per-owner events,
// earliest-first resolution with the stable-sort tie rule, prune + re-scan, one goal per step), but
//   * time is step-relative (t_ = 0 at the start of every step);
//   * the 50-sub-step RK4 chain (populateStateGrid :173-217, RK4_player/RK4_puck :70-116) is replaced
//     by its own closed form: the dynamics are linear with a constant command, so i RK4 sub-steps
//     are the affine map  v_i = u + (v0-u) R^i,  p_i = p0 + u i h + (v0-u) a (1-R^i)/(1-R)
//     with R = 1+z+z^2/2+z^3/6+z^4/24, a = h (1+z/2+z^2/6+z^3/24), z = -h drag/tau  (exactly what
//     RK4 computes, up to rounding).  Every column is therefore independent of the previous one
//     and the 32 lanes of the warp own COLUMNS: lane l evaluates columns l+1 and l+33;
//   * the shared-memory grid is only materialised in steps where some column shows an overlap
//     (wall or pair); the common no-contact step runs entirely in registers:
//     2 FMA per coordinate per column, |p| >= limit wall tests, squared-distance pair tests,
//     one ballot, closed-form commit of column 50.
// Template T = float is the throughput mode (RGS_MODE_FAST_F32); T = double (RGS_MODE_FAST_F64)
// exists to validate this formulation against the oracle event for event.
//
// All fused multiply-adds are spelled out (fm()), everything else is a single IEEE operation and
// the file is compiled with -fmad=false, so the arithmetic is the same wherever this source is
// built (nvcc for sm_100a; g++ -ffp-contract=off for the CPU-side emulator check under tests/emu).
#pragma once
#include "rgs_device.cuh"

namespace rgs {

#ifdef RGS_HOST_EMU
static inline float fm(float a, float b, float c) { return std::fmaf(a, b, c); }
static inline double fm(double a, double b, double c) { return std::fma(a, b, c); }
#else
RGS_DEV float fm(float a, float b, float c) { return __fmaf_rn(a, b, c); }
RGS_DEV double fm(double a, double b, double c) { return __fma_rn(a, b, c); }
#endif

// Per-type ([0] player, [1] puck) closed-form coefficients, built on the host (rgs_host_params.h).
template <typename T>
struct FastParams {
    DevParams<T> d;
    T neg_inv_tau[2];       // -drag/tau : z = h * neg_inv_tau
    T tab_g[2][kCols];      // R^i
    T tab_k[2][kCols];      // a (1-R^i)/(1-R)
    T lim_x[2], lim_y[2];   // arena/2 - r (folded wall-overlap limits for the register phase)
    T lim_x_in[2], lim_y_in[2];  // largest value strictly below lim (lane kernel: a corrected body never rests ON the limit)
};

template <typename T>
struct FastTables {         // block-shared copy of the coefficient tables (lane-indexed reads)
    T g[2][kCols];
    T k[2][kCols];
};

// advance one axis by an arbitrary h (time-of-impact split), same RK4 polynomial
template <typename T>
RGS_DEV void advance_axis(T &p, T &v, T u, T h, T neg_inv_tau) {
    const T z = h * neg_inv_tau;
    T q = fm(z, T(0.25), T(1));
    q = fm(z * T(1.0 / 3.0), q, T(1));
    q = fm(z * T(0.5), q, T(1));      // 1 + z/2 + z^2/6 + z^3/24
    const T R = fm(z, q, T(1));
    const T a = h * q;
    const T w = v - u;
    p = fm(w, a, fm(u, h, p));
    v = fm(w, R, u);
}

// Re-integrate bodies after an event (reference src/GameSim.cpp:285-297, 329-340): columns
// c0+1..50 of up to two bodies from their column-c0 states, lanes over columns, re-starting from
// the first column where correctOverlap (:581-623) moves a position.
//   pair: bi, bj both >= 0; wall: bj = -1 and `wall` gives the wall whose overlap is corrected.
template <typename T, int NT>
RGS_DEV void reintegrate(WarpSlice<T, NT> *W, const FastTables<T> *tab, const FastParams<T> &P, int lane, int c0, int bi,
                         int bj, int wall) {
    constexpr int PK = 2 * NT;
    T *grid = W->grid;
    const DevParams<T> &D = P.d;
    const int tyi = (bi == PK), tyj = (bj == PK);
    const bool pair = bj >= 0;
    const int bjj = pair ? bj : bi;
    const T uix = W->cmd[2 * bi], uiy = W->cmd[2 * bi + 1], ujx = W->cmd[2 * bjj], ujy = W->cmd[2 * bjj + 1];
    int cs = c0;
#pragma unroll 1
    while (cs < kGrid) {
        // start state (already overlap-corrected) at column cs
        const T pix = G(grid, bi, PXc, cs), piy = G(grid, bi, PYc, cs), vix = G(grid, bi, VXc, cs), viy = G(grid, bi, VYc, cs);
        const T pjx = G(grid, bjj, PXc, cs), pjy = G(grid, bjj, PYc, cs), vjx = G(grid, bjj, VXc, cs), vjy = G(grid, bjj, VYc, cs);
        const T wix = vix - uix, wiy = viy - uiy, wjx = vjx - ujx, wjy = vjy - ujy;
        int first_fix = kGrid + 1;
#pragma unroll 1
        for (int cb = cs + 1; cb <= kGrid; cb += 32) {
            const int c = cb + lane;
            const bool in = c <= kGrid;
            const int j = in ? c - cs : 1;
            const T tj = D.dt_col * T(j);
            const T gi = tab->g[tyi][j], ki = tab->k[tyi][j];
            T nix = fm(wix, ki, fm(uix, tj, pix)), niy = fm(wiy, ki, fm(uiy, tj, piy));
            const T nvix = fm(wix, gi, uix), nviy = fm(wiy, gi, uiy);
            T njx = T(0), njy = T(0), nvjx = T(0), nvjy = T(0);
            bool fix = false;
            if (pair) {
                const T gj = tab->g[tyj][j], kj = tab->k[tyj][j];
                njx = fm(wjx, kj, fm(ujx, tj, pjx));
                njy = fm(wjy, kj, fm(ujy, tj, pjy));
                nvjx = fm(wjx, gj, ujx);
                nvjy = fm(wjy, gj, ujy);
                const T dx = njx - nix, dy = njy - niy;
                fix = in && (dx * dx + dy * dy <= D.ov_lt[tyi][tyj]);
            } else {
                T ov;
                switch (wall) {
                case WALL_UP: ov = niy + D.rad[tyi] - D.half_y; break;
                case WALL_DOWN: ov = -niy + D.rad[tyi] - D.half_y; break;
                case WALL_RIGHT: ov = nix + D.rad[tyi] - D.half_x; break;
                default: ov = -nix + D.rad[tyi] - D.half_x; break;
                }
                fix = in && (ov > T(0));
            }
            const unsigned m = __ballot_sync(kFull, fix);
            const int stop = m ? cb + (__ffs(m) - 1) : kGrid + 1;  // first column needing a correction
            if (in && c <= stop) {
                if (c == stop) {  // apply correctOverlap on this column
                    if (pair) {
                        const T dx = njx - nix, dy = njy - niy;
                        const T overlap = D.sig[tyi][tyj] - sqrt(dx * dx + dy * dy);
                        const T si = D.fix_coef[tyi][tyj] * overlap, sj = D.fix_coef[tyj][tyi] * overlap;
                        nix -= si * dx; niy -= si * dy;
                        njx += sj * dx; njy += sj * dy;
                    } else {
                        switch (wall) {
                        case WALL_UP: niy -= D.wall_fix * (niy + D.rad[tyi] - D.half_y); break;
                        case WALL_DOWN: niy += D.wall_fix * (-niy + D.rad[tyi] - D.half_y); break;
                        case WALL_RIGHT: nix -= D.wall_fix * (nix + D.rad[tyi] - D.half_x); break;
                        default: nix += D.wall_fix * (-nix + D.rad[tyi] - D.half_x); break;
                        }
                    }
                    W->ev[RGS_EV_OVERLAP_FIXES]++;
                }
                G(grid, bi, PXc, c) = nix; G(grid, bi, PYc, c) = niy; G(grid, bi, VXc, c) = nvix; G(grid, bi, VYc, c) = nviy;
                if (pair) {
                    G(grid, bjj, PXc, c) = njx; G(grid, bjj, PYc, c) = njy; G(grid, bjj, VXc, c) = nvjx; G(grid, bjj, VYc, c) = nvjy;
                }
            }
            __syncwarp();
            if (m) { first_fix = stop; break; }
        }
        if (first_fix > kGrid) break;
        cs = first_fix;  // continue from the corrected column
    }
}

// trajectory record [t, state (reference order), damage] of step k, from the register state
template <typename T, int NT>
RGS_DEV void write_traj_fast(const RunArgs &A, int64_t game, int64_t k, double t_clock, double sa, double sb,
                             const T (&px)[2 * NT + 1], const T (&py)[2 * NT + 1], const T (&vx)[2 * NT + 1],
                             const T (&vy)[2 * NT + 1], int lane) {
    constexpr int E = 2 * NT + 1, PK = 2 * NT, SREC = 2 + 4 * E, NP = 2 * NT, TREC = 1 + SREC + NP * NP;
    double *rec = A.traj + (game * (A.n_steps / A.log_stride) + ((k + 1) / A.log_stride - 1)) * TREC;
    __syncwarp();
    for (int idx = lane; idx < TREC; idx += 32) {
        double val = 0.0;
        if (idx == 0) val = t_clock;
        else if (idx == 1) val = sa;
        else if (idx == 2) val = sb;
        else if (idx < 1 + SREC) {
            const int slot = (idx - 3) >> 2, c = (idx - 3) & 3;
            const int eb = slot == 0 ? PK : slot - 1;
#pragma unroll
            for (int e = 0; e < E; ++e)
                if (e == eb) val = (double)(c == 0 ? px[e] : c == 1 ? py[e] : c == 2 ? vx[e] : vy[e]);
        } else val = A.damage[game * (NP * NP) + (idx - 1 - SREC)];
        rec[idx] = val;
    }
}

template <typename T, int NT>
RGS_DEV void rollout_fast_warp(const FastParams<T> &P, const RunArgs &A, WarpSlice<T, NT> *W, const FastTables<T> *tab,
                               int lane, int64_t game0, int64_t game_stride) {
    constexpr int E = 2 * NT + 1, PK = 2 * NT, SREC = 2 + 4 * E, NP = 2 * NT;
    const DevParams<T> &D = P.d;
    T *grid = W->grid;

    // this lane's two columns and their coefficients, [pass][type]
    const int col0 = 1 + lane, col1 = 33 + lane;
    const bool has1 = col1 <= kGrid;
    const int c1 = has1 ? col1 : kGrid;
    const T tA = D.dt_col * T(col0), tB = D.dt_col * T(c1);
    const T gA[2] = {tab->g[0][col0], tab->g[1][col0]}, kA[2] = {tab->k[0][col0], tab->k[1][col0]};
    const T gB[2] = {tab->g[0][c1], tab->g[1][c1]}, kB[2] = {tab->k[0][c1], tab->k[1][c1]};
    const T gE[2] = {P.tab_g[0][kGrid], P.tab_g[1][kGrid]}, kE[2] = {P.tab_k[0][kGrid], P.tab_k[1][kGrid]};
    const T tE = D.dt;

    for (int64_t game = game0; game < A.n_games; game += game_stride) {
        // ---- load the game record: every lane keeps the whole state in registers ----
        double *srec = A.state + game * SREC;
        for (int idx = lane; idx < SREC; idx += 32) {
            const double val = srec[idx];
            if (idx >= 2) {
                const int slot = (idx - 2) >> 2, c = (idx - 2) & 3;
                G(grid, slot == 0 ? PK : slot - 1, c, 0) = T(val);
            }
        }
        double sa = srec[0], sb = srec[1];
        double t_clock = A.clock[game];
        long long *gev = A.events + game * RGS_N_EVENTS;
        if (lane < RGS_N_EVENTS) W->ev[lane] = gev[lane];
        if (lane < 32) W->cmd[lane] = T(0);
        __syncwarp();
        T px[E], py[E], vx[E], vy[E], ux[E], uy[E];
#pragma unroll
        for (int e = 0; e < E; ++e) {
            px[e] = G(grid, e, PXc, 0); py[e] = G(grid, e, PYc, 0); vx[e] = G(grid, e, VXc, 0); vy[e] = G(grid, e, VYc, 0);
            ux[e] = T(0); uy[e] = T(0);
        }
        long long live = W->ev[RGS_EV_LIVE_STEPS], loop_iters = W->ev[RGS_EV_LOOP_ITERS];
        long long decided = W->ev[RGS_EV_DECIDED_STEP];
        const T *actions = static_cast<const T *>(A.actions);
        int64_t cur_aidx = -1;
        const double wsc = (double)D.winning_score;

        int64_t k = 0;
        for (; k < A.n_steps; ++k) {
            if (!(sa < wsc && sb < wsc)) break;  // undecided(), :57-60

            // ---- commanded velocities (lane c < 4n fetches component c, then broadcast) ----
            {
                int64_t aidx;
                if (A.synth) aidx = (A.step0 + k) / A.hold;
                else { aidx = k / A.hold; if (aidx >= A.n_act) aidx = A.n_act - 1; }
                if (aidx != cur_aidx) {
                    T mine = T(0);
                    if (lane < 2 * NP) {
                        if (A.synth) mine = T(synth_action(A.seed, A.first_game + game, aidx, lane, A.u_max));
                        else mine = actions[game * A.act_game_stride + aidx * (2 * NP) + lane];
                    }
#pragma unroll
                    for (int e = 0; e < NP; ++e) { ux[e] = shfl(mine, 2 * e); uy[e] = shfl(mine, 2 * e + 1); }
                    if (lane < 2 * NP) W->cmd[lane] = mine;
                    cur_aidx = aidx;
                    __syncwarp();
                }
            }

            // ---- register phase: positions of every body at this lane's two columns, overlap flags ----
            T wx[E], wy[E];
#pragma unroll
            for (int e = 0; e < E; ++e) { wx[e] = vx[e] - ux[e]; wy[e] = vy[e] - uy[e]; }
            unsigned flagA = 0, flagB = 0;  // bit e: body e overlaps a wall or another body at this column
            {
                T qx[E], qy[E];
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    const int ty = (e == PK);
                    qx[e] = fm(wx[e], kA[ty], fm(ux[e], tA, px[e]));
                    qy[e] = fm(wy[e], kA[ty], fm(uy[e], tA, py[e]));
                    if (fabs(qx[e]) >= P.lim_x[ty] || fabs(qy[e]) >= P.lim_y[ty]) flagA |= 1u << e;
                }
#pragma unroll
                for (int a = 0; a < E; ++a)
#pragma unroll
                    for (int b = a + 1; b < E; ++b) {
                        const T dx = qx[a] - qx[b], dy = qy[a] - qy[b];
                        if (dx * dx + dy * dy <= D.ov_le[a == PK][b == PK]) flagA |= (1u << a) | (1u << b);
                    }
            }
            {
                T qx[E], qy[E];
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    const int ty = (e == PK);
                    qx[e] = fm(wx[e], kB[ty], fm(ux[e], tB, px[e]));
                    qy[e] = fm(wy[e], kB[ty], fm(uy[e], tB, py[e]));
                    if (fabs(qx[e]) >= P.lim_x[ty] || fabs(qy[e]) >= P.lim_y[ty]) flagB |= 1u << e;
                }
#pragma unroll
                for (int a = 0; a < E; ++a)
#pragma unroll
                    for (int b = a + 1; b < E; ++b) {
                        const T dx = qx[a] - qx[b], dy = qy[a] - qy[b];
                        if (dx * dx + dy * dy <= D.ov_le[a == PK][b == PK]) flagB |= (1u << a) | (1u << b);
                    }
                if (!has1) flagB = 0;
            }
            const bool contact = __any_sync(kFull, (flagA | flagB) != 0);
            int counter = 1;

            if (!contact) {
                // ---- no overlap at any column: commit column 50 in closed form (:153-157) ----
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    const int ty = (e == PK);
                    px[e] = fm(wx[e], kE[ty], fm(ux[e], tE, px[e]));
                    py[e] = fm(wy[e], kE[ty], fm(uy[e], tE, py[e]));
                    vx[e] = fm(wx[e], gE[ty], ux[e]);
                    vy[e] = fm(wy[e], gE[ty], uy[e]);
                }
            } else {
                // ---- materialise the grid (populateStateGrid) and run the event loop (:136-150) ----
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    const int ty = (e == PK);
                    if (lane == 0) {
                        G(grid, e, PXc, 0) = px[e]; G(grid, e, PYc, 0) = py[e]; G(grid, e, VXc, 0) = vx[e]; G(grid, e, VYc, 0) = vy[e];
                    }
                    G(grid, e, PXc, col0) = fm(wx[e], kA[ty], fm(ux[e], tA, px[e]));
                    G(grid, e, PYc, col0) = fm(wy[e], kA[ty], fm(uy[e], tA, py[e]));
                    G(grid, e, VXc, col0) = fm(wx[e], gA[ty], ux[e]);
                    G(grid, e, VYc, col0) = fm(wy[e], gA[ty], uy[e]);
                    if (has1) {
                        G(grid, e, PXc, col1) = fm(wx[e], kB[ty], fm(ux[e], tB, px[e]));
                        G(grid, e, PYc, col1) = fm(wy[e], kB[ty], fm(uy[e], tB, py[e]));
                        G(grid, e, VXc, col1) = fm(wx[e], gB[ty], ux[e]);
                        G(grid, e, VYc, col1) = fm(wy[e], gB[ty], uy[e]);
                    }
                }
                __syncwarp();
                unsigned owners = 0;  // bodies with an overlap flag at some column
                {
                    const unsigned f = flagA | flagB;
#pragma unroll
                    for (int e = 0; e < E; ++e)
                        if (__any_sync(kFull, (f >> e) & 1u)) owners |= 1u << e;
                }
                unsigned qmask = 0;
                int seq = 0, base = 1;
                bool scored = false;
#pragma unroll 1
                for (int id = 0; id < E; ++id) {
                    if (!((owners >> id) & 1u)) continue;
                    T et; int ej, ek;
                    if (scan_owner<T, NT>(grid, D, id, base, T(0), lane, et, ej, ek)) queue_push(W, qmask, seq, id, et, ej, ek, lane);
                }
#pragma unroll 1
                while (qmask != 0 && counter < kCounterLim) {
                    int ei, ej, ekind;
                    T et;
                        queue_pop(W, qmask, ei, ej, ekind, et);
                    if (lane == 0) RGS_TRACE("step %lld pop kind %d i %d j %d t %.17g\n", live, ekind, ei, ej, (double)et);
                    if ((ei == PK || ej == PK) && scored) {  // :230-231
                        if (lane == 0) { W->ev[RGS_EV_DISCARDED]++; hash_event(W->ev, live, DISCARD, ei, ej, 0); }
                        ++counter;
                        continue;
                    }
                    base = (int)floor(et / D.dt_col) + 1;  // :237
                    if (base > kGrid) { base = kGrid; if (lane == 0) W->ev[RGS_EV_BASE_CLAMP]++; }
                    const T dt_t = et - D.dt_col * T(base - 1);  // :254 / :308
                    const T h_rest = D.dt_col - dt_t;
                    int goal = 0;
                    const int tyi = (ei == PK);
                    T pix = G(grid, ei, PXc, base - 1), piy = G(grid, ei, PYc, base - 1);
                    T vix = G(grid, ei, VXc, base - 1), viy = G(grid, ei, VYc, base - 1);
                    const T uix = W->cmd[2 * ei], uiy = W->cmd[2 * ei + 1];
                    advance_axis(pix, vix, uix, dt_t, P.neg_inv_tau[tyi]);
                    advance_axis(piy, viy, uiy, dt_t, P.neg_inv_tau[tyi]);

                    if (ekind == PAIR) {  // :242-299
                        const int tyj = (ej == PK);
                        T pjx = G(grid, ej, PXc, base - 1), pjy = G(grid, ej, PYc, base - 1);
                        T vjx = G(grid, ej, VXc, base - 1), vjy = G(grid, ej, VYc, base - 1);
                        const T ujx = W->cmd[2 * ej], ujy = W->cmd[2 * ej + 1];
                        advance_axis(pjx, vjx, ujx, dt_t, P.neg_inv_tau[tyj]);
                        advance_axis(pjy, vjy, ujy, dt_t, P.neg_inv_tau[tyj]);
                        {  // residual overlap of the constant-velocity impact-time estimate :259-265
                            const T dx = pjx - pix, dy = pjy - piy;
                            const T s = dx * dx + dy * dy;
                            if (s <= D.ov_lt[tyi][tyj]) {
                                const T overlap = D.sig[tyi][tyj] - sqrt(s);
                                const T si = D.hit_coef[tyi][tyj] * overlap, sj = D.hit_coef[tyj][tyi] * overlap;
                                pix -= si * dx; piy -= si * dy;
                                pjx += sj * dx; pjy += sj * dy;
                            }
                        }
                        {  // elastic impulse :268-274
                            const T dvx = vix - vjx, dvy = viy - vjy, dpx = pix - pjx, dpy = piy - pjy;
                            const T dot = fm(dvx, dpx, dvy * dpy), nn = fm(dpx, dpx, dpy * dpy);
                            const T fi = D.imp_coef[tyi][tyj] * dot / nn, fj = D.imp_coef[tyj][tyi] * dot / nn;
                            if (lane == 0) {
                                if (ei != PK && ej != PK) {  // damage :277-282
                                    const double dv_sq = 0.01 * ((double)dvx * (double)dvx + (double)dvy * (double)dvy);
                                    double *dmg = A.damage + game * (NP * NP);
                                    dmg[ei * NP + ej] += (double)D.mass[tyi] * dv_sq;
                                    dmg[ej * NP + ei] += (double)D.mass[tyj] * dv_sq;
                                    W->ev[RGS_EV_PLAYER_PLAYER]++;
                                } else {
                                    W->ev[RGS_EV_PLAYER_PUCK]++;
                                }
                            }
                            vix -= fi * dpx; viy -= fi * dpy;
                            vjx += fj * dpx; vjy += fj * dpy;
                        }
                        advance_axis(pix, vix, uix, h_rest, P.neg_inv_tau[tyi]);
                        advance_axis(piy, viy, uiy, h_rest, P.neg_inv_tau[tyi]);
                        advance_axis(pjx, vjx, ujx, h_rest, P.neg_inv_tau[tyj]);
                        advance_axis(pjy, vjy, ujy, h_rest, P.neg_inv_tau[tyj]);
                        {  // correctOverlap at column base :287
                            const T dx = pjx - pix, dy = pjy - piy;
                            const T s = dx * dx + dy * dy;
                            if (s <= D.ov_lt[tyi][tyj]) {
                                const T overlap = D.sig[tyi][tyj] - sqrt(s);
                                const T si = D.fix_coef[tyi][tyj] * overlap, sj = D.fix_coef[tyj][tyi] * overlap;
                                pix -= si * dx; piy -= si * dy;
                                pjx += sj * dx; pjy += sj * dy;
                                if (lane == 0) W->ev[RGS_EV_OVERLAP_FIXES]++;
                            }
                        }
                        if (lane == 0) {
                            G(grid, ei, PXc, base) = pix; G(grid, ei, PYc, base) = piy; G(grid, ei, VXc, base) = vix; G(grid, ei, VYc, base) = viy;
                            G(grid, ej, PXc, base) = pjx; G(grid, ej, PYc, base) = pjy; G(grid, ej, VXc, base) = vjx; G(grid, ej, VYc, base) = vjy;
                        }
                        __syncwarp();
                        reintegrate<T, NT>(W, tab, P, lane, base, ei, ej, PAIR);
                    } else {  // wall :300-343
                        switch (ekind) {  // flip only while still moving into the wall :312-318
                        case WALL_UP: if (viy > T(0)) viy = -viy; break;
                        case WALL_DOWN: if (viy < T(0)) viy = -viy; break;
                        case WALL_RIGHT: if (vix > T(0)) vix = -vix; break;
                        default: if (vix < T(0)) vix = -vix; break;
                        }
                        const bool is_goal = (ei == PK) && (ekind == WALL_LEFT || ekind == WALL_RIGHT) && piy < D.half_goal &&
                                             piy > -D.half_goal;  // :321-322
                        if (is_goal) {
                            if (!scored) {
                                if (ekind == WALL_RIGHT) { sa += 1.0; goal = 1; } else { sb += 1.0; goal = 2; }
                                if (lane == 0) W->ev[goal == 1 ? RGS_EV_GOALS_A : RGS_EV_GOALS_B]++;
                            }
                            scored = true;
                            pix = piy = vix = viy = T(0);  // puck reset :325
                        } else {
                            advance_axis(pix, vix, uix, h_rest, P.neg_inv_tau[tyi]);
                            advance_axis(piy, viy, uiy, h_rest, P.neg_inv_tau[tyi]);
                            T ov;  // correctOverlap(wall) at column base :334
                            switch (ekind) {
                            case WALL_UP: ov = piy + D.rad[tyi] - D.half_y; if (ov > T(0)) piy -= D.wall_fix * ov; break;
                            case WALL_DOWN: ov = -piy + D.rad[tyi] - D.half_y; if (ov > T(0)) piy += D.wall_fix * ov; break;
                            case WALL_RIGHT: ov = pix + D.rad[tyi] - D.half_x; if (ov > T(0)) pix -= D.wall_fix * ov; break;
                            default: ov = -pix + D.rad[tyi] - D.half_x; if (ov > T(0)) pix += D.wall_fix * ov; break;
                            }
                            if (ov > T(0) && lane == 0) W->ev[RGS_EV_OVERLAP_FIXES]++;
                        }
                        if (lane == 0) {
                            G(grid, ei, PXc, base) = pix; G(grid, ei, PYc, base) = piy; G(grid, ei, VXc, base) = vix; G(grid, ei, VYc, base) = viy;
                            W->ev[RGS_EV_WALL_UP + ekind]++;
                        }
                        __syncwarp();
                        reintegrate<T, NT>(W, tab, P, lane, base, ei, -1, ekind);
                    }
                    if (lane == 0) hash_event(W->ev, live, ekind, ei, ej, goal);
                    __syncwarp();

                    queue_prune(W, qmask, ei, ej);  // :347-352
                    {                               // re-scan the bodies involved :355-357, :147
                        T nt; int nj, nk;
                        if (scan_owner<T, NT>(grid, D, ei, base, T(0), lane, nt, nj, nk)) queue_push(W, qmask, seq, ei, nt, nj, nk, lane);
                        if (ej >= 0 && scan_owner<T, NT>(grid, D, ej, base, T(0), lane, nt, nj, nk))
                            queue_push(W, qmask, seq, ej, nt, nj, nk, lane);
                    }
                    ++counter;
                }
                if (qmask != 0 && lane == 0) W->ev[RGS_EV_COUNTER_LIMIT]++;
                __syncwarp();
#pragma unroll
                for (int e = 0; e < E; ++e) {  // commit column 50 (:153-157)
                    px[e] = G(grid, e, PXc, kGrid); py[e] = G(grid, e, PYc, kGrid);
                    vx[e] = G(grid, e, VXc, kGrid); vy[e] = G(grid, e, VYc, kGrid);
                }
                __syncwarp();
            }
            loop_iters += counter;
            t_clock += (double)D.dt;  // t_ += dt_ (:158), kept in fp64 like the reference clock
            ++live;
            if (decided == 0 && !(sa < wsc && sb < wsc)) decided = live;

            // ---- optional strided trajectory record [t, state, damage] ----
            if (A.traj != nullptr && A.log_stride > 0 && ((k + 1) % A.log_stride) == 0)
                write_traj_fast<T, NT>(A, game, k, t_clock, sa, sb, px, py, vx, vy, lane);
        }
        // frozen steps of a decided game (GameSim::run :121-123) repeat the frozen state
        if (A.traj != nullptr && A.log_stride > 0)
            for (; k < A.n_steps; ++k)
                if (((k + 1) % A.log_stride) == 0) write_traj_fast<T, NT>(A, game, k, t_clock, sa, sb, px, py, vx, vy, lane);

        // ---- store the game record ----
        for (int idx = lane; idx < SREC; idx += 32) {
            double val = 0.0;
            if (idx == 0) val = sa;
            else if (idx == 1) val = sb;
            else {
                const int slot = (idx - 2) >> 2, c = (idx - 2) & 3;
                const int eb = slot == 0 ? PK : slot - 1;
#pragma unroll
                for (int e = 0; e < E; ++e)
                    if (e == eb) val = (double)(c == 0 ? px[e] : c == 1 ? py[e] : c == 2 ? vx[e] : vy[e]);
            }
            srec[idx] = val;
        }
        if (lane == 0) {
            A.clock[game] = t_clock;
            W->ev[RGS_EV_LIVE_STEPS] = live;
            W->ev[RGS_EV_LOOP_ITERS] = loop_iters;
            W->ev[RGS_EV_DECIDED_STEP] = decided;
        }
        __syncwarp();
        if (lane < RGS_N_EVENTS) gev[lane] = W->ev[lane];
        __syncwarp();
    }
}

}  // namespace rgs
