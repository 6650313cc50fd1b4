// rgs_host_params.h -- host-side construction of the kernels' parameter block.
//
// Constant sub-expressions of the reference (src/GameSim.cpp) are folded HERE, in fp64 and in the
// reference's own operator order, so the exact kernel reproduces every rounding:
//   dt_col_ = dt / 50 (:28), dt/2.0 and dt/6.0 of RK4 (:76-87), arena/2 -+ r (:385,408,431,454),
//   r_i + r_j (:260,471,484), OVERLAP_BUFFER * r_i/(r_i+r_j) (:587-588), 1.01 * r_i/(r_i+r_j)
//   (:263-264), 2*m_j/(m_i+m_j) (:273-274).
// The sqrt-free overlap thresholds are exact: ov_le is the largest s with sqrt_rn(s) <= sigma and
// ov_lt the largest s with sqrt_rn(s) < sigma, so (s <= ov_le) <=> (norm <= sigma) bit for bit.
#pragma once
#include <cmath>
#include <limits>

#include "rgs_device.cuh"
#include "rgs_fast.cuh"

namespace rgs {

template <typename T>
inline T sqrt_threshold(T sigma, bool strict) {
    // largest s >= 0 such that sqrt_rn(s) <= sigma (or < sigma when strict)
    const T inf = std::numeric_limits<T>::infinity();
    T s = sigma * sigma;
    auto ok = [&](T x) { const T q = std::sqrt(x); return strict ? (q < sigma) : (q <= sigma); };
    while (!ok(s)) s = std::nextafter(s, T(0));
    while (ok(std::nextafter(s, inf))) s = std::nextafter(s, inf);
    return s;
}

template <typename T>
inline DevParams<T> make_dev_params(const rgs_params &p) {
    DevParams<T> d;
    const double dt_col = p.dt / RGS_GRID_POINTS;
    d.dt = T(p.dt);
    d.dt_col = T(dt_col);
    d.dt_col_h2 = T(dt_col / 2.0);
    d.dt_col_h6 = T(dt_col / 6.0);
    const double tau[2] = {p.tau_player, p.tau_puck}, drag[2] = {1.0, 0.2};
    const double mass[2] = {p.player_mass, p.puck_mass}, rad[2] = {p.player_radius, p.puck_radius};
    d.half_x = T(p.arena_x / 2.0);
    d.half_y = T(p.arena_y / 2.0);
    d.half_goal = T(p.goal_height / 2.0);
    for (int a = 0; a < 2; ++a) {
        d.tau[a] = T(tau[a]);
        // the reciprocal form of x / tau is exact for binary64 operands away from the exponent limits
        d.inv_tau[a] = (tau[a] >= 0x1p-20 && tau[a] <= 0x1p20) ? T(1.0 / tau[a]) : T(0);
        d.drag[a] = T(drag[a]);
        d.mass[a] = T(mass[a]);
        d.rad[a] = T(rad[a]);
        d.lim_hi_y[a] = T(p.arena_y / 2.0 - rad[a]);
        d.lim_lo_y[a] = T(-p.arena_y / 2.0 + rad[a]);
        d.lim_hi_x[a] = T(p.arena_x / 2.0 - rad[a]);
        d.lim_lo_x[a] = T(-p.arena_x / 2.0 + rad[a]);
        for (int b = 0; b < 2; ++b) {
            const double sig = rad[a] + rad[b];
            d.sig[a][b] = T(sig);
            d.ov_le[a][b] = sqrt_threshold<T>(T(sig), false);
            d.ov_lt[a][b] = sqrt_threshold<T>(T(sig), true);
            d.fix_coef[a][b] = T(1.1 * rad[a] / (rad[a] + rad[b]));
            d.hit_coef[a][b] = T(1.01 * rad[a] / (rad[a] + rad[b]));
            d.imp_coef[a][b] = T(2 * mass[b] / (mass[a] + mass[b]));
        }
    }
    d.wall_fix = T(1.1);
    d.dmg_coef = T(0.01);
    d.winning_score = p.winning_score;
    d.n_team = p.players_per_team;
    return d;
}

// Closed-form coefficients of i consecutive RK4 sub-steps (see rgs_fast.cuh), evaluated in long
// double and rounded once to T.
template <typename T>
inline FastParams<T> make_fast_params(const rgs_params &p) {
    FastParams<T> f;
    f.d = make_dev_params<T>(p);
    const long double h = (long double)p.dt / RGS_GRID_POINTS;
    const long double tau[2] = {p.tau_player, p.tau_puck}, drag[2] = {1.0L, 0.2L};
    const double rad[2] = {p.player_radius, p.puck_radius};
    for (int ty = 0; ty < 2; ++ty) {
        const long double nit = -drag[ty] / tau[ty];
        f.neg_inv_tau[ty] = T(nit);
        const long double z = h * nit;
        const long double q = 1.0L + z / 2 + z * z / 6 + z * z * z / 24;
        const long double R = 1.0L + z * q, a = h * q;
        long double Ri = 1.0L, S = 0.0L;  // R^i and a * sum_{m<i} R^m
        for (int i = 0; i < kCols; ++i) {
            f.tab_g[ty][i] = T(Ri);
            f.tab_k[ty][i] = T(S);
            S += a * Ri;
            Ri *= R;
        }
        f.lim_x[ty] = T(p.arena_x / 2.0 - rad[ty]);
        f.lim_y[ty] = T(p.arena_y / 2.0 - rad[ty]);
        f.lim_x_in[ty] = std::nextafter(f.lim_x[ty], T(0));
        f.lim_y_in[ty] = std::nextafter(f.lim_y[ty], T(0));
    }
    {   // process noise (players only): per-sub-step increment and its accumulation over the 50 sub-steps of a step,
        // covariance recurrence S' = M S M^T + Q with M = [[R, 0], [a, 1]] in (v, p) order, evaluated in fp64
        const double sg = p.noise, hd = (double)h, nit = -1.0 / p.tau_player;
        const NoiseCoef<double> q = noise_coef<double>(hd, nit);
        const double z = hd * nit, qq = 1.0 + z / 2 + z * z / 6 + z * z * z / 24, R = 1.0 + z * qq, a = hd * qq;
        const double Qvv = q.l11 * q.l11, Qpv = q.l21 * q.l11, Qpp = q.l21 * q.l21 + q.l22 * q.l22;
        double Svv = 0, Spv = 0, Spp = 0;
        for (int i = 0; i < RGS_GRID_POINTS; ++i) {
            const double nvv = R * R * Svv + Qvv;
            const double npv = R * (a * Svv + Spv) + Qpv;
            const double npp = a * a * Svv + 2 * a * Spv + Spp + Qpp;
            Svv = nvv; Spv = npv; Spp = npp;
        }
        const double e11 = std::sqrt(Svv), e21 = e11 > 0 ? Spv / e11 : 0.0, e22 = std::sqrt(std::fmax(0.0, Spp - e21 * e21));
        f.noise = T(sg);
        f.nz_sub[0] = T(sg * q.l11); f.nz_sub[1] = T(sg * q.l21); f.nz_sub[2] = T(sg * q.l22);
        f.nz_end[0] = T(sg * e11); f.nz_end[1] = T(sg * e21); f.nz_end[2] = T(sg * e22);
        f.nz_env = T(6.5 * std::fabs(sg) * std::sqrt(Spp));
        f.dt_col64 = p.dt / RGS_GRID_POINTS;
        f.inv_dt_col64 = 1.0 / f.dt_col64;
        f.seed_lo = (uint32_t)p.seed;
        f.seed_hi = 0x5EEDu;
    }
    return f;
}

}  // namespace rgs
