// rgs_fast.cuh -- arithmetic shared by the K2 (throughput formulation) kernel body in rgs_lane.cuh.
//
// The reference integrates every body with 50 RK4 sub-steps per game-step (populateStateGrid, reference
// src/GameSim.cpp:173-217; RK4_player / RK4_puck :70-116).  The dynamics are linear with a constant command inside a
// step, so i RK4 sub-steps are the affine map  v_i = u + (v0-u) R^i,  p_i = p0 + u i h + (v0-u) a (1-R^i)/(1-R)
// with R = 1+z+z^2/2+z^3/6+z^4/24, a = h (1+z/2+z^2/6+z^3/24), z = -h drag/tau  (exactly what RK4 computes, up to
// rounding).  This header holds the coefficient block (FastParams, built on the host in rgs_host_params.h), the
// explicit fused multiply-add fm() and the arbitrary-h advance used at a time of impact.
//
// All fused multiply-adds are spelled out (fm()), everything else is a single IEEE operation and the file is compiled
// with -fmad=false, so the arithmetic is the same wherever this source is built (nvcc for sm_100a; g++
// -ffp-contract=off for the CPU-side emulator check under tests/emu).
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
    // process noise on the player accelerations (reference src/GameSim.cpp:62-68), see NoiseCoef below; all scaled by sigma
    T noise;                     // sigma = w_stdev_ (0: none)
    T nz_sub[3];                 // Cholesky factor {l11, l21, l22} of the (dv, dp) increment of ONE sub-step (h = dt_col)
    T nz_end[3];                 // same for the noise accumulated over a whole game-step (50 sub-steps)
    T nz_env;                    // 6.5 standard deviations of the accumulated position noise (broad-phase envelope)
    uint32_t seed_lo, seed_hi;   // reset()'s seed
    double dt_col64, inv_dt_col64;  // dt / 50 in fp64 (the reference's dt_col_, :28) and RN(1 / dt_col_): base_idx decision of resolve()
};

// The reference adds sigma * N(0,1) to both player accelerations on EVERY f_player call, i.e. on each of the four
// stages of each RK4 sub-step (:66-67, :73-85).  The dynamics are linear, so the four stage draws e1..e4 of one axis
// enter the sub-step result linearly (z = -h / tau):
//   dv = (h/6)   [ e1 (1 + z + z^2/2 + z^3/4) + e2 (2 + z + z^2/2) + e3 (2 + z) + e4 ]
//   dp = (h^2/6) [ e1 (1 + z/2 + z^2/4)       + e2 (1 + z/2)       + e3 ]
// -- a 2-D Gaussian increment per sub-step and axis, sampled EXACTLY with two normals (Cholesky factor l11, l21, l22
// of its covariance) instead of the four stage draws.  Per unit sigma.
template <typename T>
struct NoiseCoef { T l11, l21, l22; };
template <typename T>
__host__ __device__ inline NoiseCoef<T> noise_coef(T h, T neg_inv_tau) {
    const T z = h * neg_inv_tau;
    const T c1 = T(1) + z + z * z * T(0.5) + z * z * z * T(0.25), c2 = T(2) + z + z * z * T(0.5), c3 = T(2) + z, c4 = T(1);
    const T d1 = T(1) + z * T(0.5) + z * z * T(0.25), d2 = T(1) + z * T(0.5), d3 = T(1);
    const T sv = h / T(6), sp = h * h / T(6);
    const T qvv = sv * sv * (c1 * c1 + c2 * c2 + c3 * c3 + c4 * c4);
    const T qpp = sp * sp * (d1 * d1 + d2 * d2 + d3 * d3);
    const T qpv = sv * sp * (c1 * d1 + c2 * d2 + c3 * d3);
    NoiseCoef<T> r;
    r.l11 = T(sqrt(qvv));
    r.l21 = r.l11 > T(0) ? qpv / r.l11 : T(0);
    const T rem = qpp - r.l21 * r.l21;
    r.l22 = rem > T(0) ? T(sqrt(rem)) : T(0);
    return r;
}

// four independent standard normals from one Philox4x32-10 block (Box-Muller), keyed by the reset seed and counted
// by (global game index, game-step, draw id)
RGS_DEV void box_muller(uint32_t a, uint32_t b, float &g0, float &g1) {
    const float u1 = ((float)(a >> 8) + 1.0f) * 5.9604645e-8f;   // (0, 1]
    const float u2 = (float)(b >> 8) * 5.9604645e-8f;            // [0, 1)
    const float th = 6.2831853071795865f * u2;
#ifdef RGS_HOST_EMU
    const float r = sqrtf(-2.0f * logf(u1));
    g0 = r * cosf(th);
    g1 = r * sinf(th);
#else
    // device: the special-function-unit forms (absolute error ~2^-21 on [0, 2 pi) and on (0, 1]: far below what any
    // statistic of the noise can see); the noise paths of a flagged step are 50 of these per player
    const float r = sqrtf(-2.0f * __logf(u1));
    float sn, cs;
    __sincosf(th, &sn, &cs);
    g0 = r * cs;
    g1 = r * sn;
#endif
}
RGS_DEV void box_muller(uint32_t a, uint32_t b, double &g0, double &g1) {
    // the fp64 instantiation draws the same single-precision normals (24-bit uniforms, tails to 5.8 sigma) and widens
    // them: what it validates is the formulation, and double-precision log / sin / cos made its noise paths 2x dearer
    float f0, f1;
    box_muller(a, b, f0, f1);
    g0 = (double)f0;
    g1 = (double)f1;
}
template <typename T>
RGS_DEV void normals4(uint32_t seed_lo, uint32_t seed_hi, long long game, uint32_t step, uint32_t id, T g[4]) {
    uint32_t o[4];
    philox4x32_10((uint32_t)game, (uint32_t)((unsigned long long)game >> 32), step, id, seed_lo, seed_hi ^ 0x4E5A4E5Au, o);
    box_muller(o[0], o[1], g[0], g[1]);
    box_muller(o[2], o[3], g[2], g[3]);
}

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

}  // namespace rgs
