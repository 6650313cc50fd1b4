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

}  // namespace rgs
