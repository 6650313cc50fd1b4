// rgs_device.cuh -- device-side building blocks shared by the rollout kernels (sm_100a).
//
// One warp owns one game.  The per-step collision grid of the reference (51 columns of
// position/velocity per body, reference include/GameSim.h:46) lives in the warp's slice of shared
// memory as rows [body][px,py,vx,vy][column]; lanes map to COLUMNS for every scan (consecutive
// lanes -> consecutive words, conflict free) and the first overlapping column of an owner is
// found with one __ballot_sync.  The event queue (<= one event per owner body, reference
// src/GameSim.cpp:136-150) is a tiny per-warp table read by broadcast.
//
// The kernel BODIES in rgs_*.cuh are plain C++ over a small set of warp intrinsics
// (__shfl_sync, __ballot_sync, __syncwarp) so that tests/emu can compile the very same source with
// g++ and step a warp on the CPU (fiber-per-lane emulator; test infrastructure only, the product
// has no CPU path).  RGS_HOST_EMU is defined only by tests/emu/warp_emu.h.
#pragma once
#include <cstdint>
#ifndef RGS_HOST_EMU
#include <cuda_runtime.h>
#define RGS_DEV __device__ __forceinline__
#endif

#ifndef RGS_TRACE
#define RGS_TRACE(...) ((void)0)   // tests/emu may define it to print the event sequence
#endif

#include "../../include/rgs_capi.h"

namespace rgs {

constexpr int kGrid = RGS_GRID_POINTS;     // 50
constexpr int kCols = kGrid + 1;           // columns 0..50
constexpr int kGS = 51;                    // row stride (elements)
constexpr int kCounterLim = RGS_COUNTER_LIM;
constexpr unsigned kFull = 0xffffffffu;

enum { WALL_UP = 0, WALL_DOWN = 1, WALL_RIGHT = 2, WALL_LEFT = 3, PAIR = 4, DISCARD = 5 };  // Collision.h:5
enum { PXc = 0, PYc = 1, VXc = 2, VYc = 3 };

// Everything a kernel needs, pre-combined on the host IN THE REFERENCE'S OPERATOR ORDER so the
// device never re-associates a constant expression.  Index [0] = player, [1] = puck.
template <typename T>
struct DevParams {
    T dt, dt_col, dt_col_h2, dt_col_h6;
    T tau[2], drag[2], mass[2], rad[2];
    T inv_tau[2];    // RN(1 / tau): the reciprocal K1's division by tau is built from (rgs_exact.cuh div_tau); 0 = use '/' 
    T half_x, half_y, half_goal;
    T lim_hi_y[2], lim_lo_y[2], lim_hi_x[2], lim_lo_x[2];  // arena/2 - r ; -arena/2 + r
    T sig[2][2];     // r_a + r_b
    T ov_le[2][2];   // sqrt_rn(s) <= sig  <=>  s <= ov_le   (exact threshold, found on the host)
    T ov_lt[2][2];   // sqrt_rn(s) <  sig  <=>  s <= ov_lt
    T fix_coef[2][2];  // (1.1  * r_a) / (r_a + r_b)   OVERLAP_BUFFER, GameSim.cpp:587-588
    T hit_coef[2][2];  // (1.01 * r_a) / (r_a + r_b)   GameSim.cpp:263-264
    T imp_coef[2][2];  // (2 * m_b) / (m_a + m_b)      GameSim.cpp:273-274
    T wall_fix;        // 1.1
    T dmg_coef;        // 0.01, GameSim.cpp:221
    int winning_score;
    int n_team;
};

// on-device surrogate policy (rgs_policy.cuh); mirrors rgs_policy of include/rgs_capi.h
struct PolicyArgs {
    double kick_velocity, gain, input_limit, safety_radius, barrier_gain;
    int cbf, sweeps, kind;
};

struct RunArgs {
    int64_t n_games;
    double *state;      // [N, 2+4E]
    double *clock;      // [N]
    double *damage;     // [N, (2n)^2]
    long long *events;  // [N, 16]
    const void *actions;
    int64_t n_act, hold, act_game_stride;
    int64_t n_steps, log_stride;
    double *traj;
    int64_t traj_recs;  // trajectory records per game = n_steps / log_stride (0 when not logging)
    // trajectory layout: field f of record r of game g is traj[(g * traj_recs + r) * TREC + f] (records, the default:
    // traj_field_stride = 1) or traj[(r * TREC + f) * traj_field_stride + g] (field-major, traj_field_stride = the
    // batch's game count: the 32 lanes of a warp -- 32 consecutive games -- store 256 contiguous bytes per field)
    int64_t traj_field_stride;
    // command source: 0 the action tensor, 1 on-chip synthetic actions, 2 the on-device policy `pol` (K2 lane kernel)
    int synth;
    uint64_t seed;
    int64_t first_game, step0;
    double u_max;
    PolicyArgs pol;
    // K2 lane kernel: per-resident-lane global scratch (correctOverlap tracks + jam snapshot, rgs_lane.cuh)
    void *lane_scratch;
    size_t lane_scratch_bytes;
    int64_t noise_game0;   // global index of game 0 of this launch for the process-noise streams
};

// ---- Philox4x32-10 (Salmon et al. 2011), counter-based: keyed by (seed, game, action index) ----
__host__ __device__ inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll 1   // rolled: the generator sits in the rollout kernels' step loop but runs once per `hold` steps
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// U(-a, a) from 32 random bits; every operation is exact or a single fp64 rounding, so the host
// (numpy) and the device produce the same double, and the same float after one more rounding.
__host__ __device__ inline double sym_uniform(uint32_t bits, double a) {
    const double u = (double)bits * 2.3283064365386963e-10;  // 2^-32, exact
    return (2.0 * u - 1.0) * a;
}

// Component `comp` (0 .. 4n-1) of the synthetic action record `aidx` of global game `game`.
RGS_DEV double synth_action(uint64_t seed, int64_t game, int64_t aidx, int comp, double u_max) {
    uint32_t o[4];
    philox4x32_10((uint32_t)game, (uint32_t)((uint64_t)game >> 32), (uint32_t)aidx,
                  ((uint32_t)((uint64_t)aidx >> 32) << 8) | (uint32_t)(comp >> 2), (uint32_t)seed,
                  (uint32_t)(seed >> 32), o);
    return sym_uniform(o[comp & 3], u_max);
}

// ---- per-warp shared-memory slice -----------------------------------------------------------
template <typename T, int NT>
struct WarpSlice {
    static constexpr int E = 2 * NT + 1;
    T grid[4 * E * kGS];
    T cmd[32];       // commanded velocity of body e, axis a at [2e+a] (fast kernel: run-time indexed reads)
    T qt[16];        // queued event time per owner body
    int qinfo[16];   // (seq << 8) | (kind << 4) | (j & 15), j = 15 for walls
    long long ev[RGS_N_EVENTS];
};

template <typename T>
RGS_DEV T &G(T *grid, int e, int c, int col) { return grid[(e * 4 + c) * kGS + col]; }

template <typename T>
RGS_DEV T shfl(T v, int src) { return __shfl_sync(kFull, v, src); }

RGS_DEV void hash_event(long long *ev, long long live, int kind, int i, int j, int goal) {
    unsigned long long h = (unsigned long long)ev[RGS_EV_HASH];
    const unsigned long long w = ((unsigned long long)live << 16) | ((unsigned long long)kind << 12) |
                                 ((unsigned long long)i << 8) | ((unsigned long long)(j < 0 ? 15 : j) << 4) |
                                 (unsigned long long)goal;
    h ^= w;
    h *= 0x100000001B3ULL;
    ev[RGS_EV_HASH] = (long long)h;
}

// Inner body of insertCollisions for ONE owner (reference src/GameSim.cpp:362-507), lanes over
// columns: every lane builds the candidate event of its column (walls UP, DOWN, RIGHT, LEFT, then
// the other bodies in entity order, strict '<' keeps the first of a tie), one ballot picks the
// first column whose candidate falls inside this step, its lane broadcasts the event.
// t0 is the simulator clock t_ (0 in step-relative mode).  Uniform return value.
template <typename T, int NT>
RGS_DEV bool scan_owner(T *grid, const DevParams<T> &P, int id, int base, T t0, int lane, T &out_t,
                                           int &out_j, int &out_kind) {
    constexpr int E = 2 * NT + 1, PK = 2 * NT;
    const int ty = (id == PK);
    const T r = P.rad[ty];
    for (int c0 = base; c0 <= kGrid; c0 += 32) {
        const int col = c0 + lane;
        T bt = T(1.0e9);
        int bj = -1, bk = -1;
        bool valid = false;
        if (col <= kGrid) {
            const T px = G(grid, id, PXc, col), py = G(grid, id, PYc, col);
            const T tm = t0 + P.dt_col * T(col - 1);  // T[i-1] = t_ + dt_col_*(i-1), GameSim.cpp:201
            if (py + r >= P.half_y) {                 // WALL_UP :374-394
                const T vm = G(grid, id, VYc, col - 1);
                if (vm > T(0)) {
                    const T tc = tm + fmax(T(0), (P.lim_hi_y[ty] - G(grid, id, PYc, col - 1)) / vm);
                    if (tc < bt) { bt = tc; bj = -1; bk = WALL_UP; }
                }
            }
            if (py - r <= -P.half_y) {  // WALL_DOWN :397-417
                const T vm = G(grid, id, VYc, col - 1);
                if (vm < T(0)) {
                    const T tc = tm + fmax(T(0), (P.lim_lo_y[ty] - G(grid, id, PYc, col - 1)) / vm);
                    if (tc < bt) { bt = tc; bj = -1; bk = WALL_DOWN; }
                }
            }
            if (px + r >= P.half_x) {  // WALL_RIGHT :420-440
                const T vm = G(grid, id, VXc, col - 1);
                if (vm > T(0)) {
                    const T tc = tm + fmax(T(0), (P.lim_hi_x[ty] - G(grid, id, PXc, col - 1)) / vm);
                    if (tc < bt) { bt = tc; bj = -1; bk = WALL_RIGHT; }
                }
            }
            if (px - r <= -P.half_x) {  // WALL_LEFT :443-463
                const T vm = G(grid, id, VXc, col - 1);
                if (vm < T(0)) {
                    const T tc = tm + fmax(T(0), (P.lim_lo_x[ty] - G(grid, id, PXc, col - 1)) / vm);
                    if (tc < bt) { bt = tc; bj = -1; bk = WALL_LEFT; }
                }
            }
#pragma unroll
            for (int o = 0; o < E; ++o) {  // INTER_AGENT :466-498
                if (o == id) continue;
                const int tyo = (o == PK);
                const T ex = px - G(grid, o, PXc, col), ey = py - G(grid, o, PYc, col);
                if (ex * ex + ey * ey <= P.ov_le[ty][tyo]) {  // <=> norm <= r + r_o, exactly
                    const T dpx = G(grid, o, PXc, col - 1) - G(grid, id, PXc, col - 1);
                    const T dpy = G(grid, o, PYc, col - 1) - G(grid, id, PYc, col - 1);
                    const T dvx = G(grid, o, VXc, col - 1) - G(grid, id, VXc, col - 1);
                    const T dvy = G(grid, o, VYc, col - 1) - G(grid, id, VYc, col - 1);
                    const T sigma = P.sig[ty][tyo];
                    const T dpdv = dpx * dvx + dpy * dvy;
                    const T dvdv = dvx * dvx + dvy * dvy;
                    const T dpdp = dpx * dpx + dpy * dpy;
                    const T d = dpdv * dpdv - dvdv * (dpdp - sigma * sigma);
                    if (dpdv < T(0) && d >= T(0)) {
                        const T tc = tm + fmax(T(0), -(dpdv + sqrt(d)) / dvdv);
                        if (tc < bt) { bt = tc; bj = o; bk = PAIR; }
                    }
                }
            }
            valid = (bt - t0 < P.dt);  // :501
        }
        const unsigned m = __ballot_sync(kFull, valid);
        if (m) {
            const int src = __ffs(m) - 1;
            out_t = shfl(bt, src);
            out_j = __shfl_sync(kFull, bj, src);
            out_kind = __shfl_sync(kFull, bk, src);
            return true;
        }
    }
    return false;
}

// Event table helpers.  Vector order of the reference queue is (t descending, insertion order
// ascending) because std::sort on <= 16 elements is a stable insertion sort (GameSim.cpp:510,
// Collision.h:24-30); the element executed next is the LAST one: smallest t, and among equal
// times the most recently inserted.
template <typename T, int NT>
RGS_DEV void queue_push(WarpSlice<T, NT> *W, unsigned &qmask, int &seq, int owner, T t, int j, int kind,
                                           int lane) {
    if (lane == 0) {
        W->qt[owner] = t;
        W->qinfo[owner] = (seq << 8) | (kind << 4) | (j & 15);
    }
    qmask |= 1u << owner;
    ++seq;
    __syncwarp();
}

template <typename T, int NT>
RGS_DEV void queue_pop(WarpSlice<T, NT> *W, unsigned &qmask, int &ei, int &ej, int &ekind, T &et) {
    constexpr int E = 2 * NT + 1;
    int best = -1, binfo = 0;
    T bt = T(0);
#pragma unroll
    for (int e = 0; e < E; ++e) {
        if ((qmask >> e) & 1u) {
            const T te = W->qt[e];
            const int info = W->qinfo[e];
            if (best < 0 || te < bt || (te == bt && (info >> 8) > (binfo >> 8))) { best = e; bt = te; binfo = info; }
        }
    }
    ei = best;
    et = bt;
    ekind = (binfo >> 4) & 15;
    ej = binfo & 15;
    if (ej == 15) ej = -1;
    qmask &= ~(1u << best);
}

// Drop queued events sharing a body with the executed one (GameSim.cpp:347-352).
template <typename T, int NT>
RGS_DEV void queue_prune(WarpSlice<T, NT> *W, unsigned &qmask, int ei, int ej) {
    constexpr int E = 2 * NT + 1;
#pragma unroll
    for (int e = 0; e < E; ++e) {
        if ((qmask >> e) & 1u) {
            int cj = W->qinfo[e] & 15;
            if (cj == 15) cj = -1;
            const bool drop = (ei == e) || (ei == cj && cj != -1) || (ej == e) || (ej == cj && cj != -1);
            if (drop) qmask &= ~(1u << e);
        }
    }
}

}  // namespace rgs
