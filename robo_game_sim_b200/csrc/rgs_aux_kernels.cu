// rgs_aux_kernels.cu -- K4 (statistics reduction) and K5 (batched reset, synthetic "S0" inputs).
//
// K5/reset replaces GameSim::reset (reference src/GameSim.cpp:21-55) for a whole batch.
// K4 has no reference counterpart: the reference leaves win/score statistics to an offline MATLAB
// script over per-run log files (matlab/interpret_mc_results.m:29-50); here they are reduced on
// the device into int64[RGS_N_STATS] so a multi-GPU Monte Carlo run needs ONE tiny all-reduce.
#include "rgs_kernels.h"

namespace rgs {

struct ResetArgs {
    double start[2 * RGS_MAX_TEAM][2];  // player start positions, computed on the host (:35-38)
    int n_team;
};

__global__ void __launch_bounds__(256) reset_kernel(const __grid_constant__ ResetArgs R, int64_t n_games,
                                                    const double *__restrict__ x0_ball, int64_t x0_stride,
                                                    const double *__restrict__ state0, int64_t state0_stride, double *__restrict__ state,
                                                    double *__restrict__ clock, double *__restrict__ damage,
                                                    long long *__restrict__ events) {
    const int E = 2 * R.n_team + 1, SREC = 2 + 4 * E, NP = 2 * R.n_team, DREC = NP * NP;
    // one thread per output word: fully coalesced stores over the game-major records
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int64_t w = tid; w < n_games * SREC; w += stride) {
        const int64_t g = w / SREC;
        const int idx = (int)(w - g * SREC);
        double v;
        if (state0) v = state0[g * state0_stride + idx];
        else if (idx < 2) v = 0.0;
        else if (idx < 6) v = x0_ball ? x0_ball[g * x0_stride + (idx - 2)] : 0.0;
        else {
            const int pl = (idx - 6) >> 2, c = (idx - 6) & 3;
            v = c < 2 ? R.start[pl][c] : 0.0;
        }
        state[w] = v;
    }
    for (int64_t w = tid; w < n_games * DREC; w += stride) damage[w] = 0.0;
    for (int64_t w = tid; w < n_games * RGS_N_EVENTS; w += stride) events[w] = 0;
    for (int64_t w = tid; w < n_games; w += stride) clock[w] = 0.0;
}

cudaError_t launch_reset(int n_team, int64_t n_games, const rgs_params &p, const double *d_x0_ball, int64_t x0_stride,
                         const double *d_state0, int64_t state0_stride, double *state, double *clock, double *damage,
                         long long *events, cudaStream_t stream) {
    ResetArgs R;
    R.n_team = n_team;
    for (int team = 0; team < 2; ++team)
        for (int k = 0; k < n_team; ++k) {
            double *s = R.start[team * n_team + k];
            s[0] = team == 0 ? -p.arena_x / 4.0 : p.arena_x / 4.0;
            // n = 2: +-arena_Y/4 exactly as the reference writes it; general n: evenly spread
            s[1] = (n_team == 2) ? (k == 0 ? p.arena_y / 4.0 : -p.arena_y / 4.0)
                                 : p.arena_y / 2.0 - (k + 0.5) * p.arena_y / n_team;
        }
    const int threads = 256;
    int64_t blocks = (n_games * (2 + 4 * (2 * n_team + 1)) + threads - 1) / threads;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    reset_kernel<<<(unsigned)blocks, threads, 0, stream>>>(R, n_games, d_x0_ball, x0_stride, d_state0, state0_stride, state, clock,
                                                           damage, events);
    return cudaGetLastError();
}


// ---- K3 epilogue: terminal cost of every rolled-out candidate + arg-min ------------------------
// cost[g] = sum over players p, components c of w[p][c] * (x_g[p][c] - x_des[p][c])^2: the terminal
// term of the reference's DMPC objective, (x_N - x_des)^T Omega_N (x_N - x_des) with
// Omega_N_max = diag(10, 10, 20, 20) (reference py/src/DMPCTeam.py:56-59, py/src/DMPC.py:16-54),
// evaluated on the SIMULATED final state instead of the controller's internal linear model.
// best[0] = arg-min (lowest index among equal costs), best[1] = its cost (double bits).
__global__ void __launch_bounds__(256) cost_kernel(int64_t n_games, int srec, int np, const double *__restrict__ state,
                                                   const double *__restrict__ x_des, const double *__restrict__ w,
                                                   double *__restrict__ cost, unsigned long long *__restrict__ block_best) {
    __shared__ double sc[256];
    __shared__ long long si[256];
    double bc = 1.0e300;
    long long bi = -1;
    for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < n_games; g += (int64_t)gridDim.x * blockDim.x) {
        const double *x = state + g * srec + 6;  // players follow [TAS, TBS, PK(4)]
        double c = 0.0;
        for (int k = 0; k < 4 * np; ++k) {
            const double d = x[k] - x_des[k];
            c += w[k] * d * d;
        }
        cost[g] = c;
        if (c < bc) { bc = c; bi = g; }  // g increases: the first minimum of this thread is kept
    }
    sc[threadIdx.x] = bc;
    si[threadIdx.x] = bi;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
            const double c2 = sc[threadIdx.x + o];
            const long long i2 = si[threadIdx.x + o];
            if (i2 >= 0 && (si[threadIdx.x] < 0 || c2 < sc[threadIdx.x] || (c2 == sc[threadIdx.x] && i2 < si[threadIdx.x]))) {
                sc[threadIdx.x] = c2;
                si[threadIdx.x] = i2;
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        block_best[2 * blockIdx.x] = (unsigned long long)si[0];
        block_best[2 * blockIdx.x + 1] = (unsigned long long)__double_as_longlong(sc[0]);
    }
}
__global__ void argmin_final_kernel(int n_blocks, const unsigned long long *__restrict__ block_best, long long *__restrict__ best) {
    double bc = 1.0e300;
    long long bi = -1;
    for (int b = 0; b < n_blocks; ++b) {
        const long long i = (long long)block_best[2 * b];
        const double c = __longlong_as_double((long long)block_best[2 * b + 1]);
        if (i >= 0 && (bi < 0 || c < bc || (c == bc && i < bi))) { bc = c; bi = i; }
    }
    best[0] = bi;
    best[1] = __double_as_longlong(bc);
}

cudaError_t launch_terminal_cost(int n_team, int64_t n_games, const double *state, const double *d_x_des, const double *d_w,
                                 double *d_cost, unsigned long long *scratch, long long *d_best, cudaStream_t stream) {
    const int threads = 256;
    int64_t blocks = (n_games + threads - 1) / threads;
    if (blocks > kCostBlocks) blocks = kCostBlocks;
    if (blocks < 1) blocks = 1;
    cost_kernel<<<(unsigned)blocks, threads, 0, stream>>>(n_games, 2 + 4 * (2 * n_team + 1), 2 * n_team, state, d_x_des, d_w, d_cost,
                                                          scratch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    argmin_final_kernel<<<1, 1, 0, stream>>>((int)blocks, scratch, d_best);
    return cudaGetLastError();
}

// ---- synthetic inputs "S0" -----------------------------------------------------------------
// x0_ball of game g: Philox counter (g_lo, g_hi, 0xFFFFFFFF, 0xFFFFFFFF) -> px~U(-4,4), py~U(-2,2),
// vx,vy~U(-3,3).  Actions: rgs_device.cuh synth_action (counter (g_lo, g_hi, aidx_lo, aidx_hi<<8|word)).
__global__ void __launch_bounds__(256) synth_kernel(uint64_t seed, int64_t first_game, int64_t n, int n_team,
                                                    int64_t n_act, double u_max, double *__restrict__ x0_ball,
                                                    void *__restrict__ actions, int dtype) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x0_ball) {
        for (int64_t g = tid; g < n; g += stride) {
            const int64_t game = first_game + g;
            uint32_t o[4];
            philox4x32_10((uint32_t)game, (uint32_t)((uint64_t)game >> 32), 0xFFFFFFFFu, 0xFFFFFFFFu, (uint32_t)seed,
                          (uint32_t)(seed >> 32), o);
            double4 v;
            v.x = sym_uniform(o[0], 4.0);
            v.y = sym_uniform(o[1], 2.0);
            v.z = sym_uniform(o[2], 3.0);
            v.w = sym_uniform(o[3], 3.0);
            reinterpret_cast<double4 *>(x0_ball)[g] = v;
        }
    }
    if (actions) {
        // one thread per Philox block = 4 consecutive components of one action record (4n is a multiple of 4)
        const int quads = n_team;  // 4n components / 4
        const int64_t total = n * n_act * quads;
        for (int64_t w = tid; w < total; w += stride) {
            const int q = (int)(w % quads);
            const int64_t rec = w / quads;
            const int64_t aidx = rec % n_act, g = rec / n_act;
            const int64_t game = first_game + g;
            uint32_t o[4];
            philox4x32_10((uint32_t)game, (uint32_t)((uint64_t)game >> 32), (uint32_t)aidx,
                          ((uint32_t)((uint64_t)aidx >> 32) << 8) | (uint32_t)q, (uint32_t)seed, (uint32_t)(seed >> 32),
                          o);
            if (dtype == 64) {
                double *dst = static_cast<double *>(actions) + w * 4;
                reinterpret_cast<double2 *>(dst)[0] = make_double2(sym_uniform(o[0], u_max), sym_uniform(o[1], u_max));
                reinterpret_cast<double2 *>(dst)[1] = make_double2(sym_uniform(o[2], u_max), sym_uniform(o[3], u_max));
            } else {
                float4 f;
                f.x = (float)sym_uniform(o[0], u_max);
                f.y = (float)sym_uniform(o[1], u_max);
                f.z = (float)sym_uniform(o[2], u_max);
                f.w = (float)sym_uniform(o[3], u_max);
                reinterpret_cast<float4 *>(actions)[w] = f;
            }
        }
    }
}

cudaError_t launch_synth(uint64_t seed, int64_t first_game, int64_t n, int n_team, int64_t n_act, double u_max,
                         double *d_x0_ball, void *d_actions, int dtype, cudaStream_t stream) {
    const int threads = 256;
    int64_t work = d_actions ? n * n_act * n_team : n;
    int64_t blocks = (work + threads - 1) / threads;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    synth_kernel<<<(unsigned)blocks, threads, 0, stream>>>(seed, first_game, n, n_team, n_act, u_max, d_x0_ball,
                                                           d_actions, dtype);
    return cudaGetLastError();
}

// ---- K4: statistics ------------------------------------------------------------------------
// Damage (reference src/GameSim.cpp:277-282; histogrammed per run by matlab/interpret_mc_results.m:37-48) is summed
// in FIXED POINT (RGS_DAMAGE_FIXED_SCALE units, rounded per game) so that the whole vector stays one integer
// all-reduce and the sum does not depend on the order of the games.
__global__ void __launch_bounds__(256) stats_kernel(int64_t n_games, int srec, int np, const double *__restrict__ state,
                                                    const long long *__restrict__ events, const double *__restrict__ damage,
                                                    int winning_score, unsigned long long *__restrict__ stats) {
    __shared__ unsigned long long sh[RGS_N_STATS];
    if (threadIdx.x < RGS_N_STATS) sh[threadIdx.x] = 0ull;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    // per-thread partial sums of the 14 additive event fields (lane-private, reduced by shuffles)
    long long acc[14];
#pragma unroll
    for (int k = 0; k < 14; ++k) acc[k] = 0;
    for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < n_games; g += stride) {
        const double sa = state[g * srec], sb = state[g * srec + 1];
        const bool won_a = sa >= (double)winning_score, won_b = sb >= (double)winning_score;
        atomicAdd(&sh[RGS_ST_GAMES], 1ull);
        // the reference stops a game as soon as either score reaches winning_score (:57-60)
        if (won_a) atomicAdd(&sh[RGS_ST_WINS_A], 1ull);
        else if (won_b) atomicAdd(&sh[RGS_ST_WINS_B], 1ull);
        else atomicAdd(&sh[RGS_ST_UNDECIDED], 1ull);
        atomicAdd(&sh[sa > sb ? RGS_ST_LEAD_A : (sb > sa ? RGS_ST_LEAD_B : RGS_ST_TIED)], 1ull);
        const int ia = sa < 4.0 ? (int)sa : 4, ib = sb < 4.0 ? (int)sb : 4;
        atomicAdd(&sh[RGS_ST_SCORE_HIST + 5 * ia + ib], 1ull);
        const longlong2 *ev = reinterpret_cast<const longlong2 *>(events + g * RGS_N_EVENTS);
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            const longlong2 e = ev[k];
            acc[2 * k] += e.x;
            acc[2 * k + 1] += e.y;
        }
        if (ev[7].y != 0) atomicAdd(&sh[RGS_ST_OVERLAP_FIXES], (unsigned long long)ev[7].y);
        if (ev[3].x != 0) {  // only games with player-player impacts carry damage
            const double *dm = damage + g * (np * np);
            const int nt = np >> 1;
            for (int a = 0; a < np; ++a)
                for (int v = 0; v < np; ++v) {
                    const double d = dm[a * np + v];
                    if (d != 0.0) {
                        const unsigned long long q = (unsigned long long)llrint(d * (double)RGS_DAMAGE_FIXED_SCALE);
                        atomicAdd(&sh[RGS_ST_DAMAGE_TEAM + 2 * (a >= nt) + (v >= nt)], q);
                        if (np == 4) atomicAdd(&sh[RGS_ST_DAMAGE + 4 * a + v], q);
                    }
                }
        }
    }
#pragma unroll
    for (int k = 0; k < 14; ++k) {
        long long v = acc[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (lane == 0) atomicAdd(&sh[RGS_ST_EVENTS0 + k], (unsigned long long)v);
    }
    __syncthreads();
    if (threadIdx.x < RGS_N_STATS && sh[threadIdx.x] != 0ull) atomicAdd(&stats[threadIdx.x], sh[threadIdx.x]);
}

cudaError_t launch_stats(int n_team, int64_t n_games, const double *state, const long long *events, const double *damage,
                         int winning_score, long long *d_stats, cudaStream_t stream, int sm_count) {
    cudaError_t err = cudaMemsetAsync(d_stats, 0, sizeof(long long) * RGS_N_STATS, stream);
    if (err != cudaSuccess) return err;
    const int threads = 256;
    int64_t blocks = (n_games + threads - 1) / threads;
    if (blocks > (int64_t)sm_count * 4) blocks = (int64_t)sm_count * 4;
    if (blocks < 1) blocks = 1;
    stats_kernel<<<(unsigned)blocks, threads, 0, stream>>>(n_games, 2 + 4 * (2 * n_team + 1), 2 * n_team, state, events, damage,
                                                           winning_score, reinterpret_cast<unsigned long long *>(d_stats));
    return cudaGetLastError();
}

// ---- issue-roofline denominators: dependent-FMA-chain micro-kernels ------------------------------
// 8 independent chains per thread, 1024 threads x 2 CTAs per SM: enough ILP x TLP to saturate the
// FP32 (128 lanes/SM) or FP64 pipe.  Reported as FMA TFLOP/s (2 flop per FMA); the non-FMA lane-op
// rate that bounds the fmad=false parity kernel is half of it.
template <typename T>
__global__ void __launch_bounds__(1024) fma_peak_kernel(T *out, int iters, T a, T b) {
    T x0 = T(threadIdx.x), x1 = x0 + T(1), x2 = x0 + T(2), x3 = x0 + T(3), x4 = x0 + T(4), x5 = x0 + T(5), x6 = x0 + T(6),
      x7 = x0 + T(7);
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    const T s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == T(-12345)) out[0] = s;  // never true; keeps the chains alive
}

cudaError_t measure_fma_peak(int fp64, int sm_count, double *tflops) {
    void *scratch = nullptr;
    cudaError_t e = cudaMalloc(&scratch, 64);
    if (e != cudaSuccess) return e;
    cudaEvent_t t0, t1;
    cudaEventCreate(&t0);
    cudaEventCreate(&t1);
    const int blocks = sm_count * 2, threads = 1024, iters = fp64 ? 2000 : 4000;
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(t0);
        if (fp64) fma_peak_kernel<double><<<blocks, threads>>>(static_cast<double *>(scratch), iters, 0.999, 0.001);
        else fma_peak_kernel<float><<<blocks, threads>>>(static_cast<float *>(scratch), iters, 0.999f, 0.001f);
        cudaEventRecord(t1);
        e = cudaEventSynchronize(t1);
        if (e != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, t0, t1);
        const double flops = 2.0 * 64.0 * (double)iters * (double)blocks * (double)threads;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(t0);
    cudaEventDestroy(t1);
    cudaFree(scratch);
    if (e == cudaSuccess) e = cudaGetLastError();
    *tflops = best;
    return e;
}

}  // namespace rgs
