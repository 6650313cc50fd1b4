// rgs_lane_launch.cuh -- K2 kernel + launch templates: fused multi-step rollout, throughput formulation, one
// LANE per game (see rgs_lane.cuh).  fp32 = throughput mode, fp64 = validation of the formulation.  Instantiated by
// the four translation units rgs_lane_{f32,f64}{,_noise}.cu (so they compile in parallel); rgs_lane_kernel.cu dispatches.
//
// Shared memory per CTA: the closed-form coefficient table + one scratch record per thread, laid
// out word-major ([word][thread]) so that the run-time body indices of the event path never cause
// a bank conflict (bank = lane) and never force per-thread arrays into local memory.
// Global memory per resident lane: LaneScratch::SCR_WORDS words (correctOverlap tracks + jam snapshot),
// contiguous per lane; touched by the event path only.
#pragma once
#include <type_traits>

#include "rgs_lane.cuh"
#include "rgs_host_params.h"
#include "rgs_kernels.h"

namespace rgs {

template <typename T, int NT, int THREADS, bool kPolicy = false, bool kNoise = false>
// fp32: at most 128 registers per lane, i.e. 16 resident warps per SM whatever the CTA size (the kernel is bound by
// instruction-fetch latency, which only resident warps hide: 12 instead of 16 warps cost 40 %, profiles/r02_*)
__global__ void __launch_bounds__(THREADS, sizeof(T) == 4 ? 512 / THREADS : 1) rollout_lane_kernel(const __grid_constant__ FastParams<T> P,
                                                                  const __grid_constant__ RunArgs A, const int lanes) {
    // `lanes` = games per CTA.  It equals THREADS except in the one-warp instantiation used for small batches
    // (launch_lane_nt), where a CTA may carry fewer games than it has threads: a latency-bound batch finishes sooner
    // on many thin warps (less divergence per warp, more warps per scheduler) than on a few full ones.
    using S = LaneScratch<T, NT>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LaneTab<T> *tab = reinterpret_cast<LaneTab<T> *>(smem_raw);
    constexpr size_t tab_bytes = (sizeof(LaneTab<T>) + 15) & ~size_t(15);
    // the record layout always has THREADS slots (a compile-time word stride keeps the address arithmetic out of the
    // code); in the one-warp instantiation only the first `lanes` of them carry a game
    const int per_cta = THREADS == 32 ? lanes : THREADS;
    T *fw = reinterpret_cast<T *>(smem_raw + tab_bytes);
    int *iw = reinterpret_cast<int *>(fw + (size_t)S::NF * THREADS);
    for (int i = threadIdx.x; i < 2 * kCols; i += THREADS) {
        (&tab->e[0][0])[i].g = (&P.tab_g[0][0])[i];
        (&tab->e[0][0])[i].k = (&P.tab_k[0][0])[i];
    }
    __syncthreads();
    if ((int)threadIdx.x >= per_cta) return;
    constexpr size_t scr_words = kNoise ? S::SCR_WORDS_NOISE : S::SCR_WORDS;
    T *scr = static_cast<T *>(A.lane_scratch) + ((size_t)blockIdx.x * THREADS + threadIdx.x) * scr_words;
    const S s{fw + threadIdx.x, iw + threadIdx.x, THREADS, scr};
    const int64_t stride = (int64_t)gridDim.x * per_cta;
    for (int64_t game = (int64_t)blockIdx.x * per_cta + threadIdx.x; game < A.n_games; game += stride)
        rollout_lane_game<T, NT, kPolicy, kNoise>(P, A, tab, s, game);
}

template <typename T, int NT>
constexpr size_t lane_smem_bytes(int threads) {
    return ((sizeof(LaneTab<T>) + 15) & ~size_t(15)) +
           ((size_t)LaneScratch<T, NT>::NF * sizeof(T) + (size_t)LaneScratch<T, NT>::NI * sizeof(int)) * threads;
}

// CTA size: measured on B200 (profiles/r01_ab_variants.md), 128 lanes per CTA (3 CTAs = 12 warps per SM for
// 2v2 fp32) run 10 % faster than 64 (7 CTAs); 256 and 512 are level with 128, 192 / 320 / 384 clearly worse.
// Records too large for two 128-lane CTAs per SM keep 64.
template <typename T, int NT>
constexpr int lane_big_threads() {
#ifdef RGS_AB_THREADS
    return RGS_AB_THREADS;
#else
    return lane_smem_bytes<T, NT>(128) <= 110 * 1024 ? 128 : 64;
#endif
}

#ifdef RGS_AB_SMALL_WARPS
constexpr int kSmallWarpsPerSm = RGS_AB_SMALL_WARPS;
#else
constexpr int kSmallWarpsPerSm = 7;   // one-warp CTAs per SM aimed at by the small-batch path (see launch_lane_nt)
#endif

template <typename T, int NT, int kThreads, bool kPolicy, bool kNoise>
static cudaError_t lane_resident_ctas(int lanes, int *per_sm) {
    (void)lanes;
    const size_t smem = lane_smem_bytes<T, NT>(kThreads);
    auto kern = rollout_lane_kernel<T, NT, kThreads, kPolicy, kNoise>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(per_sm, kern, kThreads, smem);
    if (err == cudaSuccess && *per_sm < 1) *per_sm = 1;
    return err;
}

template <typename T, int NT, int kThreads, bool kPolicy = false, bool kNoise = false>
static cudaError_t launch_lane_sized(const FastParams<T> &P, const RunArgs &A, cudaStream_t stream, int sm_count,
                                     int lanes = kThreads) {
    int per_sm = 0;
    cudaError_t err = lane_resident_ctas<T, NT, kThreads, kPolicy, kNoise>(lanes, &per_sm);
    if (err != cudaSuccess) return err;
    int64_t blocks = (A.n_games + lanes - 1) / lanes;
    const int64_t cap = (int64_t)sm_count * per_sm;  // persistent grid: resident CTAs x SM count
    if (blocks > cap) blocks = cap;  // (equalising the rounds per lane was measured: no gain, profiles/r01_ab_variants.md)
    if (blocks < 1) blocks = 1;
    const size_t words = kNoise ? LaneScratch<T, NT>::SCR_WORDS_NOISE : LaneScratch<T, NT>::SCR_WORDS;
    const size_t need = (size_t)blocks * kThreads * words * sizeof(T);
    if (!A.lane_scratch || A.lane_scratch_bytes < need) return cudaErrorInvalidValue;
    rollout_lane_kernel<T, NT, kThreads, kPolicy, kNoise><<<(unsigned)blocks, kThreads, lane_smem_bytes<T, NT>(kThreads), stream>>>(P, A, lanes);
    return cudaGetLastError();
}

// Batches smaller than one big CTA per SM (the K3 sampler's 4096 candidates) use one-warp CTAs carrying only a few
// games each, so that they spread over all SMs.
template <typename T, int NT, bool kNoise>
static cudaError_t launch_lane_nt(const FastParams<T> &P, const RunArgs &A, cudaStream_t stream, int sm_count) {
    constexpr int kBig = lane_big_threads<T, NT>();
    if (A.synth == 2) {  // closed loop on the on-device policy (with process noise: the reference's 2v2 only)
        if constexpr (NT == 2 || !kNoise) return launch_lane_sized<T, NT, kBig, true, kNoise>(P, A, stream, sm_count);
        else return cudaErrorNotSupported;
    }
    if (A.n_games < (int64_t)sm_count * kBig) {
        // games per one-warp CTA: aim at 7 warps per SM.  Measured on the 4096 x 50-step sampler tick
        // (profiles/r01_configs.json): 32 lanes 2.36 ms, 8 lanes 1.84, 4 lanes (6.9 warps/SM) 1.54, 2 lanes 1.87,
        // 1 lane (28 warps/SM, instruction-fetch-bound) 2.64.
        int lanes = (int)((A.n_games + (int64_t)sm_count * kSmallWarpsPerSm - 1) / ((int64_t)sm_count * kSmallWarpsPerSm));
        if (lanes < 1) lanes = 1;
        if (lanes > 32) lanes = 32;
        return launch_lane_sized<T, NT, 32, false, kNoise>(P, A, stream, sm_count, lanes);
    }
    return launch_lane_sized<T, NT, kBig, false, kNoise>(P, A, stream, sm_count);
}

// upper bound of the scratch any launch of this (T, NT) can need on a device with sm_count SMs
template <typename T, int NT, bool kNoise>
static cudaError_t lane_scratch_nt(int sm_count, size_t *bytes) {
    constexpr int kBig = lane_big_threads<T, NT>();
    int per_sm = 0;
    cudaError_t err = lane_resident_ctas<T, NT, kBig, false, kNoise>(kBig, &per_sm);
    if (err != cudaSuccess) return err;
    int64_t threads = (int64_t)per_sm * kBig;
    if constexpr (NT == 2 || !kNoise) {
        err = lane_resident_ctas<T, NT, kBig, true, kNoise>(kBig, &per_sm);
        if (err != cudaSuccess) return err;
        if ((int64_t)per_sm * kBig > threads) threads = (int64_t)per_sm * kBig;
    }
    // small-batch path: at most kSmallWarpsPerSm (+1 for the rounding of games per CTA) one-warp CTAs per SM
    if ((int64_t)(kSmallWarpsPerSm + 1) * 32 > threads) threads = (int64_t)(kSmallWarpsPerSm + 1) * 32;
    const size_t words = kNoise ? LaneScratch<T, NT>::SCR_WORDS_NOISE : LaneScratch<T, NT>::SCR_WORDS;
    *bytes = (size_t)sm_count * (size_t)threads * words * sizeof(T);
    return cudaSuccess;
}

template <typename T, typename F>
static cudaError_t by_team(int n_team, F &&f) {
    switch (n_team) {
#ifdef RGS_AB_ONLY_2V2   // size / A-B builds only (tools/sass_size.sh): compile the reference's team size alone
    case 2: return f(std::integral_constant<int, 2>{});
    default: return cudaErrorInvalidValue;
#else
    case 1: return f(std::integral_constant<int, 1>{});
    case 2: return f(std::integral_constant<int, 2>{});
    case 3: return f(std::integral_constant<int, 3>{});
    case 4: return f(std::integral_constant<int, 4>{});
    case 5: return f(std::integral_constant<int, 5>{});
    case 6: return f(std::integral_constant<int, 6>{});
    case 7: return f(std::integral_constant<int, 7>{});
    default: return cudaErrorInvalidValue;
#endif
    }
}

// one translation unit = one (T, kNoise): launch_lane_<tag> / scratch_lane_<tag>
#define RGS_LANE_TU(TAG, TYPE, NOISE)                                                                                     \
    cudaError_t launch_lane_##TAG(int n_team, const FastParams<TYPE> &P, const RunArgs &A, cudaStream_t stream, int sm_count) { \
        return by_team<TYPE>(n_team, [&](auto nt) { return launch_lane_nt<TYPE, decltype(nt)::value, NOISE>(P, A, stream, sm_count); }); \
    }                                                                                                                     \
    cudaError_t scratch_lane_##TAG(int n_team, int sm_count, size_t *bytes) {                                            \
        return by_team<TYPE>(n_team, [&](auto nt) { return lane_scratch_nt<TYPE, decltype(nt)::value, NOISE>(sm_count, bytes); }); \
    }

}  // namespace rgs
