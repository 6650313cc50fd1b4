// rgs_lane_kernel.cu -- K2 launch wrapper: fused multi-step rollout, throughput formulation, one
// LANE per game (see rgs_lane.cuh).  fp32 = throughput mode, fp64 = validation of the formulation.
//
// Shared memory per CTA: the closed-form coefficient table + one scratch record per thread, laid
// out word-major ([word][thread]) so that the run-time body indices of the event path never cause
// a bank conflict (bank = lane) and never force per-thread arrays into local memory.
#include <cstdlib>

#include "rgs_lane.cuh"
#include "rgs_host_params.h"
#include "rgs_kernels.h"

namespace rgs {

template <typename T, int NT, int THREADS, bool kPolicy = false>
#ifndef RGS_LANE_MINBLOCKS
#define RGS_LANE_MINBLOCKS 1
#endif
__global__ void __launch_bounds__(THREADS, RGS_LANE_MINBLOCKS) rollout_lane_kernel(const __grid_constant__ FastParams<T> P,
                                                               const __grid_constant__ RunArgs A, const int lanes, const int batch) {
    // `lanes` = games per CTA.  It equals THREADS except in the one-warp instantiation used for small batches
    // (launch_lane_nt), where a CTA may carry fewer games than it has threads: a latency-bound batch finishes sooner
    // on many thin warps (less divergence per warp, more warps per scheduler) than on a few full ones.
    using S = LaneScratch<T, NT>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    LaneTab<T> *tab = reinterpret_cast<LaneTab<T> *>(smem_raw);
    constexpr size_t tab_bytes = (sizeof(LaneTab<T>) + 15) & ~size_t(15);
    const int per_cta = THREADS == 32 ? lanes : THREADS;
    T *fw = reinterpret_cast<T *>(smem_raw + tab_bytes);
    int *iw = reinterpret_cast<int *>(fw + (size_t)S::NF * per_cta);
    for (int i = threadIdx.x; i < 2 * kCols; i += THREADS) {
        (&tab->e[0][0])[i].g = (&P.tab_g[0][0])[i];
        (&tab->e[0][0])[i].k = (&P.tab_k[0][0])[i];
    }
    __syncthreads();
    const S s{fw + threadIdx.x, iw + threadIdx.x, per_cta};
    const int64_t stride = (int64_t)gridDim.x * per_cta;
#ifdef RGS_LANE_DEB
    // warp-scheduled form: every thread of the warp votes, threads beyond per_cta own no game
    const int64_t game = (int)threadIdx.x < per_cta ? (int64_t)blockIdx.x * per_cta + threadIdx.x : A.n_games;
    rollout_lane_warp<T, NT>(P, A, tab, s, game, stride, batch);
#else
    if ((int)threadIdx.x >= per_cta) return;
    for (int64_t game = (int64_t)blockIdx.x * per_cta + threadIdx.x; game < A.n_games; game += stride)
        rollout_lane_game<T, NT, kPolicy>(P, A, tab, s, game);
#endif
}

template <typename T, int NT>
constexpr size_t lane_smem_bytes(int threads) {
    return ((sizeof(LaneTab<T>) + 15) & ~size_t(15)) +
           ((size_t)LaneScratch<T, NT>::NF * sizeof(T) + (size_t)LaneScratch<T, NT>::NI * sizeof(int)) * threads;
}

template <typename T, int NT, int kThreads, bool kPolicy = false>
static cudaError_t launch_lane_sized(const FastParams<T> &P, const RunArgs &A, cudaStream_t stream, int sm_count,
                                     int lanes = kThreads) {
    size_t smem = lane_smem_bytes<T, NT>(lanes);
#ifdef RGS_EXP_ONLY_2V2
    if (const char *pad = std::getenv("RGS_EXP_SMEM_PAD")) smem += (size_t)std::atoi(pad);  // occupancy sensitivity experiments
#endif
    auto kern = rollout_lane_kernel<T, NT, kThreads, kPolicy>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    int per_sm = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) per_sm = 1;
    int64_t blocks = (A.n_games + lanes - 1) / lanes;
    const int64_t cap = (int64_t)sm_count * per_sm;  // persistent grid: resident CTAs x SM count
    if (blocks > cap) blocks = cap;  // (equalising the rounds per lane was measured: no gain, profiles/r01_ab_variants.md)
    if (blocks < 1) blocks = 1;
    int batch = 16;  // lanes parked in the event loop before the warp runs an event iteration (RGS_LANE_DEB builds)
    if (const char *e = std::getenv("RGS_DEB_BATCH")) batch = std::atoi(e);
    kern<<<(unsigned)blocks, kThreads, smem, stream>>>(P, A, lanes, batch);
    return cudaGetLastError();
}

// CTA size: measured on B200 (profiles/r01_ab_variants.md), 128 lanes per CTA (3 CTAs = 12 warps per SM for
// 2v2 fp32) run 10 % faster than 64 (7 CTAs); 256 and 512 are level with 128, 192 / 320 / 384 clearly worse.
// Records too large for two 128-lane CTAs per SM keep 64; batches smaller than one 128-lane CTA per SM (the
// K3 sampler's 4096 candidates) use one-warp CTAs carrying only a few games each, so that they spread over all SMs.
template <typename T, int NT>
static cudaError_t launch_lane_nt(const FastParams<T> &P, const RunArgs &A, cudaStream_t stream, int sm_count) {
#ifdef RGS_LANE_THREADS
    return launch_lane_sized<T, NT, RGS_LANE_THREADS>(P, A, stream, sm_count);
#else
    constexpr int kBig = lane_smem_bytes<T, NT>(128) <= 110 * 1024 ? 128 : 64;
    if (A.synth == 2) {  // closed loop on the on-device policy: instantiated for the reference's 2v2 only
        if constexpr (NT == 2) return launch_lane_sized<T, NT, kBig, true>(P, A, stream, sm_count);
        else return cudaErrorNotSupported;
    }
    if (A.n_games < (int64_t)sm_count * kBig) {
        // games per one-warp CTA: aim at 7 warps per SM.  Measured on the 4096 x 50-step sampler tick
        // (profiles/r01_configs.json): 32 lanes 2.36 ms, 8 lanes 1.84, 4 lanes (6.9 warps/SM) 1.54, 2 lanes 1.87,
        // 1 lane (28 warps/SM, instruction-fetch-bound) 2.64.
        int lanes = (int)((A.n_games + (int64_t)sm_count * 7 - 1) / ((int64_t)sm_count * 7));
        if (const char *e = std::getenv("RGS_SMALL_LANES")) lanes = std::atoi(e);  // tuning knob
        if (lanes < 1) lanes = 1;
        if (lanes > 32) lanes = 32;
        return launch_lane_sized<T, NT, 32>(P, A, stream, sm_count, lanes);
    }
    return launch_lane_sized<T, NT, kBig>(P, A, stream, sm_count);
#endif
}

template <typename T>
static cudaError_t launch_lane(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    const FastParams<T> P = make_fast_params<T>(p);
    switch (n_team) {
#ifndef RGS_EXP_ONLY_2V2
    case 1: return launch_lane_nt<T, 1>(P, A, stream, sm_count);
#endif
    case 2: return launch_lane_nt<T, 2>(P, A, stream, sm_count);
#ifndef RGS_EXP_ONLY_2V2
    case 3: return launch_lane_nt<T, 3>(P, A, stream, sm_count);
    case 4: return launch_lane_nt<T, 4>(P, A, stream, sm_count);
    case 5: return launch_lane_nt<T, 5>(P, A, stream, sm_count);
    case 6: return launch_lane_nt<T, 6>(P, A, stream, sm_count);
    case 7: return launch_lane_nt<T, 7>(P, A, stream, sm_count);
#endif
    default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_rollout_lane_f32(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    return launch_lane<float>(n_team, p, A, stream, sm_count);
}
cudaError_t launch_rollout_lane_f64(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    return launch_lane<double>(n_team, p, A, stream, sm_count);
}

}  // namespace rgs
