// rgs_exact_kernel.cu -- K1 launch wrapper: fused multi-step fp64 rollout (see rgs_exact.cuh).
// Built with -fmad=false: one IEEE-754 rounding per source operator, as the reference evaluates.
#include "rgs_exact.cuh"
#include "rgs_kernels.h"

namespace rgs {

// resident CTAs per SM: the RK4 chains are latency-bound (largest stall: `wait` on dependent fp64 operations), so for
// the small grids (1v1, 2v2: 35 KB of shared memory per CTA) five CTAs at <= 96 registers beat four at 112
// (+7 %, profiles/r02_ab_variants.md); six (<= 80 registers, spills) do not.  Larger teams are shared-memory-limited.
#ifndef RGS_AB_K1_MINBLOCKS
#define RGS_AB_K1_MINBLOCKS (NT <= 2 ? 5 : 1)
#endif
template <int NT>
__global__ void __launch_bounds__(128, RGS_AB_K1_MINBLOCKS) rollout_exact_kernel(const __grid_constant__ DevParams<double> P,
                                                            const __grid_constant__ RunArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    WarpSlice<double, NT> *W = reinterpret_cast<WarpSlice<double, NT> *>(smem_raw) + warp;
    rollout_exact_warp<NT>(P, A, W, lane, (int64_t)blockIdx.x * wpb + warp, (int64_t)gridDim.x * wpb);
}

// ---- host-side launcher (called from rgs_capi.cu) ------------------------------------------
template <int NT>
static cudaError_t launch_exact_nt(const DevParams<double> &P, const RunArgs &A, cudaStream_t stream, int sm_count) {
    constexpr int kThreads = 128;
    const size_t smem = sizeof(WarpSlice<double, NT>) * (kThreads / 32);
    auto kern = rollout_exact_kernel<NT>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    int per_sm = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) per_sm = 1;
    int64_t blocks = (A.n_games + (kThreads / 32) - 1) / (kThreads / 32);
    const int64_t cap = (int64_t)sm_count * per_sm;  // persistent: a whole number of waves of the SM count
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    kern<<<(unsigned)blocks, kThreads, smem, stream>>>(P, A);
    return cudaGetLastError();
}

cudaError_t launch_rollout_exact(int n_team, const DevParams<double> &P, const RunArgs &A, cudaStream_t stream,
                                 int sm_count) {
    switch (n_team) {
    case 1: return launch_exact_nt<1>(P, A, stream, sm_count);
    case 2: return launch_exact_nt<2>(P, A, stream, sm_count);
    case 3: return launch_exact_nt<3>(P, A, stream, sm_count);
    case 4: return launch_exact_nt<4>(P, A, stream, sm_count);
    case 5: return launch_exact_nt<5>(P, A, stream, sm_count);
    case 6: return launch_exact_nt<6>(P, A, stream, sm_count);
    case 7: return launch_exact_nt<7>(P, A, stream, sm_count);
    default: return cudaErrorInvalidValue;
    }
}

}  // namespace rgs
