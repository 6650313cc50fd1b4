// rgs_fast_kernel.cu -- K2 launch wrapper: fused multi-step rollout, closed-form sub-steps,
// lanes over grid columns (see rgs_fast.cuh).  fp32 = throughput mode, fp64 = validation mode.
#include "rgs_fast.cuh"
#include "rgs_host_params.h"
#include "rgs_kernels.h"

namespace rgs {

template <typename T, int NT, int THREADS>
__global__ void __launch_bounds__(THREADS) rollout_fast_kernel(const __grid_constant__ FastParams<T> P,
                                                               const __grid_constant__ RunArgs A) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FastTables<T> *tab = reinterpret_cast<FastTables<T> *>(smem_raw);
    constexpr size_t tab_bytes = (sizeof(FastTables<T>) + 15) & ~size_t(15);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = THREADS >> 5;
    for (int i = threadIdx.x; i < 2 * kCols; i += THREADS) {
        (&tab->g[0][0])[i] = (&P.tab_g[0][0])[i];
        (&tab->k[0][0])[i] = (&P.tab_k[0][0])[i];
    }
    __syncthreads();
    WarpSlice<T, NT> *W = reinterpret_cast<WarpSlice<T, NT> *>(smem_raw + tab_bytes) + warp;
    rollout_fast_warp<T, NT>(P, A, W, tab, lane, (int64_t)blockIdx.x * wpb + warp, (int64_t)gridDim.x * wpb);
}

template <typename T, int NT>
static cudaError_t launch_fast_nt(const FastParams<T> &P, const RunArgs &A, cudaStream_t stream, int sm_count) {
    constexpr int kThreads = 128;
    constexpr size_t tab_bytes = (sizeof(FastTables<T>) + 15) & ~size_t(15);
    const size_t smem = tab_bytes + sizeof(WarpSlice<T, NT>) * (kThreads / 32);
    auto kern = rollout_fast_kernel<T, NT, kThreads>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    int per_sm = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) per_sm = 1;
    int64_t blocks = (A.n_games + (kThreads / 32) - 1) / (kThreads / 32);
    const int64_t cap = (int64_t)sm_count * per_sm;  // persistent grid: resident CTAs x SM count
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    kern<<<(unsigned)blocks, kThreads, smem, stream>>>(P, A);
    return cudaGetLastError();
}

template <typename T>
static cudaError_t launch_fast(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    const FastParams<T> P = make_fast_params<T>(p);
    switch (n_team) {
    case 1: return launch_fast_nt<T, 1>(P, A, stream, sm_count);
    case 2: return launch_fast_nt<T, 2>(P, A, stream, sm_count);
    case 3: return launch_fast_nt<T, 3>(P, A, stream, sm_count);
    case 4: return launch_fast_nt<T, 4>(P, A, stream, sm_count);
    case 5: return launch_fast_nt<T, 5>(P, A, stream, sm_count);
    case 6: return launch_fast_nt<T, 6>(P, A, stream, sm_count);
    case 7: return launch_fast_nt<T, 7>(P, A, stream, sm_count);
    default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_rollout_fast_f32(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    return launch_fast<float>(n_team, p, A, stream, sm_count);
}
cudaError_t launch_rollout_fast_f64(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    return launch_fast<double>(n_team, p, A, stream, sm_count);
}

}  // namespace rgs
