// rgs_phase_kernel.cu -- K2 launch wrapper, phased block kernel (see rgs_phase.cuh): the throughput
// formulation with the narrow-phase items and the popped events of a whole thread block compacted
// onto consecutive threads.  fp32 = throughput mode, fp64 = validation of the formulation.
//
// Shared memory per CTA: coefficient table + a pool of scratch records (word-major, bank = record
// id) + the item / event lists.  The game a thread owns stays in registers between steps.
#include "rgs_phase.cuh"
#include "rgs_host_params.h"
#include "rgs_kernels.h"

namespace rgs {

template <typename T, int NT, int THREADS>
__global__ void __launch_bounds__(THREADS) rollout_phase_kernel(const __grid_constant__ FastParams<T> P,
                                                                const __grid_constant__ RunArgs A, int n_rec, int cap_items) {
    using Sh = PhaseShared<T, NT>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Sh sh;
    sh.carve(smem_raw, n_rec, cap_items);
    const int tid = threadIdx.x;
    {
        LaneTab<T> *tab = reinterpret_cast<LaneTab<T> *>(smem_raw);
        for (int i = tid; i < 2 * kCols; i += THREADS) {
            (&tab->e[0][0])[i].g = (&P.tab_g[0][0])[i];
            (&tab->e[0][0])[i].k = (&P.tab_k[0][0])[i];
        }
        for (int i = tid; i < n_rec; i += THREADS) sh.free_stack[i] = i;
        if (tid < 8) sh.ctr[tid] = tid == Sh::FREE_TOP ? n_rec : 0;
    }
    __syncthreads();
    const PhaseKernel<T, NT> K(P, A, sh);
    PhaseSlot<T, NT> ts;
    K.init_slot(ts, (int64_t)blockIdx.x * THREADS + tid);
    const int64_t stride = (int64_t)gridDim.x * THREADS;
    for (;;) {
        K.head(ts, stride);
        int more = __syncthreads_or(ts.rec >= 0);
        K.release_held(ts);  // pool pushes: SEARCH / POP / RESOLVE only; pool pops: HEAD only
        if (!more) {
            if (__syncthreads_and(K.slot_done(ts))) break;
            continue;
        }
#pragma unroll 1
        for (int r = 0; r < kPhaseRounds && more; ++r) {
            K.search(tid, THREADS);
            __syncthreads();
            if (tid == 0) { sh.ctr[Sh::N_WALL] = 0; sh.ctr[Sh::N_PAIR] = 0; }
            K.pop_phase(ts);
            __syncthreads();
            K.resolve_phase(tid, THREADS);
            more = __syncthreads_or(ts.rec >= 0);
            if (tid == 0) { sh.ctr[Sh::N_WALL_EV] = 0; sh.ctr[Sh::N_PAIR_EV] = 0; }
        }
    }
}

template <typename T, int NT>
static cudaError_t launch_phase_nt(const FastParams<T> &P, const RunArgs &A, cudaStream_t stream, int sm_count) {
    constexpr int kThreads = 256;
    constexpr int kRec = kThreads / 2;        // 31 % of the slots hold a record in an average S0 step
    constexpr int kCapItems = 2 * kThreads;
    if (A.n_games >= (int64_t(1) << 31)) return cudaErrorInvalidValue;
    const size_t smem = PhaseShared<T, NT>::bytes(kRec, kCapItems);
    auto kern = rollout_phase_kernel<T, NT, kThreads>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    int per_sm = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) per_sm = 1;
    int64_t blocks = (A.n_games + kThreads - 1) / kThreads;
    const int64_t cap = (int64_t)sm_count * per_sm;  // persistent grid: resident CTAs x SM count
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    kern<<<(unsigned)blocks, kThreads, smem, stream>>>(P, A, kRec, kCapItems);
    return cudaGetLastError();
}

template <typename T>
static cudaError_t launch_phase(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    const FastParams<T> P = make_fast_params<T>(p);
    switch (n_team) {
#ifndef RGS_EXP_ONLY_2V2
    case 1: return launch_phase_nt<T, 1>(P, A, stream, sm_count);
#endif
    case 2: return launch_phase_nt<T, 2>(P, A, stream, sm_count);
#ifndef RGS_EXP_ONLY_2V2
    case 3: return launch_phase_nt<T, 3>(P, A, stream, sm_count);
#endif
    default: return cudaErrorInvalidValue;  // larger teams run the one-lane-per-game kernel
    }
}

cudaError_t launch_rollout_phase_f32(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    return launch_phase<float>(n_team, p, A, stream, sm_count);
}
cudaError_t launch_rollout_phase_f64(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    return launch_phase<double>(n_team, p, A, stream, sm_count);
}

}  // namespace rgs
