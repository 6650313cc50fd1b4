// rgs_kernels.h -- launchers of the sm_100a kernels (one translation unit per kernel family).
#pragma once
#include <cuda_runtime.h>

#include "rgs_device.cuh"

namespace rgs {

// K1: fp64, reference operator order (rgs_exact_kernel.cu)
cudaError_t launch_rollout_exact(int n_team, const DevParams<double> &P, const RunArgs &A, cudaStream_t stream,
                                 int sm_count);

// K2, one lane per game (rgs_lane_kernel.cu): the throughput kernel the FAST modes run; fp32 (throughput) and fp64
// (validation of the formulation against the oracle) instantiations.  A.lane_scratch must point at
// lane_scratch_bytes() bytes of device memory that no other launch in flight uses.
cudaError_t launch_rollout_lane_f32(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count);
cudaError_t launch_rollout_lane_f64(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count);
cudaError_t lane_scratch_bytes(int n_team, int fp64, int noise, int sm_count, size_t *bytes);

// K4: per-batch statistics vector int64[RGS_N_STATS] (rgs_aux_kernels.cu)
cudaError_t launch_stats(int n_team, int64_t n_games, const double *state, const long long *events, const double *damage,
                         int winning_score, long long *d_stats, cudaStream_t stream, int sm_count);

// K5: batched reset + synthetic S0 inputs (rgs_aux_kernels.cu)
cudaError_t launch_reset(int n_team, int64_t n_games, const rgs_params &p, const double *d_x0_ball, int64_t x0_stride,
                         const double *d_state0, int64_t state0_stride, double *state, double *clock, double *damage,
                         long long *events, cudaStream_t stream);

// K3 epilogue: terminal cost of every candidate + arg-min (rgs_aux_kernels.cu); scratch = 2 * kCostBlocks words
constexpr int kCostBlocks = 148 * 4;
cudaError_t launch_terminal_cost(int n_team, int64_t n_games, const double *state, const double *d_x_des, const double *d_w,
                                 double *d_cost, unsigned long long *scratch, long long *d_best, cudaStream_t stream);
cudaError_t launch_synth(uint64_t seed, int64_t first_game, int64_t n, int n_team, int64_t n_act, double u_max,
                         double *d_x0_ball, void *d_actions, int dtype, cudaStream_t stream);

// FMA-chain micro-benchmark: the measured FP32 / FP64 issue peak used as the roofline denominator
cudaError_t measure_fma_peak(int fp64, int sm_count, double *tflops);

// rgs_capi.cu's argument check, error message (rgs_last_error) and launch counter, for entry points in other files
int capi_check_params(const rgs_params *p);
int capi_fail(int code, const char *msg);
int capi_cuda_fail(cudaError_t e, const char *what);
void capi_count_launches(int n);

}  // namespace rgs
