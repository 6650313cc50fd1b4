// rgs_lane_kernel.cu -- K2 launch dispatch (the kernels are instantiated in rgs_lane_{f32,f64}{,_noise}.cu).
// Process noise (rgs_params.noise != 0) selects the kNoise instantiations, so that the noise-free kernels do not carry
// the generator's code.
#include "rgs_fast.cuh"
#include "rgs_host_params.h"
#include "rgs_kernels.h"

namespace rgs {

#define RGS_LANE_DECL(TAG, TYPE)                                                                                          \
    cudaError_t launch_lane_##TAG(int n_team, const FastParams<TYPE> &P, const RunArgs &A, cudaStream_t stream, int sm_count); \
    cudaError_t scratch_lane_##TAG(int n_team, int sm_count, size_t *bytes);
RGS_LANE_DECL(f32, float)
RGS_LANE_DECL(f64, double)
RGS_LANE_DECL(f32_noise, float)
RGS_LANE_DECL(f64_noise, double)

cudaError_t launch_rollout_lane_f32(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    const FastParams<float> P = make_fast_params<float>(p);
    return p.noise != 0.0 ? launch_lane_f32_noise(n_team, P, A, stream, sm_count) : launch_lane_f32(n_team, P, A, stream, sm_count);
}
cudaError_t launch_rollout_lane_f64(int n_team, const rgs_params &p, const RunArgs &A, cudaStream_t stream, int sm_count) {
    const FastParams<double> P = make_fast_params<double>(p);
    return p.noise != 0.0 ? launch_lane_f64_noise(n_team, P, A, stream, sm_count) : launch_lane_f64(n_team, P, A, stream, sm_count);
}
cudaError_t lane_scratch_bytes(int n_team, int fp64, int noise, int sm_count, size_t *bytes) {
    if (fp64) return noise ? scratch_lane_f64_noise(n_team, sm_count, bytes) : scratch_lane_f64(n_team, sm_count, bytes);
    return noise ? scratch_lane_f32_noise(n_team, sm_count, bytes) : scratch_lane_f32(n_team, sm_count, bytes);
}

}  // namespace rgs
