// rgs_lane_f32_noise.cu -- K2 lane kernels, T = float, process noise on (see rgs_lane_launch.cuh)
#include "rgs_lane_launch.cuh"

namespace rgs {
RGS_LANE_TU(f32_noise, float, true)
}  // namespace rgs
