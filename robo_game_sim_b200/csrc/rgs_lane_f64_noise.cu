// rgs_lane_f64_noise.cu -- K2 lane kernels, T = double, process noise on (see rgs_lane_launch.cuh)
#include "rgs_lane_launch.cuh"

namespace rgs {
RGS_LANE_TU(f64_noise, double, true)
}  // namespace rgs
