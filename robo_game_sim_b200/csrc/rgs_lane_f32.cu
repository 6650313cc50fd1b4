// rgs_lane_f32.cu -- K2 lane kernels, T = float, process noise off (see rgs_lane_launch.cuh)
#include "rgs_lane_launch.cuh"

namespace rgs {
RGS_LANE_TU(f32, float, false)
}  // namespace rgs
