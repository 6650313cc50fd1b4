// rgs_lane_f64.cu -- K2 lane kernels, T = double, process noise off (see rgs_lane_launch.cuh)
#include "rgs_lane_launch.cuh"

namespace rgs {
RGS_LANE_TU(f64, double, false)
}  // namespace rgs
