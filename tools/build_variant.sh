#!/bin/bash
# tools/build_variant.sh NAME "FLAGS" : A/B build of the K2 lane kernel with extra -D flags -> exp/librgs_NAME.so
# (exp/ is git-ignored; the shipped library never carries A/B switches at run time).  Use with RGS_B200_LIB=...
set -e
N=$1; FLAGS=$2
cd "$(dirname "$0")/../robo_game_sim_b200/csrc"
mkdir -p ../../exp/$N
nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr -fmad=false $FLAGS -c rgs_lane_f32.cu -o ../../exp/$N/lane.o
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../exp/librgs_$N.so ../../exp/$N/lane.o build/rgs_lane_kernel.o build/rgs_lane_f64.o build/rgs_lane_f32_noise.o build/rgs_lane_f64_noise.o build/rgs_exact_kernel.o build/rgs_aux_kernels.o build/rgs_noalloc.o build/rgs_capi.o
cuobjdump -res-usage ../../exp/librgs_$N.so 2>/dev/null | grep -A1 -E "rollout_lane_kernelIfLi2ELi128ELb0ELb0" | grep -E "REG" | cut -c1-60
