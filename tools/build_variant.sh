#!/bin/bash
# tools/build_variant.sh NAME "-DFLAG ..."  -> exp/librgs_NAME.so : an A/B build of the two K2 kernels (2v2 only),
# linked with the already built objects of the other translation units.  Kernel experiments only.
set -e
cd "$(dirname "$0")/../robo_game_sim_b200/csrc"
N=$1; shift
mkdir -p ../../exp/$N
FLAGS="-std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr -fmad=false -DRGS_EXP_ONLY_2V2 $*"
nvcc $FLAGS -c rgs_lane_kernel.cu -o ../../exp/$N/lane.o &
nvcc $FLAGS -c rgs_phase_kernel.cu -o ../../exp/$N/phase.o &
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../exp/librgs_$N.so ../../exp/$N/lane.o ../../exp/$N/phase.o build/rgs_exact_kernel.o build/rgs_fast_kernel.o build/rgs_aux_kernels.o build/rgs_capi.o
cuobjdump -res-usage ../../exp/librgs_$N.so 2>/dev/null | grep -A1 -E "rollout_(lane|phase)_kernelIfLi2E" | grep -E "REG" | cut -c1-60
