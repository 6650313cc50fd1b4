#!/bin/bash
# tools/sass_size.sh [FLAGS] : SASS instruction count of the 2v2 fp32 K2 kernels (quick: compiles the 2v2 instantiations only)
cd "$(dirname "$0")/../robo_game_sim_b200/csrc"
nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr -fmad=false -DRGS_AB_ONLY_2V2 $1 -c rgs_lane_f32.cu -o /tmp/sass_size.o || exit 1
cuobjdump -sass /tmp/sass_size.o 2>/dev/null | awk '/Function :/{name=$3} /^[ \t]+\/\*[0-9a-f]+\*\/[ \t]+[A-Z@]/{n[name]++} END{for(k in n) printf "%6d %s\n", n[k], substr(k,1,100)}' | sort -rn
cuobjdump -res-usage /tmp/sass_size.o 2>/dev/null | grep -A1 "Li128ELb0ELb0" | grep REG | cut -c1-40
