#!/bin/bash
# tools/build_variant.sh NAME [-DFLAG ...]: alternative fp32 build of the kernel source for timing experiments
# (tools/prof_step.py reads GO1_LIB).  Output: tools/variants/libgo1sim_NAME.so (git-ignored, travels with gpurun).
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p tools/variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared --expt-relaxed-constexpr \
  -prec-div=false -prec-sqrt=false -ftz=true -DGO1_VARIANT "$@" -o tools/variants/libgo1sim_$name.so quadruped_rl_locomotion_b200/csrc/go1sim.cu
echo built tools/variants/libgo1sim_$name.so
