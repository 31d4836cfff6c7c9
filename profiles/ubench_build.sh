#!/bin/bash
# build the micro-benchmarks (profiles/ubench_*.cu -> profiles/ubench_*.bin); extra nvcc flags pass through
cd "$(dirname "$0")/.."
F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I eam_rl4co_b200/csrc -I include"
nvcc $F "$@" -o profiles/ubench_phases.bin profiles/ubench_phases.cu eam_rl4co_b200/csrc/capi.cu eam_rl4co_b200/csrc/decode_kernels.cu
