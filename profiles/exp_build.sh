#!/bin/bash
# build an experiment variant of the library: profiles/exp_build.sh -DR4_X_...   (no flags = the product build)
cd "$(dirname "$0")/.."
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared "$@" \
  -o eam_rl4co_b200/librl4co_b200.so eam_rl4co_b200/csrc/*.cu
