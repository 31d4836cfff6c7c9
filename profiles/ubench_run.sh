#!/bin/bash
# profiles/ubench_run.sh "<flags variant 1>" "<flags variant 2>" ... : build the phase micro-benchmark per variant and run on the GPU box
cd "$(dirname "$0")/.."
i=0
cmd=""
for v in "$@"; do
  F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I eam_rl4co_b200/csrc -I include"
  nvcc $F $v -o profiles/ubench_v$i.bin profiles/ubench_phases.cu eam_rl4co_b200/csrc/capi.cu eam_rl4co_b200/csrc/decode_kernels.cu || exit 1
  cmd="$cmd echo '=== variant: $v'; timeout 60 ./profiles/ubench_v$i.bin;"
  i=$((i+1))
done
/usr/local/graft/bin/gpurun --timeout 300 -- "$cmd" 2>&1 | grep -E "===|softmax|select|rror"
rm -f profiles/ubench_v*.bin
