#!/bin/bash
# profiles/variants.sh name1:"-DFLAG ..." name2:"..."  -> build_variants/<name>.so (experiment builds of the library,
# git-ignored, shipped to the GPU box).  Run one with RL4CO_B200_LIB=build_variants/<name>.so python ...
cd "$(dirname "$0")/.."
mkdir -p build_variants
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  [ "$flags" == "$spec" ] && flags=""
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared $flags \
    -o build_variants/$name.so eam_rl4co_b200/csrc/*.cu 2>&1 | grep -v "warning\|Remark\|^$\|\^\|declared but never" &
done
wait
ls -la build_variants/
