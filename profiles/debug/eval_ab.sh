#!/bin/bash
# A/B of the evaluate forward: pointer on tcgen05 (default) against the CUDA-core pointer inside the forward kernel
R4_EVAL_PTR_FWD=tc timeout 300 python -m pytest tests/test_train_gpu.py -x -q -m gpu 2>&1 | tail -4
for m in tc cuda; do
  echo "== R4_EVAL_PTR_FWD=$m"
  R4_EVAL_PTR_FWD=$m timeout 100 python profiles/debug/train_kernels.py 2>&1 | grep "device time\|r4::ev\|r4::evq" | cut -c1-100
done
