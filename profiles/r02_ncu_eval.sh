#!/bin/bash
# round 2: ncu --set full of the four evaluate kernels (c5 training step, CVRP100 POMO, B = 64), after the same command exited 0
mkdir -p gpurun_out
timeout 300 python profiles/train_timing.py 64 > gpurun_out/train_timing.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:eval_ -s 12 -c 4 -f -o gpurun_out/eval_r02 \
    python profiles/train_timing.py 64 > gpurun_out/ncu_eval.log 2>&1
echo "ncu exit $?"; grep -v Warn gpurun_out/train_timing.log
