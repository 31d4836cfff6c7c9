#!/bin/bash
# round 2, final kernels: ncu --set full of every evaluate kernel of ONE c5 training step (CVRP100 POMO, B = 64), after the same
# command exited 0 without ncu.  Ten launches per step match regex:eval_ (mask words, glimpse, pointer, query forward; pointer,
# glimpse row pass, dlkf, dgk / dgv, query backward); three warm-up steps are skipped.
mkdir -p gpurun_out
timeout 300 python profiles/train_timing.py 64 > gpurun_out/train_timing.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:eval_ -s 27 -c 9 -f -o gpurun_out/eval2_r02 \
    python profiles/train_timing.py 64 > gpurun_out/ncu_eval2.log 2>&1
echo "ncu exit $?"; grep -v Warn gpurun_out/train_timing.log | tail -8
