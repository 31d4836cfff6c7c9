mkdir -p gpurun_out
export RL4CO_TC_WIDE=1
timeout 200 python profiles/tc2_timing.py 296 > gpurun_out/tc3_pre.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:rollout_tc3 -s 3 -c 1 -f -o gpurun_out/rollout_tc3 \
    python profiles/tc2_timing.py 296 > gpurun_out/ncu_tc3.log 2>&1
echo "ncu exit $?"
