mkdir -p gpurun_out
export RL4CO_TC_WIDE=1
timeout 300 python -m pytest tests/test_tc_gpu.py -x -q -k "streaming" > gpurun_out/tc3_test.log 2>&1; echo "test rc=$?" >> gpurun_out/tc3_test.log
tail -15 gpurun_out/tc3_test.log
timeout 200 python profiles/tc2_timing.py 296 1024 > gpurun_out/tc3_timing_wide.log 2>&1; echo "rc=$?" >> gpurun_out/tc3_timing_wide.log
cat gpurun_out/tc3_timing_wide.log | grep -v Warn
