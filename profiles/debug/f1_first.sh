mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_train_gpu.py -x -q > gpurun_out/f1_test.log 2>&1; echo "test rc=$?" >> gpurun_out/f1_test.log
tail -25 gpurun_out/f1_test.log
timeout 300 python profiles/train_timing.py 64 > gpurun_out/f1_timing.log 2>&1; echo "rc=$?" >> gpurun_out/f1_timing.log
grep -v Warn gpurun_out/f1_timing.log
