#!/bin/bash
# round 2, first GPU call: full gpu test-suite (no -x, to see every failure), smoke, bench (N=1, strong == weak),
# bench at the 8-GPU shard size (128 instances) on one GPU, parity report
set -o pipefail
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --timeout=900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/smoke.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
python bench.py --steps 20 --warmup 3 --batch 128 --no-extras --no-cpu-baseline > gpurun_out/bench_b128.json 2> gpurun_out/bench_b128.err; echo "bench128 exit $?"
cat gpurun_out/bench_b128.json
python profiles/parity_report.py > gpurun_out/parity.log 2>&1; echo "parity exit $?"; tail -30 gpurun_out/parity.log
