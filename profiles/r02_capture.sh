#!/bin/bash
# round 2 evidence in one GPU call: gpu tests, smoke, bench line (N=1), bench at the 8-GPU shard size, reference arm,
# parity report, and ONE ncu pass (launch list was taken by profiles/r02_launches.sh): --set full of the rollout kernel
set -o pipefail
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout=900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python bench.py --steps 20 --warmup 3 --batch 128 --no-extras --no-cpu-baseline > gpurun_out/bench_b128.json 2>> gpurun_out/bench.err; echo "bench128 exit $?"
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"
python profiles/parity_report.py > gpurun_out/parity.log 2>&1; echo "parity exit $?"
python profiles/tc2_timing.py 128 148 256 296 512 1024 > gpurun_out/tc2_timing.log 2>&1; echo "tc2 timing exit $?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:rollout_tc -s 3 -c 1 -f -o gpurun_out/rollout_tc_r02 \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
