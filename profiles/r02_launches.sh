#!/bin/bash
# ncu launch list of the bench command (device-resident leg + e2e legs), after the same command exited 0 without ncu
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/plain_launch.log 2>&1 || { tail -5 gpurun_out/plain_launch.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/ncu_launches.log 2>&1
echo "launch list exit $?"
