#!/bin/bash
# ncu --set full of the single-start rollout kernels: c2 / c3 on the streaming kernel, c1 on the cooperative kernel
mkdir -p gpurun_out
python profiles/s1_timing.py c1 c2 c3 --impl auto --reps 2 > gpurun_out/s1_plain.log 2>&1 || exit 1
cat gpurun_out/s1_plain.log
for c in c1 c2 c3; do
  ncu --set full --clock-control none --import-source on -k regex:rollout_ -s 2 -c 1 -f -o gpurun_out/s1_$c \
      python profiles/s1_timing.py $c --impl auto --reps 2 > gpurun_out/ncu_s1_$c.log 2>&1
  echo "ncu $c exit $?"
done
