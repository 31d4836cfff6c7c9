mkdir -p gpurun_out
timeout 300 python profiles/train_timing.py 64 > gpurun_out/f1_pre.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:eval_ -s 9 -c 3 -f -o gpurun_out/f1_eval \
    python profiles/train_timing.py 64 > gpurun_out/ncu_f1.log 2>&1
echo "ncu exit $?"
