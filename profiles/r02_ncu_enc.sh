#!/bin/bash
mkdir -p gpurun_out
python profiles/enc_timing.py > gpurun_out/enc_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:enc_ffn_kernel -s 3 -c 1 -f -o gpurun_out/enc_ffn python profiles/enc_timing.py > gpurun_out/ncu_enc_ffn.log 2>&1; echo "ncu ffn exit $?"
ncu --set full --clock-control none --import-source on -k regex:enc_attn_block_kernel -s 3 -c 1 -f -o gpurun_out/enc_attn python profiles/enc_timing.py > gpurun_out/ncu_enc_attn.log 2>&1; echo "ncu attn exit $?"
