#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout=900 2>&1 | tail -5
python bench.py --no-cpu-baseline --no-extras > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err; tail -3 gpurun_out/bench_g.err
python bench.py --no-cpu-baseline --no-extras --batch 128 > gpurun_out/bench_g128.json 2>> gpurun_out/bench_g.err
python - <<PY
import json
for f in ("gpurun_out/bench_g.json","gpurun_out/bench_g128.json"):
    d=json.load(open(f)); print(f, "value_ms", round(d["ms_per_step"],3), "kernel_ms", round(d["roofline"]["kernel_ms"],3), "e2e_ms", round(d["e2e"]["ms_per_step"],3), "uncaptured", round(d["e2e"]["uncaptured_ms_per_step"],3))
PY
