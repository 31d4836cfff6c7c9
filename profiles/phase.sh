#!/bin/bash
# profiles/phase.sh [B] [tsp|cvrp] [decode_type] : in-kernel phase timing.  The counters are compiled in only with
# -DR4_PHASE_PROF (they cost registers the product kernel needs): build that variant, run profiles/phase_timing.py
# on the GPU box, restore the product build.
cd "$(dirname "$0")/.."
bash profiles/exp_build.sh -DR4_PHASE_PROF || exit 1
/usr/local/graft/bin/gpurun --timeout 400 -- "timeout 300 python profiles/phase_timing.py $* 2>&1 | tail -22" 2>&1 | tail -26
bash profiles/exp_build.sh
