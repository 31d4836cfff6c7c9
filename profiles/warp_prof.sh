# per-warp timing of the selection / env phases (experiment build, then the product build is restored by the caller)
python profiles/phase_timing.py 2>&1 | tail -36
