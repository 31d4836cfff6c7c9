python profiles/phase_timing.py 2>&1 | tail -19
