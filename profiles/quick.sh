python -m pytest tests/test_tc_gpu.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('kernel_ms',d['roofline']['kernel_ms'],'frac',d['roofline']['frac'],'e2e_ms',d['e2e']['ms_per_step'])"
