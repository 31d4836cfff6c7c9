#!/bin/bash
# evaluate kernels: parity tests, then the per-kernel device times of the c5 training step
timeout 300 python -m pytest tests/test_train_gpu.py -x -q -m gpu 2>&1 | tail -4
timeout 100 python profiles/debug/train_kernels.py 2>&1 | grep "device time\|r4::" | cut -c1-100
