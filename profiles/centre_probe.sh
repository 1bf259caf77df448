#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -12
for w in C5 C4; do timeout 100 python profiles/run_once.py $w 3 2>&1 | tail -1; done
timeout 100 python profiles/run_once.py C5 3 strong=8 2>&1 | tail -1
