#!/bin/bash
# one-off: narrow-window Taylor sums of the sorted kernel (c') -- parity and timing at C5
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q -k "sorted or narrow or regime or C5 or options" 2>&1 | tail -5
for c in 0 1 2; do timeout 300 python profiles/run_once.py C5 3 centre=$c 2>&1 | tail -1; done
timeout 300 python profiles/run_once.py C5 3 strong=8 2>&1 | tail -1
timeout 300 python profiles/run_once.py C5 3 strong=8 centre=0 2>&1 | tail -1
