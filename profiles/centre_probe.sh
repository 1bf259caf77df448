#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "narrow" 2>&1 | tail -40
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_cells.json 2> $O/bench_cells.err; echo "bench rc=$?"; tail -3 $O/bench_cells.err
