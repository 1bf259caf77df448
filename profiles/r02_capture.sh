#!/bin/bash
# Round-2 ncu captures (one GPU).  Run on the GPU box from the repo root:  bash profiles/r02_capture.sh
# Every program is run once WITHOUT ncu first (a number printed under ncu is never a bench value); each report is
# summarised on the box with profiles/summarize.py (the text comes back in gpurun_out/ and is copied to profiles/).
set -u
O=gpurun_out
mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on"
run() { # name, kernel regex, count, run_once arguments...   (ONLY="name name" in the environment selects captures)
    local name=$1 re=$2 cnt=$3; shift 3
    if [ -n "${ONLY:-}" ] && ! echo " $ONLY " | grep -q " $name "; then return; fi
    timeout 600 python profiles/run_once.py "$@" > $O/once_$name.log 2>&1 || { echo "run_once $name failed"; tail -3 $O/once_$name.log; return; }
    timeout 900 $NCU -k regex:"$re" -c $cnt -f -o $O/r02_$name python profiles/run_once.py "$@" > $O/ncu_$name.log 2>&1
    echo "$name: rc=$? $(tail -1 $O/once_$name.log | cut -c1-200)"
    # the reports together exceed what gpurun brings back (64 MiB): summarise here, keep the text
    python profiles/summarize.py $O/r02_$name.ncu-rep > $O/r02_${name}_full.txt 2>/dev/null && rm -f $O/r02_$name.ncu-rep
}
# launch list of the bench command itself
timeout 900 python bench.py --steps 2 --warmup 3 --cpu-seconds 0 > $O/bench_short.json 2> $O/bench_short.err
echo "bench short rc=$?"
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r02_launches_bench_C4.csv \
    python bench.py --steps 2 --warmup 3 --cpu-seconds 0 > $O/ncu_bench.log 2>&1
echo "bench launch list rc=$?"
run C4_production  "cell_|transit_kernel|column_cta|compact|sigma_phase|finalize|bucket" 10 C4 1
run C4_bruteforce  "transit_kernel" 1 C4 1 cull=0 mirror=0 thin=0
run C4_kernel_a    "number_density_rows" 1 C4 1 kernel_a
run C5_production  "cell_|transit_kernel|column_cta|compact|sigma_phase" 8 C5 1
run C5_sorted      "transit_kernel" 1 C5 1 centre=2
run C4_windows     "transit_kernel" 1 C4 1 sort=0
run C2_production  "moment|column_cta|fout" 8 C2 1
run C3_production  "transit_kernel|column_cta|compact" 6 C3 1
run RET_batch64    "transit_kernel|column_cta|compact|lightcurve" 8 RET 1 batch=64
ls -la $O/r02_*_full.txt
