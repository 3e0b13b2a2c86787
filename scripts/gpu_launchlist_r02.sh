#!/bin/bash
# ncu launch list of the bench command with the final library (one ncu pass; the plain run first, as the recipe asks).
# Usage: gpurun --timeout 700 -- 'bash scripts/gpu_launchlist_r02.sh TAG'
TAG=${1:-r02v}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $OUT/bench_short.json 2> $OUT/bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $OUT/ncu_launches.log 2>&1
echo "exit $?"; cat $OUT/bench_short.json; tail -2 $OUT/ncu_launches.log; wc -l $OUT/launches.csv
