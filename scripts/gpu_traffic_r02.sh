#!/bin/bash
# DRAM traffic of ONE eval_kernel launch at the bench size (4M x 4M) with the final library: the roofline.traffic figure of bench.py.
# Usage: gpurun --timeout 500 -- 'bash scripts/gpu_traffic_r02.sh TAG'
TAG=${1:-r02w}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python scripts/prof_eval.py 3 f64 4000000 1 > $OUT/prof_4m_plain.log 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:eval_kernel -c 1 --csv \
    --log-file $OUT/traffic_4m.csv python scripts/prof_eval.py 3 f64 4000000 1 > $OUT/ncu_traffic.log 2>&1
echo "exit $?"; cat $OUT/prof_4m_plain.log; tail -4 $OUT/traffic_4m.csv
