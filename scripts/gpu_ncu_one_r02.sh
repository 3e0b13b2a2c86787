#!/bin/bash
# ONE ncu --set full capture of one eval_kernel instantiation: gpu_ncu_one_r02.sh TAG NAME KID DTYPE N DIGITS
TAG=${1:-r02z}; NAME=${2:-laplace_fxu_f64_digits10}; KID=${3:-0}; DT=${4:-f64}; N=${5:-262144}; DIG=${6:-10}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python scripts/prof_eval.py $KID $DT $N 2 $DIG > $OUT/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -c 1 -o $OUT/prof_$NAME \
    python scripts/prof_eval.py $KID $DT $N 2 $DIG > $OUT/ncu.log 2>&1
echo "exit $?"; cat $OUT/prof_plain.log; tail -n 3 $OUT/ncu.log
