#!/bin/bash
# ONE ncu --set full invocation over the f32 kernel and the digits-10 Laplace kernel of the final library.
TAG=${1:-r02y}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python scripts/prof_two.py > $OUT/prof_two_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -c 2 -o $OUT/prof_f32_and_digits10 \
    python scripts/prof_two.py > $OUT/ncu_two.log 2>&1
echo "exit $?"; cat $OUT/prof_two_plain.log; tail -n 3 $OUT/ncu_two.log; ls -la $OUT
