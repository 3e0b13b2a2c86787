#!/bin/bash
# Runs scripts/prof_eval.py for every library variant under sctl_b200/lib/variants (GPU box). Usage: sweep_variants.sh [N] [digits] [f64|f32] [kernel_id]
N=${1:-262144}; DIG=${2:--1}; DT=${3:-f64}; KID=${4:-3}
for so in sctl_b200/lib/variants/libsctl_b200_*.so; do
  v=$(basename $so .so); v=${v#libsctl_b200_}
  printf "%-10s " $v
  SCTL_B200_LIB=$PWD/$so python scripts/prof_eval.py $KID $DT $N 3 $DIG 2>&1 | tail -1
done
