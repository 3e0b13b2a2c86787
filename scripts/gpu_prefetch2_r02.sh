#!/bin/bash
# Second look at the f32 prefetch: the kernels / digits levels the first sweep did not time.
TAG=${1:-r02p2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
C="2:f32:524288:4,1:f32:524288:-1,0:f32:524288:-1,0:f32:524288:4,6:f32:524288:-1,7:f32:524288:-1"
python scripts/time_f32.py 4 $C > $OUT/time_main.txt 2>&1; cat $OUT/time_main.txt
SCTL_B200_LIB=$PWD/sctl_b200/lib/variants/libsctl_b200_prefetch.so python scripts/time_f32.py 4 $C > $OUT/time_prefetch.txt 2>&1; cat $OUT/time_prefetch.txt
