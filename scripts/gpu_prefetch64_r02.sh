#!/bin/bash
# FP64 prefetch (all f64 kernels, -DSCTL_B200_PREFETCH_F64=1) against the main library, with output checksums, then the parity and
# edge-case modules with the Python binding on the variant.
TAG=${1:-r02p3}
OUT=gpurun_out/$TAG
mkdir -p $OUT
C="0:f64:262144:10,0:f64:262144:-1,2:f64:262144:-1,1:f64:262144:-1,3:f64:262144:10,3:f64:262144:-1,4:f64:262144:-1"
python scripts/time_f32.py 3 $C > $OUT/time_main.txt 2>&1; cat $OUT/time_main.txt
V=$PWD/sctl_b200/lib/variants/libsctl_b200_prefetch64.so
SCTL_B200_LIB=$V python scripts/time_f32.py 3 $C > $OUT/time_prefetch64.txt 2>&1; cat $OUT/time_prefetch64.txt
SCTL_B200_LIB=$V timeout 60 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge_cases.py -m gpu -x -q > $OUT/pytest_prefetch64.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_prefetch64.log; tail -2 $OUT/pytest_prefetch64.log
