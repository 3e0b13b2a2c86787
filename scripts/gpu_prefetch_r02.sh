#!/bin/bash
# FP32 prefetch variant (sctl_b200/lib/variants/libsctl_b200_prefetch.so, -DSCTL_B200_PREFETCH_F32=1) against the main library,
# then the whole GPU suite with the Python binding pointed at the variant.
TAG=${1:-r02p}
OUT=gpurun_out/$TAG
mkdir -p $OUT
V=$PWD/sctl_b200/lib/variants/libsctl_b200_prefetch.so
python scripts/time_f32.py 5 > $OUT/time_main.txt 2>&1; cat $OUT/time_main.txt
SCTL_B200_LIB=$V python scripts/time_f32.py 5 > $OUT/time_prefetch.txt 2>&1; cat $OUT/time_prefetch.txt
SCTL_B200_LIB=$V timeout 200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_prefetch.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu_prefetch.log
tail -3 $OUT/pytest_gpu_prefetch.log
