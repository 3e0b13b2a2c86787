#!/bin/bash
# GPU-box pass 2: FP64-pipe microbenchmarks, compile-time variant sweep, the C++ shim test and the full-size tests.
TAG=${1:-r01b}
OUT=gpurun_out/$TAG
mkdir -p $OUT
./build/fp64_micro > $OUT/fp64_micro.log 2>&1; cat $OUT/fp64_micro.log
bash scripts/sweep_variants.sh 262144 -1 > $OUT/sweep_full.log 2>&1; cat $OUT/sweep_full.log
bash scripts/sweep_variants.sh 262144 11 > $OUT/sweep_d11.log 2>&1; cat $OUT/sweep_d11.log
python -m pytest tests/test_cpp_shim.py -m gpu -x -q -s > $OUT/pytest_shim.log 2>&1; tail -60 $OUT/pytest_shim.log
python -m pytest tests/test_gpu_full_size.py -m gpu -x -q -s --durations=10 > $OUT/pytest_full.log 2>&1; tail -25 $OUT/pytest_full.log
