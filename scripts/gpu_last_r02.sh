#!/bin/bash
# Last check of the round with the final main library: f32 timings of the kernels that read sources one iteration ahead,
# smoke, the parity module and the allocation-free entry point.
TAG=${1:-r02q3}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python scripts/time_f32.py 4 "0:f32:524288:-1,0:f32:524288:4,2:f32:524288:-1,2:f32:524288:4,6:f32:524288:-1,3:f32:524288:-1,3:f64:262144:-1" > $OUT/time_main_final.txt 2>&1; cat $OUT/time_main_final.txt
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -4 $OUT/smoke.log
timeout 60 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/pytest_parity.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_parity.log; tail -3 $OUT/pytest_parity.log
