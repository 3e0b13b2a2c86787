#!/bin/bash
# Last check of the round with the final main library: f32 timings, smoke, the parity module.
TAG=${1:-r02q2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python scripts/time_f32.py 5 > $OUT/time_main_final.txt 2>&1; cat $OUT/time_main_final.txt
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -4 $OUT/smoke.log
timeout 100 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/pytest_parity.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_parity.log; tail -3 $OUT/pytest_parity.log
