#!/bin/bash
# GPU-box pass: parity tests on the main library, then the compile-time variant sweep. Usage: gpu_sweep.sh TAG [N]
TAG=${1:-sweep}; N=${2:-262144}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $OUT/pytest_parity.log 2>&1; tail -5 $OUT/pytest_parity.log
bash scripts/sweep_variants.sh $N -1 > $OUT/sweep_full.log 2>&1; cat $OUT/sweep_full.log
bash scripts/sweep_variants.sh $((N*2)) -1 f32 > $OUT/sweep_f32.log 2>&1; cat $OUT/sweep_f32.log
python scripts/quick_gpu.py > $OUT/quick.log 2>&1; cat $OUT/quick.log
