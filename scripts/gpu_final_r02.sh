#!/bin/bash
# Final 1-GPU pass of round 2 with the library as committed: the allocation-free entry point's test alone in a fresh process,
# the whole GPU suite (that test last), smoke, both bench arms, and ONE ncu capture (--set full of the north-star kernel).
# Usage (from the repo root): gpurun --timeout 1500 -- 'bash scripts/gpu_final_r02.sh TAG'
TAG=${1:-r02t}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $OUT/smi.txt 2>&1
lscpu | head -20 > $OUT/lscpu.txt 2>&1
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -rA > $OUT/pytest_no_alloc_alone.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_no_alloc_alone.log
tail -4 $OUT/pytest_no_alloc_alone.log
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -5 $OUT/smoke.log
python bench.py --impl reference > $OUT/bench_ref.json 2> $OUT/bench_ref.err; cat $OUT/bench_ref.json
python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json
python scripts/prof_eval.py 3 f64 262144 2 > $OUT/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -c 1 -o $OUT/prof_stokes_fxu_f64 \
    python scripts/prof_eval.py 3 f64 262144 2 > $OUT/ncu_full.log 2>&1
tail -2 $OUT/prof_plain.log $OUT/ncu_full.log
ls -la $OUT
