#!/bin/bash
# One GPU-box pass: parity tests, the bench line, the ncu launch list of the bench command and one
# `--set full` capture of the north-star kernel. Usage (from the repo root): gpurun --timeout 1500 -- 'bash scripts/gpu_round.sh TAG'
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $OUT/smi.txt 2>&1
lscpu | head -20 > $OUT/lscpu.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
python scripts/quick_gpu.py > $OUT/quick.log 2>&1; cat $OUT/quick.log
python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; cat $OUT/bench_ref.json
python scripts/prof_eval.py 3 f64 262144 2 > $OUT/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -c 2 -o $OUT/prof_stokes_fxu_f64 \
    python scripts/prof_eval.py 3 f64 262144 2 > $OUT/ncu_full.log 2>&1
tail -2 $OUT/ncu_full.log
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $OUT/bench_short.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $OUT/ncu_launches.log 2>&1
tail -2 $OUT/ncu_launches.log
# DRAM traffic of ONE eval_kernel launch at the bench size (4M x 4M): a metrics-only pass (few replays of a 21 s kernel)
python scripts/prof_eval.py 3 f64 4000000 1 > $OUT/prof_4m_plain.log 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:eval_kernel -c 1 --csv \
    --log-file $OUT/traffic_4m.csv python scripts/prof_eval.py 3 f64 4000000 1 > $OUT/ncu_traffic.log 2>&1
tail -4 $OUT/traffic_4m.csv
# the one-launch small kernel and the FP32 kernel, for the record
python scripts/prof_eval.py 3 f64 2000 3 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:eval_small_kernel -c 1 -o $OUT/prof_small_stokes_fxu_f64_n2000 \
    python scripts/prof_eval.py 3 f64 2000 3 > $OUT/ncu_small.log 2>&1
python scripts/prof_eval.py 3 f32 262144 2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -c 1 -o $OUT/prof_stokes_fxu_f32 \
    python scripts/prof_eval.py 3 f32 262144 2 > $OUT/ncu_f32.log 2>&1
ls -la $OUT
