#!/bin/bash
# One GPU-box pass for round 2: parity tests, the bench line (ours + reference arm), the ncu launch list of the bench command,
# one `--set full` capture of the north-star kernel, the DRAM traffic of one 4M x 4M launch, and the f32 / digits-10 kernels.
# Usage (from the repo root): gpurun --timeout 2400 -- 'bash scripts/gpu_round_r02.sh TAG'
TAG=${1:-r02}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $OUT/smi.txt 2>&1
lscpu | head -20 > $OUT/lscpu.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -5 $OUT/smoke.log
python bench.py --impl reference > $OUT/bench_ref.json 2> $OUT/bench_ref.err; cat $OUT/bench_ref.json
python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json
python scripts/prof_eval.py 3 f64 262144 2 > $OUT/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -c 2 -o $OUT/prof_stokes_fxu_f64 \
    python scripts/prof_eval.py 3 f64 262144 2 > $OUT/ncu_full.log 2>&1
tail -2 $OUT/ncu_full.log
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $OUT/bench_short.json 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $OUT/ncu_launches.log 2>&1
tail -2 $OUT/ncu_launches.log
python scripts/prof_eval.py 3 f64 4000000 1 > $OUT/prof_4m_plain.log 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:eval_kernel -c 1 --csv \
    --log-file $OUT/traffic_4m.csv python scripts/prof_eval.py 3 f64 4000000 1 > $OUT/ncu_traffic.log 2>&1
tail -4 $OUT/traffic_4m.csv
python scripts/prof_eval.py 3 f32 524288 2 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -c 1 -o $OUT/prof_stokes_fxu_f32 \
    python scripts/prof_eval.py 3 f32 524288 2 > $OUT/ncu_f32.log 2>&1
python scripts/prof_eval.py 0 f64 262144 2 10 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -c 1 -o $OUT/prof_laplace_fxu_f64_digits10 \
    python scripts/prof_eval.py 0 f64 262144 2 10 > $OUT/ncu_d10.log 2>&1
python scripts/bench_kmat.py 8192 > $OUT/kmat_8192.txt 2>&1; python scripts/bench_kmat.py 16384 > $OUT/kmat_16384.txt 2>&1; cat $OUT/kmat_8192.txt $OUT/kmat_16384.txt
python tests/tools/bench_configs.py > $OUT/configs_1gpu.txt 2>&1; tail -12 $OUT/configs_1gpu.txt
ls -la $OUT
