#!/bin/bash
# One torchrun bench line at C3 on N GPUs (gpurun --gpus N -- 'bash scripts/gpu_bench_n_r02.sh TAG N').
TAG=${1:-r02x}; N=${2:-4}
OUT=gpurun_out/$TAG
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus $N --steps 2 --warmup 3 --cpu-seconds 4 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench exit $?"; cat $OUT/bench_n$N.json
