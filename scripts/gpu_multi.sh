#!/bin/bash
# Multi-GPU pass (gpurun --gpus N): new tests, then bench.py under torchrun at N GPUs (small problem first, then config C3).
TAG=${1:-multi}; N=${2:-2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt 2>&1; cat $OUT/gpus.txt
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
SCTL_B200_BENCH_VERBOSE=1 SCTL_B200_BENCH_N=1000000 run --steps 3 --warmup 3 > $OUT/bench_n${N}_1M.json 2> $OUT/bench_n${N}_1M.err; echo "exit $?"; grep -E "rank|Error|error" $OUT/bench_n${N}_1M.err | tail; cat $OUT/bench_n${N}_1M.json
if [ -n "$FULL" ]; then SCTL_B200_BENCH_VERBOSE=1 run --steps 2 --warmup 3 > $OUT/bench_n${N}.json 2> $OUT/bench_n${N}.err; echo "exit $?"; grep -E "rank|Error|error" $OUT/bench_n${N}.err | tail; cat $OUT/bench_n${N}.json; fi
