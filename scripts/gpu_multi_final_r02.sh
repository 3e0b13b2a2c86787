#!/bin/bash
# Compact 8-GPU pass with the final library (gpurun --gpus 8 --timeout 330 -- 'bash scripts/gpu_multi_final_r02.sh TAG 8'):
# the C-ABI NCCL driver at 8 ranks (tests/cpp/sharded_test), SCTL_B200_ALL_GPUS over 1..8 GPUs, the bench line at C3.
TAG=${1:-r02u}; N=${2:-8}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt 2>&1
timeout 150 tests/cpp/sharded_test $N > $OUT/sharded_test_n$N.txt 2>&1; echo "sharded_test exit $?"; tail -6 $OUT/sharded_test_n$N.txt
timeout 150 python scripts/bench_all_gpus.py 1000000 3 3 > $OUT/all_gpus_n$N.txt 2>&1; echo "all_gpus exit $?"; cat $OUT/all_gpus_n$N.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT SCTL_B200_BENCH_VERBOSE=1 timeout 240 $TR bench.py --gpus $N --steps 2 --warmup 3 --cpu-seconds 4 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench C3 exit $?"; cat $OUT/bench_n$N.json
grep -E "NCCL INFO (ncclCommInitRank |comm 0x)" $OUT/bench_n$N.err | head -24 > $OUT/bench_n${N}_nccl_rank_lines.txt; wc -l $OUT/bench_n${N}_nccl_rank_lines.txt
