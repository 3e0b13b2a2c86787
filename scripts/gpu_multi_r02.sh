#!/bin/bash
# Round-2 multi-GPU pass (gpurun --gpus N -- 'bash scripts/gpu_multi_r02.sh TAG N'): the NCCL driver inside libsctl_b200.so.
TAG=${1:-r02multi}; N=${2:-2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi -L > $OUT/gpus.txt 2>&1; cat $OUT/gpus.txt
timeout 600 tests/cpp/sharded_test $N > $OUT/sharded_test_n$N.txt 2>&1; echo "sharded_test exit $?"; tail -25 $OUT/sharded_test_n$N.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR scripts/check_sharded_fmm.py 200003 > $OUT/sharded_fmm_n$N.txt 2>&1; echo "check_sharded_fmm exit $?"; grep -E "rank|Error" $OUT/sharded_fmm_n$N.txt | tail -12
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT SCTL_B200_BENCH_VERBOSE=1 SCTL_B200_BENCH_N=1000000 timeout 900 $TR bench.py --gpus $N --steps 3 --warmup 3 --cpu-seconds 4 > $OUT/bench_n${N}_1M.json 2> $OUT/bench_n${N}_1M.err; echo "bench exit $?"
grep -E "NCCL INFO (comm|ncclCommInitRank|Init COMPLETE|NVLS)|per-step|Error|error" $OUT/bench_n${N}_1M.err | head -30; cat $OUT/bench_n${N}_1M.json
SCTL_B200_BENCH_N=1000000 timeout 600 $TR bench.py --impl reference --gpus $N --steps 1 --warmup 1 --cpu-seconds 4 > $OUT/bench_ref_n${N}_1M.json 2> $OUT/bench_ref_n${N}_1M.err; cat $OUT/bench_ref_n${N}_1M.json
timeout 900 python scripts/bench_all_gpus.py ${ALLGPU_N:-1000000} 3 3 > $OUT/all_gpus_n$N.txt 2>&1; echo "all_gpus exit $?"; cat $OUT/all_gpus_n$N.txt
if [ -n "$FULL" ]; then
  SCTL_B200_BENCH_VERBOSE=1 timeout 1500 $TR bench.py --gpus $N --steps 2 --warmup 3 > $OUT/bench_n${N}.json 2> $OUT/bench_n${N}.err; echo "bench C3 exit $?"; cat $OUT/bench_n${N}.json
  timeout 1500 python scripts/bench_all_gpus.py 4000000 3 1 > $OUT/all_gpus_c3_n$N.txt 2>&1; cat $OUT/all_gpus_c3_n$N.txt
  timeout 1500 $TR tests/tools/bench_configs_multi.py C4 C5 > $OUT/configs_n$N.txt 2>&1; echo "configs exit $?"; grep -vE "^\[|^W|^\*" $OUT/configs_n$N.txt | tail -12
fi
