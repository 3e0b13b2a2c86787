#!/bin/bash
# Which NVML query perturbs the multi-GPU run? (gpurun --gpus 2)
OUT=gpurun_out/${1:-samplerdiag}; mkdir -p $OUT
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 --no-cpu-baseline 2>&1 | grep -E "rank [01]\]" | tr '\n' ' '; echo; }
export SCTL_B200_BENCH_VERBOSE=1 SCTL_B200_BENCH_N=1000000
for mode in clock power reasons none; do echo "== mode $mode period 0.5"; SCTL_B200_SAMPLER_MODE=$mode run; done
echo "== mode all period 3"; SCTL_B200_SAMPLER_PERIOD=3 run
echo "== 1 process on the 2-GPU box, mode all"; python bench.py --gpus 1 --steps 4 --warmup 3 --no-cpu-baseline 2>&1 | grep -E "rank 0\]"
echo "== 1 process, CUDA_VISIBLE_DEVICES=0"; CUDA_VISIBLE_DEVICES=0 python bench.py --gpus 1 --steps 4 --warmup 3 --no-cpu-baseline 2>&1 | grep -E "rank 0\]"
