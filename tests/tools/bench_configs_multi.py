"""BASELINE.json configs at their quoted GPU count under torchrun (one process per GPU, NCCL): targets AND sources start
sharded on the ranks (the EvalDirect situation), ShardedDirectSum gathers the sources and evaluates the rank's targets.
Per config: 1 warm-up + K timed steps (CUDA events, barrier on both sides, max over ranks), whole-job pairs/s, and the
rel-L2 of a target sample on every rank against the CPU reference (max over ranks).
Usage: torchrun ... bench_configs_multi.py [C3 C4 C5] [--scale S]   (--scale divides N for a quick functional check)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import torch.distributed as dist
from oracle import loader
from sctl_b200 import KERNELS, capi, synth
from sctl_b200.distributed import ShardedDirectSum, partition

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
capi.check(capi.lib().sctl_b200_set_device(lr))
dist.init_process_group("nccl")
dev = torch.device("cuda", lr)
args = sys.argv[1:]
scale = 1
if "--scale" in args:
    i = args.index("--scale")
    scale = int(args[i + 1])
    del args[i:i + 2]
CONFIGS = [("C3", 3, 4_000_000, np.float64, 2), ("C4", 1, 2_000_000, np.float64, 3), ("C4", 4, 2_000_000, np.float64, 3), ("C5", 3, 16_000_000, np.float32, 1)]
chk = loader.best()
for name, kid, N, dt, steps in CONFIGS:
    if args and name not in args:
        continue
    N //= scale
    d = capi.kernel_dims(kid)
    t0, t1 = partition(N, rank, world)
    Xt, Xs, Xn, F = synth.make_problem(d, N, N, dt, t0=t0, t1=t1, s0=t0, s1=t1)
    T = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    dXt, dXs, dXn, dF = T(Xt), T(Xs), T(Xn), T(F)
    ker = KERNELS[kid]()
    drv = ShardedDirectSum(ker)
    counts = [partition(N, r, world)[1] - partition(N, r, world)[0] for r in range(world)]
    U = drv(None, dXt, dXs, dXn, dF, counts=counts)  # warm-up
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        U = drv(U, dXt, dXs, dXn, dF, counts=counts)
    e1.record(); torch.cuda.synchronize(); dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    # parity: a sample of this rank's targets against the CPU reference over ALL sources (regenerated here, f64 on the CPU side)
    rng = np.random.default_rng(100 + rank)
    idx = np.unique(np.concatenate([rng.integers(0, t1 - t0, 6), [0, t1 - t0 - 1]]))
    _, Xs_all, Xn_all, F_all = synth.make_problem(d, N, N, dt, t0=0, t1=1)
    f64 = lambda a: None if a is None else np.ascontiguousarray(a, np.float64)
    want = chk.eval(kid, f64(Xt.reshape(-1, 3)[idx].reshape(-1)), f64(Xs_all), f64(Xn_all), f64(F_all)).reshape(-1, d["trg_dim"])
    got = U.reshape(-1, d["trg_dim"])[torch.from_numpy(idx).to(dev)].cpu().numpy().astype(np.float64)
    err = torch.tensor([np.linalg.norm(got - want) / np.linalg.norm(want)], device=dev, dtype=torch.float64)
    dist.all_reduce(err, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"{name} {d['name']:15s} {np.dtype(dt).name} N=M={N} on {world} GPUs: {ms.item():10.3f} ms/step = {float(N) * N / ms.item() * 1e3:.3e} pairs/s whole job "
              f"({float(N) * N / ms.item() * 1e3 / world:.3e} per GPU) | rel-L2 vs {chk.kind} (8 targets per rank, max over ranks) {err.item():.2e}", flush=True)
    del dXt, dXs, dXn, dF, U, drv
    torch.cuda.empty_cache()
dist.barrier()
dist.destroy_process_group()
