"""Diagnostic (torchrun): where does the multi-GPU step lose time? Modes: overlap / no-overlap / allgather-then-one-Eval / no-comm baseline."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from sctl_b200 import Stokes3D_FxU, capi, synth
from sctl_b200.distributed import ShardedDirectSum, partition

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
torch.cuda.set_device(lr)
capi.check(capi.lib().sctl_b200_set_device(lr))
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
ker = Stokes3D_FxU()
d = capi.kernel_dims(3)
t0, t1 = partition(n, rank, world)
counts = [partition(n, r, world)[1] - partition(n, r, world)[0] for r in range(world)]
Xt_np, Xs_np, _, F_np = synth.make_problem(d, n, n, np.float64, t0=t0, t1=t1, s0=t0, s1=t1)
Xt, Xs, F = (torch.from_numpy(a).to(dev) for a in (Xt_np, Xs_np, F_np))
U = torch.zeros(3 * (t1 - t0), dtype=torch.float64, device=dev)
_, Xs_all_np, _, F_all_np = synth.make_problem(d, 1, n, np.float64, coincident=0)
Xs_all, F_all = torch.from_numpy(Xs_all_np).to(dev), torch.from_numpy(F_all_np).to(dev)


def timed(name, fn, steps=int(os.environ.get("DIAG_STEPS", "3"))):
    fn(); torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    evs[0].record()
    for i in range(steps):
        fn(); evs[i + 1].record()
    torch.cuda.synchronize()
    ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
    print(f"[rank {rank}] {name:28s} per-step ms: " + " ".join(f"{m:9.1f}" for m in ms) + f"   ideal {n * (t1 - t0) / 7.65e11 * 1e3:9.1f}", flush=True)
    dist.barrier()


s_ov = ShardedDirectSum(ker, mode="ring")
s_no = ShardedDirectSum(ker, mode="gather")
gx = torch.empty(3 * n, dtype=torch.float64, device=dev)
gf = torch.empty(3 * n, dtype=torch.float64, device=dev)


def allgather_step():
    if len(set(counts)) == 1:
        dist.all_gather_into_tensor(gx, Xs); dist.all_gather_into_tensor(gf, F)
    else:
        dist.all_gather(list(gx.split([3 * c for c in counts])), Xs); dist.all_gather(list(gf.split([3 * c for c in counts])), F)
    ker.Eval(U, Xt, gx, None, gf, overwrite=True)


timed("no-comm baseline", lambda: ker.Eval(U, Xt, Xs_all, None, F_all, overwrite=True))
timed("allgather + one Eval", allgather_step)
timed("gather mode (default)", lambda: s_no(U, Xt, Xs, None, F, counts=counts))
timed("broadcast ring, overlap", lambda: s_ov(U, Xt, Xs, None, F, counts=counts))
timed("no-comm baseline (again)", lambda: ker.Eval(U, Xt, Xs_all, None, F_all, overwrite=True))
dist.destroy_process_group()
