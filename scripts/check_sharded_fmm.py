"""torchrun check (N GPUs, NCCL): ShardedParticleFMM with skewed target counts and ragged source blocks against one
single-GPU Eval of the same global problem on every rank. Prints one OK/FAIL line per rank."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from sctl_b200 import Stokes3D_DxU, Stokes3D_FxU, capi, synth
from sctl_b200.distributed import ShardedParticleFMM

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
capi.check(capi.lib().sctl_b200_set_device(int(os.environ["LOCAL_RANK"])))
dist.init_process_group("nccl")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 200_003
dt = np.float64
sl, dl = Stokes3D_FxU(), Stokes3D_DxU()
# skewed ownership: rank r owns a share proportional to r+1
w = np.arange(1, world + 1, dtype=np.float64)
edges = np.concatenate([[0], np.round(np.cumsum(w) / w.sum() * N).astype(np.int64)])
t0, t1 = int(edges[rank]), int(edges[rank + 1])
s0, s1 = N * rank // world, N * (rank + 1) // world
Xt, Xs, Xn, F = synth.make_problem(capi.kernel_dims(4), N, N, dt, coincident=64)
dev = torch.device("cuda", torch.cuda.current_device())
T = lambda a: torch.from_numpy(a).to(dev)
fmm = ShardedParticleFMM(dt)
fmm.SetKernels(sl, sl, sl)
fmm.AddTrg("U", sl, sl)
fmm.AddSrc("SL", sl, sl)
fmm.AddSrc("DL", dl, dl)
fmm.SetKernelS2T("SL", "U", sl)
fmm.SetKernelS2T("DL", "U", dl)
fmm.SetTrgCoord("U", T(Xt[3 * t0:3 * t1]))
fmm.SetSrcCoord("SL", T(Xs[3 * s0:3 * s1]))
fmm.SetSrcCoord("DL", T(Xs[3 * s0:3 * s1]), T(Xn[3 * s0:3 * s1]))
fmm.SetSrcDensity("SL", T(F[3 * s0:3 * s1]))
fmm.SetSrcDensity("DL", T(F[3 * s0:3 * s1]))
ok = True
for it in range(3):  # density updates, as in a solver loop
    fmm.SetSrcDensity("DL", T((it + 1.0) * F[3 * s0:3 * s1]))
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    U = fmm.Eval("U")
    e1.record(); torch.cuda.synchronize()
    want = sl.Eval(None, T(Xt[3 * t0:3 * t1]), T(Xs), None, T(F))
    want = dl.Eval(want, T(Xt[3 * t0:3 * t1]), T(Xs), T(Xn), T((it + 1.0) * F))
    err = float(torch.linalg.norm(U - want) / torch.linalg.norm(want))
    ok = ok and err < 1e-12 and U.numel() == 3 * (t1 - t0)
    print(f"rank {rank} it {it}: own targets {t1 - t0} of {N} (skewed: re-partitioned inside libsctl_b200.so), eval {e0.elapsed_time(e1):.2f} ms, rel-L2 vs single-GPU {err:.2e}", flush=True)
print(f"rank {rank}: {'OK' if ok else 'FAIL'}", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
