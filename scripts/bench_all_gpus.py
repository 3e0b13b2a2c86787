"""SCTL_B200_ALL_GPUS throughput: the only multi-GPU route a C++ host-pointer caller has without bringing its own ranks.
One process, host (pageable numpy) arrays, sctl_b200_eval(..., SCTL_B200_ALL_GPUS) over 1, 2, 4, ... visible GPUs
(sctl_b200_init selects the device set), in both source-distribution modes:
  nccl      : every GPU uploads 1/G of the sources, NCCL (ncclCommInitAll communicators inside the library) replicates them
  replicate : every GPU uploads all sources itself (no collective; the round-1 form)
Usage: python scripts/bench_all_gpus.py [N=1000000] [kernel_id=3] [reps=3]"""
import ctypes
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sctl_b200 import KERNELS, capi, synth  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
kid = int(sys.argv[2]) if len(sys.argv) > 2 else 3
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
d = capi.kernel_dims(kid)
ndev = capi.device_count()
Xt, Xs, Xn, F = synth.make_problem(d, N, N, np.float64)
ker = KERNELS[kid]()
L = capi.lib()
print(f"{d['name']} f64 {N} x {N}, host pointers (pageable), {ndev} GPU(s) visible")
base = None
ref = None
g = 1
while g <= ndev:
    ids = (ctypes.c_int * g)(*range(g))
    for mode in (("nccl", "replicate") if g > 1 else ("single",)):
        os.environ["SCTL_B200_ALL_GPUS_MODE"] = "replicate" if mode == "replicate" else "nccl"
        capi.check(L.sctl_b200_init(g, ids))
        U = ker.Eval(None, Xt, Xs, Xn, F, overwrite=True, all_gpus=True)  # warm-up: contexts, communicators, pools
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            U = ker.Eval(U, Xt, Xs, Xn, F, overwrite=True, all_gpus=True)
            ts.append(time.perf_counter() - t0)
        t = min(ts)
        if ref is None:
            ref, base = U.copy(), t
        err = float(np.linalg.norm(U - ref) / np.linalg.norm(ref))
        print(f"  {g} GPU(s) {mode:9s}: {t * 1e3:9.2f} ms/call (best of {reps}; all: {[round(x * 1e3, 1) for x in ts]})  {N * N / t:.4g} pairs/s  "
              f"speed-up {base / t:5.2f}x  rel-L2 vs 1 GPU {err:.1e}", flush=True)
    g *= 2
capi.check(L.sctl_b200_finalize())
