"""Profiling target: a few launches of one Eval kernel (device-resident inputs). Usage: prof_eval.py KID DTYPE N [REPS] [DIGITS]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import KERNELS, capi, synth

kid, dt, n = int(sys.argv[1]), {"f64": np.float64, "f32": np.float32}[sys.argv[2]], int(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
digits = int(sys.argv[5]) if len(sys.argv) > 5 else -1
d = capi.kernel_dims(kid)
dev = torch.device("cuda:0")
Xt, Xs, Xn, F = synth.make_problem(d, n, n, dt)
t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
Xt, Xs, Xn, F = t(Xt), t(Xs), t(Xn), t(F)
ker = KERNELS[kid]()
U = ker.Eval(None, Xt, Xs, Xn, F, digits=digits)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    U = ker.Eval(U, Xt, Xs, Xn, F, digits=digits, overwrite=True)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"{d['name']} {np.dtype(dt).name} N={n} digits={digits} ms={ms:.3f} pairs/s={n * n / ms * 1e3:.4g}")
