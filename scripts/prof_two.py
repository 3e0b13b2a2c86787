"""Profiling target for ONE ncu invocation over two instantiations: Stokes3D_FxU f32 (N = 524288, full precision) and
Laplace3D_FxU f64 at digits = 10 (N = 262144). One launch of each eval_kernel (ncu replays it); device-resident inputs."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import KERNELS, capi, synth

dev = torch.device("cuda:0")
for kid, dt, n, digits in ((3, np.float32, 524288, -1), (0, np.float64, 262144, 10)):
    d = capi.kernel_dims(kid)
    Xt, Xs, Xn, F = synth.make_problem(d, n, n, dt)
    t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    Xt, Xs, Xn, F = t(Xt), t(Xs), t(Xn), t(F)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    U = KERNELS[kid]().Eval(None, Xt, Xs, Xn, F, digits=digits)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"{d['name']} {np.dtype(dt).name} N={n} digits={digits} ms={ms:.3f} (first call, incl. pack/reduce) pairs/s={n * n / ms * 1e3:.4g}")
