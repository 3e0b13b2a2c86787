"""Scratch GPU probe: peak DFMA/FFMA, Stokes FxU throughput at a few sizes. Not a bench (see bench.py)."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import capi, synth, KERNELS

print("peak f64 TF", capi.measure_peak(capi.F64, 1.0), "peak f32 TF", capi.measure_peak(capi.F32, 1.0))
dev = torch.device("cuda:0")
for kid, dt, N in [(3, np.float64, 262144), (3, np.float64, 1048576), (0, np.float64, 1048576), (3, np.float32, 1048576), (4, np.float64, 524288)]:
    d = capi.kernel_dims(kid)
    Xt, Xs, Xn, F = synth.make_problem(d, N, N, dt)
    t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    Xt, Xs, Xn, F = t(Xt), t(Xs), t(Xn), t(F)
    ker = KERNELS[kid]()
    U = ker.Eval(None, Xt, Xs, Xn, F)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); U = ker.Eval(U, Xt, Xs, Xn, F, overwrite=True); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(d["name"], dt.__name__, N, "ms", ms, "pairs/s %.4g" % (N * N / ms * 1e3), flush=True)
