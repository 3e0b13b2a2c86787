"""KernelMatrix throughput (HBM-write-bound): achieved GB/s of M written vs MEASURED_PEAKS.json hbm_gbs. Usage: bench_kmat.py [N]"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import KERNELS, capi, synth

N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
peaks = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
hbm = json.load(open(peaks))["hbm_gbs"] if os.path.exists(peaks) else 6650.0
dev = torch.device("cuda:0")
for kid, dt in [(0, np.float64), (3, np.float64), (4, np.float64), (5, np.float64), (3, np.float32), (0, np.float32)]:
    d = capi.kernel_dims(kid)
    n = N if d["src_dim"] * d["trg_dim"] <= 9 else N // 2
    Xt, Xs, Xn, _ = synth.make_problem(d, n, n, dt)
    t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    Xt, Xs, Xn = t(Xt), t(Xs), t(Xn)
    ker = KERNELS[kid]()
    M = ker.KernelMatrix(Xt, Xs, Xn)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        ker.KernelMatrix(Xt, Xs, Xn, out=M)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    gb = M.numel() * M.element_size() / 1e9
    print(f"{d['name']:15s} {np.dtype(dt).name} N={n} M={gb:.2f} GB  ms={ms:.3f}  write {gb / ms * 1e3:.0f} GB/s = {gb / ms * 1e3 / hbm * 100:.1f}% of measured copy bandwidth ({hbm:.0f} GB/s counts read+write)  pairs/s={n * n / ms * 1e3:.3g}", flush=True)
