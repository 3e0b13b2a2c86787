"""Times eval kernels (default: the f32 ones) of the library SCTL_B200_LIB names (device-resident inputs, CUDA events, best of REPS)."""
import hashlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import KERNELS, capi, synth

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5  # usage: time_f32.py [REPS] [kid:f32|f64:n:digits,...]
dev = torch.device("cuda:0")
print("library:", capi.LIB_PATH)
CASES = ((3, np.float32, 524288, -1), (0, np.float32, 524288, -1), (2, np.float32, 524288, -1), (4, np.float32, 524288, -1),
         (3, np.float32, 524288, 4), (3, np.float64, 262144, -1))
if len(sys.argv) > 2:  # kid:f32|f64:n:digits,...
    DT = {"f32": np.float32, "f64": np.float64}
    CASES = tuple((int(a), DT[b], int(c), int(d)) for a, b, c, d in (x.split(":") for x in sys.argv[2].split(",")))
for kid, dt, n, digits in CASES:
    d = capi.kernel_dims(kid)
    Xt, Xs, Xn, F = synth.make_problem(d, n, n, dt)
    t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    Xt, Xs, Xn, F = t(Xt), t(Xs), t(Xn), t(F)
    ker = KERNELS[kid]()
    U = ker.Eval(None, Xt, Xs, Xn, F, digits=digits)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        U = ker.Eval(U, Xt, Xs, Xn, F, digits=digits, overwrite=True)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    md5 = hashlib.md5(U.cpu().numpy().tobytes()).hexdigest()[:12]  # same bits from two builds = same arithmetic
    print(f"{d['name']:15s} {np.dtype(dt).name} N={n} digits={digits:2d} best ms={best:.3f} pairs/s={n * n / best * 1e3:.4g} md5(U)={md5}", flush=True)
