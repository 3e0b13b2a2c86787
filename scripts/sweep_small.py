"""Small-kernel tuning: device time per call by targets-per-CTA (SCTL_B200_SMALL_TPW) vs the tiled path. Usage: sweep_small.py KID DTYPE"""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import capi, synth

kid, dt = int(sys.argv[1]), {"f64": np.float64, "f32": np.float32}[sys.argv[2]]
d = capi.kernel_dims(kid)
L = capi.lib()
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
for nt, ns in [(500, 500), (1000, 1000), (1500, 1500), (2000, 2000), (3000, 3000), (4000, 4000), (5000, 5000), (6000, 6000), (7000, 7000), (8000, 8000), (300, 8000), (20000, 1000), (100000, 250)]:
    Xt, Xs, Xn, F = synth.make_problem(d, nt, ns, dt)
    t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    Xt, Xs, Xn, F = t(Xt), t(Xs), t(Xn), t(F)
    U = torch.zeros(nt * d["trg_dim"], dtype=Xt.dtype, device=dev)
    p = lambda a: None if a is None else ctypes.c_void_p(a.data_ptr())
    args = (kid, capi.F64 if dt == np.float64 else capi.F32, nt, ns, p(Xt), p(Xs), p(Xn) if d["normal_dim"] else None, p(F), p(U), -1, capi.OVERWRITE, ctypes.c_void_p(st))
    res = []
    for label, env in [("tiled", {"SCTL_B200_SMALL_MAX_PAIRS": "0"}), ("tpw8", {"SCTL_B200_SMALL_TPW": "8"}), ("tpw16", {"SCTL_B200_SMALL_TPW": "16"}), ("tpw32", {"SCTL_B200_SMALL_TPW": "32"})]:
        for k in ("SCTL_B200_SMALL_MAX_PAIRS", "SCTL_B200_SMALL_TPW"):
            os.environ.pop(k, None)
        if label != "tiled":
            os.environ["SCTL_B200_SMALL_MAX_PAIRS"] = str(2**31 - 1)
        os.environ.update(env)
        for _ in range(5):
            capi.check(L.sctl_b200_eval_async(*args))
        torch.cuda.synchronize()
        reps = 100
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            L.sctl_b200_eval_async(*args)
        e1.record()
        torch.cuda.synchronize()
        res.append(f"{label} {1e3 * e0.elapsed_time(e1) / reps:7.1f} us")
    print(f"{d['name']} {np.dtype(dt).name} Nt={nt:6d} Ns={ns:6d}: " + " | ".join(res), flush=True)
