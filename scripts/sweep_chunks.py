"""Sweep SCTL_B200_FORCE_CHUNKS for one kernel/size (device-resident inputs) to calibrate make_plan's choice.
Usage: sweep_chunks.py KID DTYPE NS NT [REPS]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import KERNELS, capi, synth

kid, dt, ns, nt = int(sys.argv[1]), {"f64": np.float64, "f32": np.float32}[sys.argv[2]], int(sys.argv[3]), int(sys.argv[4])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 20
d = capi.kernel_dims(kid)
dev = torch.device("cuda:0")
Xt, Xs, Xn, F = synth.make_problem(d, ns, nt, dt)
t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
Xt, Xs, Xn, F = t(Xt), t(Xs), t(Xn), t(F)
ker = KERNELS[kid]()
n_tiles = (ns + 255) // 256


def run():
    U = ker.Eval(None, Xt, Xs, Xn, F)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        U = ker.Eval(U, Xt, Xs, Xn, F, overwrite=True)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


os.environ.pop("SCTL_B200_FORCE_CHUNKS", None)
print(f"{d['name']} {np.dtype(dt).name} Ns={ns} Nt={nt} n_tiles={n_tiles}: planner ms={run():.4f}")
seen = set()
for tpc in list(range(1, 17)) + [20, 24, 32, 48, 64, 96, 128]:
    if tpc > n_tiles:
        break
    c = (n_tiles + tpc - 1) // tpc
    if c in seen:
        continue
    seen.add(c)
    os.environ["SCTL_B200_FORCE_CHUNKS"] = str(c)
    print(f"  tpc={tpc:4d} chunks={c:5d} ms={run():.4f}")
