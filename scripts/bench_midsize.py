"""Mid-size grids: pairs/s of the tiled grid (SCTL_B200_WQ=0) against the persistent even-split kernel (SCTL_B200_WQ=1) and
the planner's own choice (unset), device-resident inputs, CUDA events over `reps` back-to-back calls (pack + eval [+ reduce]).
Usage: python scripts/bench_midsize.py [reps=20] [modes=0,1,auto]   (mode "1<k>" = queue with k items per CTA and level)"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from sctl_b200 import KERNELS, capi, synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
SELF = bool(os.environ.get("MIDSIZE_SELF"))
SIZES = [int(x) for x in os.environ.get("MIDSIZE_N", "8192,12000,20000,30000,45000,60000,100000,200000,262144").split(",")]
MODES = [m if m != "auto" else None for m in (sys.argv[2].split(",") if len(sys.argv) > 2 else ["0", "1", "auto"])]
dev = torch.device("cuda:0")
peak = capi.measure_peak(capi.F64, 1.0)
print(f"measured DFMA-chain peak {peak:.2f} TFLOP/s; {'self-interaction (targets == sources)' if SELF else 'independent targets, first 1024 coincide with sources'}")
INSTR = {0: 12, 3: 22, 4: 25}  # FP64-pipe instructions per pair
for kid in (0, 3, 4):
    d = capi.kernel_dims(kid)
    for N in SIZES:
        Xt, Xs, Xn, F = synth.make_problem(d, N, N, np.float64)
        if SELF:  # self-interaction (N-body): targets ARE the sources, every target tile repairs the tiles on its diagonal
            Xt = Xs.copy()
        t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
        Xt, Xs, Xn, F = t(Xt), t(Xs), t(Xn), t(F)
        ker = KERNELS[kid]()
        row, ref = [], None
        for mode in MODES:
            os.environ.pop("SCTL_B200_WQ_ITEMS_PER_LEVEL", None)
            if mode is None:
                os.environ.pop("SCTL_B200_WQ", None)
            else:
                os.environ["SCTL_B200_WQ"] = mode[0]
                if len(mode) > 1:
                    os.environ["SCTL_B200_WQ_ITEMS_PER_LEVEL"] = mode[1:]
            U = ker.Eval(None, Xt, Xs, Xn, F)
            for _ in range(3):
                ker.Eval(U, Xt, Xs, Xn, F, overwrite=True)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                ker.Eval(U, Xt, Xs, Xn, F, overwrite=True)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            if ref is None:
                ref = U.clone()
            err = float(torch.linalg.norm(U - ref) / torch.linalg.norm(ref))
            rate = N * N / ms * 1e3
            row.append(f"{'grid' if mode == '0' else 'auto' if mode is None else 'queue/' + (mode[1:] or '2')} {ms:8.4f} ms {rate:.3e} p/s pipe {rate * INSTR[kid] * 2 / peak * 1e-12 * 100:5.1f}% (dev {err:.0e})")
        print(f"{d['name']:14s} N={N:7d}  " + " | ".join(row), flush=True)
os.environ.pop("SCTL_B200_WQ", None)
