"""KernelMatrix write bandwidth against the sources-per-CTA knob (SCTL_B200_KMAT_SB = 32 / 16 / 8 / 4, unset = the launcher's
choice): thin rows (one value per target) at small N are a 40-90 us kernel, where one ragged wave of CTAs costs 15 %. Also the
batched small-block form at 256 / 1024 / 4096 blocks of 96 far nodes x 300 near targets (the SetupNear shape).
Usage: python scripts/bench_kmat_sb.py"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import KERNELS, capi, synth

peaks = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
hbm = json.load(open(peaks))["hbm_gbs"] if os.path.exists(peaks) else 6650.0
dev = torch.device("cuda:0")
t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
MODES = ["32", "16", "8", "4", None]


def timed(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


print(f"measured copy bandwidth {hbm:.0f} GB/s (read+write); entries are GB/s of M written (% of that)")
for kid, dt in [(0, np.float32), (0, np.float64), (1, np.float32), (3, np.float64), (3, np.float32)]:
    d = capi.kernel_dims(kid)
    for N in (4096, 8192, 16384):
        n = N if d["src_dim"] * d["trg_dim"] == 1 else N // 2
        Xt, Xs, Xn, _ = synth.make_problem(d, n, n, dt)
        Xt, Xs, Xn = t(Xt), t(Xs), t(Xn)
        ker = KERNELS[kid]()
        M = ker.KernelMatrix(Xt, Xs, Xn)
        gb = M.numel() * M.element_size() / 1e9
        row = []
        for m in MODES:
            if m is None:
                os.environ.pop("SCTL_B200_KMAT_SB", None)
            else:
                os.environ["SCTL_B200_KMAT_SB"] = m
            ms = timed(lambda: ker.KernelMatrix(Xt, Xs, Xn, out=M))
            row.append(f"sb={m or 'auto':>4s} {gb / ms * 1e3:6.0f} ({gb / ms * 1e3 / hbm * 100:5.1f}%)")
        print(f"{d['name']:14s} {np.dtype(dt).name} N={n:6d} M={gb:5.2f} GB  " + " | ".join(row), flush=True)
os.environ.pop("SCTL_B200_KMAT_SB", None)
kid, dt = 4, np.float64
d = capi.kernel_dims(kid)
for nb in (64, 256, 1024, 4096):
    nt, ns = 300, 96
    trg_dsp = np.arange(nb + 1, dtype=np.int64) * nt
    src_dsp = np.arange(nb + 1, dtype=np.int64) * ns
    Xt, Xs, Xn, _ = synth.make_problem(d, nb * nt, nb * ns, dt)
    Xt, Xs, Xn = t(Xt), t(Xs), t(Xn)
    W = t(synth.uniform(5, 0, nb * ns, dtype=dt) + 1.0)
    ker = KERNELS[kid]()
    M, M_dsp = ker.KernelMatrixBatch(trg_dsp, src_dsp, Xt, Xs, Xn, src_wts=W)
    gb = M.numel() * M.element_size() / 1e9
    row = []
    for m in MODES:
        if m is None:
            os.environ.pop("SCTL_B200_KMAT_SB", None)
        else:
            os.environ["SCTL_B200_KMAT_SB"] = m
        ms = timed(lambda: ker.KernelMatrixBatch(trg_dsp, src_dsp, Xt, Xs, Xn, src_wts=W, out=M, M_dsp=M_dsp))
        row.append(f"sb={m or 'auto':>4s} {ms * 1e3:7.1f} us {gb / ms * 1e3:6.0f} GB/s")
    print(f"batch {nb:5d} blocks of {ns} x {nt} Stokes DxU f64, M={gb:5.2f} GB (wall per call incl. description upload + sync)  " + " | ".join(row), flush=True)
os.environ.pop("SCTL_B200_KMAT_SB", None)
