"""All BASELINE.json configs on ONE GPU (C4/C5 are 8-GPU configs: one rank's target shard is timed): kernel-only and
end-to-end (host buffers through the C ABI) pairs/s, FP pipe fraction, rel-L2 on a target subset vs the CPU reference."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from oracle import loader
from sctl_b200 import KERNELS, capi, synth

chk = loader.best()
dev = torch.device("cuda:0")
peak = {np.float64: capi.measure_peak(capi.F64, 1.0), np.float32: capi.measure_peak(capi.F32, 1.0)}
INSTR = {0: 12, 1: 17, 2: 17, 3: 22, 4: 25}  # executed FP instructions per pair (f64; f32 has one fewer: 4-op rsqrt)
print(f"peak DFMA {peak[np.float64]:.2f} TF, FFMA {peak[np.float32]:.2f} TF; checker {chk.kind}")
CONFIGS = [("C1", 0, 20_000, 20_000, np.float64, 1), ("C2", 0, 1_000_000, 1_000_000, np.float64, 1), ("C2", 2, 1_000_000, 1_000_000, np.float64, 1),
           ("C3", 3, 4_000_000, 4_000_000, np.float64, 1), ("C4", 1, 2_000_000, 2_000_000, np.float64, 8), ("C4", 4, 2_000_000, 2_000_000, np.float64, 8),
           ("C5", 3, 16_000_000, 16_000_000, np.float32, 8)]
only = sys.argv[1:]  # e.g. C1 C2
for name, kid, Ns, Nt_total, dt, ranks in CONFIGS:
    if only and name not in only:
        continue
    d = capi.kernel_dims(kid)
    Nt = Nt_total // ranks
    Xt, Xs, Xn, F = synth.make_problem(d, Nt_total, Ns, dt, t0=0, t1=Nt)
    t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    dXt, dXs, dXn, dF = t(Xt), t(Xs), t(Xn), t(F)
    ker = KERNELS[kid]()
    U = ker.Eval(None, dXt, dXs, dXn, dF)
    torch.cuda.synchronize()
    reps = 20 if Nt * Ns < 1e10 else (3 if Nt * Ns < 3e12 else 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        ker.Eval(U, dXt, dXs, dXn, dF, overwrite=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # end to end: host numpy buffers through sctl_b200_eval
    Uh = np.zeros(Nt * d["trg_dim"], dt)
    ker.Eval(Uh, Xt, Xs, Xn, F, overwrite=True)
    w0 = time.perf_counter()
    for _ in range(reps):
        ker.Eval(Uh, Xt, Xs, Xn, F, overwrite=True)
    e2e_ms = (time.perf_counter() - w0) * 1e3 / reps
    rng = np.random.default_rng(1)
    idx = np.unique(np.concatenate([rng.integers(0, Nt, 60), [0, 1, Nt - 1]]))
    f64 = lambda a: None if a is None else np.ascontiguousarray(a, np.float64)
    c0 = time.perf_counter()
    want = chk.eval(kid, f64(Xt.reshape(-1, 3)[idx].reshape(-1)), f64(Xs), f64(Xn), f64(F)).reshape(-1, d["trg_dim"])
    cpu_rate = idx.size * Ns / (time.perf_counter() - c0)
    got = Uh.reshape(-1, d["trg_dim"])[idx].astype(np.float64)
    err = np.linalg.norm(got - want) / np.linalg.norm(want)
    instr = INSTR[kid] - (1 if dt == np.float32 else 0)
    pipe = Nt * Ns * instr * 2 / (ms * 1e-3) * 1e-12 / peak[dt]
    print(f"{name} {d['name']:15s} {np.dtype(dt).name} Ns={Ns} Nt={Nt}{' (1/%d shard)' % ranks if ranks > 1 else ''}: kernel {ms:10.3f} ms = {Nt * Ns / ms * 1e3:.3e} pairs/s "
          f"(FP pipe {pipe * 100:.1f}% of measured peak, flops_ref frac {Nt * Ns * d['flops'] / (ms * 1e-3) * 1e-12 / peak[dt] * 100:.1f}%) | e2e host buffers {e2e_ms:10.3f} ms = "
          f"{Nt * Ns / e2e_ms * 1e3:.3e} pairs/s | rel-L2 vs {chk.kind} on {idx.size} targets {err:.2e} | CPU {chk.kind} ({idx.size} targets) {cpu_rate:.2e} pairs/s", flush=True)
    del dXt, dXs, dXn, dF, U
