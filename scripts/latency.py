"""Per-call cost of the C ABI at small sizes: host-side enqueue time and device time per sctl_b200_eval_async call
(device-resident inputs, torch only used to allocate). Usage: latency.py [KID] [DTYPE]"""
import ctypes, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import capi, synth

kid = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dt = {"f64": np.float64, "f32": np.float32}[sys.argv[2] if len(sys.argv) > 2 else "f64"]
d = capi.kernel_dims(kid)
L = capi.lib()
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
for n in (64, 256, 1000, 2000, 4000, 8000, 16000):
    Xt, Xs, Xn, F = synth.make_problem(d, n, n, dt)
    t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    Xt, Xs, Xn, F = t(Xt), t(Xs), t(Xn), t(F)
    U = torch.zeros(n * d["trg_dim"], dtype=Xt.dtype, device=dev)
    p = lambda a: None if a is None else ctypes.c_void_p(a.data_ptr())
    args = (kid, capi.F64 if dt == np.float64 else capi.F32, n, n, p(Xt), p(Xs), p(Xn) if d["normal_dim"] else None, p(F), p(U), -1, capi.OVERWRITE, ctypes.c_void_p(st))
    for _ in range(20):
        capi.check(L.sctl_b200_eval_async(*args))
    torch.cuda.synchronize()
    reps = 500
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter()
    e0.record()
    for _ in range(reps):
        L.sctl_b200_eval_async(*args)
    e1.record()
    w1 = time.perf_counter()
    torch.cuda.synchronize()
    w2 = time.perf_counter()
    print(f"{d['name']} {np.dtype(dt).name} N=M={n:6d}: host enqueue {1e6 * (w1 - w0) / reps:7.1f} us/call, device {1e3 * e0.elapsed_time(e1) / reps:7.1f} us/call, pairs/s {n * n / (e0.elapsed_time(e1) / reps) * 1e3:.3e}")

# host-pointer entry (what the C++ header shim calls): pageable numpy buffers in, result out, synchronous
print("host-pointer entry sctl_b200_eval (pageable buffers, synchronous):")
for n in (64, 256, 1000, 4000, 16000, 20000):
    Xt, Xs, Xn, F = synth.make_problem(d, n, n, dt)
    U = np.zeros(n * d["trg_dim"], dtype=dt)
    hp = lambda a: None if a is None else a.ctypes.data_as(ctypes.c_void_p)
    args = (kid, capi.F64 if dt == np.float64 else capi.F32, n, n, hp(Xt), hp(Xs), hp(Xn) if d["normal_dim"] else None, hp(F), hp(U), -1, capi.OVERWRITE)
    for _ in range(10):
        capi.check(L.sctl_b200_eval(*args))
    reps = 200
    w0 = time.perf_counter()
    for _ in range(reps):
        L.sctl_b200_eval(*args)
    w1 = time.perf_counter()
    tm = capi.last_timing()
    print(f"{d['name']} {np.dtype(dt).name} N=M={n:6d}: wall {1e6 * (w1 - w0) / reps:8.1f} us/call   last call split (ms): {tm}")
