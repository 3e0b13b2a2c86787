"""Measurement of the SURVEY 8(f) rows on one B200 (device-resident data, CUDA-event / wall time):
  f1/f4  GPU-aware ParticleFMM registry in a solver loop (SetSrcDensity + Eval per iteration, coordinates resident, weights
         and target-normal contraction fused) against re-sending everything through the host-pointer Eval every iteration;
  f3     batched KernelMatrix: thousands of small (far nodes x near targets) blocks in ONE launch against one
         KernelMatrix call per block (what BoundaryIntegralOp::SetupNear does, boundary_integral.txx:946-1008)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sctl_b200 import ParticleFMM, Stokes3D_DxU, Stokes3D_FxU, capi, synth

dev = torch.device("cuda:0")
sl, dl = Stokes3D_FxU(), Stokes3D_DxU()

print("== f1/f4: solver loop, Stokes3D_DxU f64, density changes every iteration ==")
for N in (2000, 20000, 200000):
    d = capi.kernel_dims(4)
    Xt, Xs, Xn, F = synth.make_problem(d, N, N, np.float64)
    W = np.abs(synth.uniform(9, 0, N, dtype=np.float64)) + 0.5
    fmm = ParticleFMM(np.float64)
    fmm.SetKernels(sl, sl, sl); fmm.AddTrg("U", sl, sl); fmm.AddSrc("DL", dl, dl); fmm.SetKernelS2T("DL", "U", dl)
    fmm.SetTrgCoord("U", Xt); fmm.SetSrcCoord("DL", Xs, Xn); fmm.SetSrcWeights("DL", W)
    U = np.zeros(3 * N)
    its = 20 if N <= 20000 else 5
    for warm in (True, False):
        t0 = time.perf_counter()
        for it in range(2 if warm else its):
            fmm.SetSrcDensity("DL", F)
            U = fmm.EvalDirect("U", U)
        t_reg = (time.perf_counter() - t0) / (2 if warm else its)
    for warm in (True, False):
        t0 = time.perf_counter()
        for it in range(2 if warm else its):
            Fw = (F.reshape(N, 3) * W[:, None]).reshape(-1)  # the host-side scaling of boundary_integral.txx:1036-1051
            U2 = dl.Eval(U, Xt, Xs, Xn, Fw, overwrite=True)
        t_host = (time.perf_counter() - t0) / (2 if warm else its)
    err = np.linalg.norm(U - U2) / np.linalg.norm(U2)
    print(f"N=M={N:7d}: registry {1e3 * t_reg:9.3f} ms/iter ({N * N / t_reg:.3e} pairs/s wall)   host-pointer Eval {1e3 * t_host:9.3f} ms/iter   speed-up {t_host / t_reg:.2f}x   rel diff {err:.1e}", flush=True)

print("== f3: batched KernelMatrix, Stokes3D_DxU f64, blocks of 96 far nodes x 300 near targets ==")
for nb in (256, 4096):
    ns_b, nt_b = 96, 300
    d = capi.kernel_dims(4)
    Xt, Xs, Xn, _ = synth.make_problem(d, nb * nt_b, nb * ns_b, np.float64)
    t = lambda a: torch.from_numpy(a).to(dev)
    Xt, Xs, Xn = t(Xt), t(Xs), t(Xn)
    trg_dsp = np.arange(nb + 1, dtype=np.int64) * nt_b
    src_dsp = np.arange(nb + 1, dtype=np.int64) * ns_b
    M, M_dsp = dl.KernelMatrixBatch(trg_dsp, src_dsp, Xt, Xs, Xn)
    torch.cuda.synchronize()
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        dl.KernelMatrixBatch(trg_dsp, src_dsp, Xt, Xs, Xn, out=M, M_dsp=M_dsp)
    torch.cuda.synchronize()
    t_b = (time.perf_counter() - t0) / reps
    nloop = min(nb, 256)
    outs = [torch.empty((ns_b * 3, nt_b * 3), dtype=torch.float64, device=dev) for _ in range(nloop)]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for b in range(nloop):
        dl.KernelMatrix(Xt[3 * trg_dsp[b]:3 * trg_dsp[b + 1]], Xs[3 * src_dsp[b]:3 * src_dsp[b + 1]], Xn[3 * src_dsp[b]:3 * src_dsp[b + 1]], out=outs[b])
    torch.cuda.synchronize()
    t_l = (time.perf_counter() - t0) / nloop * nb
    gb = M.numel() * 8 / 1e9
    ok = torch.equal(M[M_dsp[1]:M_dsp[1] + outs[1].numel()], outs[1].reshape(-1))
    print(f"{nb:5d} blocks ({gb:.2f} GB of M): batched {1e3 * t_b:8.3f} ms = {gb / t_b:7.0f} GB/s written   one call per block {1e3 * t_l:8.3f} ms   speed-up {t_l / t_b:.1f}x   identical block: {ok}", flush=True)
