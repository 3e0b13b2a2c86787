"""Small workload for compute-sanitizer (memcheck): every kernel family once at ragged sizes."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sctl_b200 import KERNELS, ParticleFMM, Stokes3D_FxU, Stokes3D_DxU, capi, synth

for kid in range(8):
    d = capi.kernel_dims(kid)
    for dt in (np.float64, np.float32):
        for Nt, Ns in ((1, 1), (130, 700), (2100, 333)):
            Xt, Xs, Xn, F = synth.make_problem(d, Nt, Ns, dt, coincident=8)
            U = KERNELS[kid]().Eval(None, Xt, Xs, Xn, F)
            U = KERNELS[kid]().Eval(U, Xt, Xs, Xn, F, digits=8)
            assert np.all(np.isfinite(U))
        M = KERNELS[kid]().KernelMatrix(Xt[:3 * 70], Xs[:3 * 45], None if Xn is None else Xn[:3 * 45])
        trg_dsp, src_dsp = np.array([0, 40, 40, 170], np.int64), np.array([0, 33, 50, 111], np.int64)
        Mb, _ = KERNELS[kid]().KernelMatrixBatch(trg_dsp, src_dsp, Xt[:3 * 170], Xs[:3 * 111], None if Xn is None else Xn[:3 * 111])
os.environ["SCTL_B200_FORCE_CHUNKS"] = "5"
d = capi.kernel_dims(3)
Xt, Xs, Xn, F = synth.make_problem(d, 3000, 5000, np.float64)
KERNELS[3]().Eval(None, Xt, Xs, Xn, F)
fmm = ParticleFMM(np.float64)
sl, dl = Stokes3D_FxU(), Stokes3D_DxU()
fmm.AddTrg("T", sl, sl); fmm.AddSrc("S", sl, sl); fmm.AddSrc("D", dl, dl)
fmm.SetKernelS2T("S", "T", sl); fmm.SetKernelS2T("D", "T", dl)
fmm.SetTrgCoord("T", Xt); fmm.SetSrcCoord("S", Xs); fmm.SetSrcDensity("S", F)
fmm.SetSrcCoord("D", Xs[:900], Xt[:900]); fmm.SetSrcDensity("D", F[:900])
fmm.Eval("T")
print("sanitize target done, launches", capi.launch_count())
