"""GPU-aware ParticleFMM registry (sctl_b200/fmm.py -> sctl_b200_fmm_* C ABI) against the CPU checker. Mirrors the
reference's own test set-up (fmm-wrapper.txx:35-93): Stokes double-layer sources -> "Velocity" targets, plus a second
(single-layer) source type summed into the same target (fmm-wrapper.txx:514-559)."""
import numpy as np
import pytest

from sctl_b200 import ParticleFMM, Stokes3D_DxU, Stokes3D_FSxU, Stokes3D_FxU, Laplace3D_FxU, capi, synth
from util import TOL, rel_l2

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("eval_path")]


def _setup(dt, N):
    m2l, sl, dl = Stokes3D_FSxU(), Stokes3D_FxU(), Stokes3D_DxU()
    fmm = ParticleFMM(dt)
    fmm.SetAccuracy(-1)
    fmm.SetKernels(m2l, m2l, sl)
    fmm.AddTrg("Velocity", m2l, sl)
    fmm.AddSrc("DoubleLayer", dl, dl)
    fmm.AddSrc("SingleLayer", sl, sl)
    fmm.SetKernelS2T("DoubleLayer", "Velocity", dl)
    fmm.SetKernelS2T("SingleLayer", "Velocity", sl)
    Xt = synth.uniform(1, 0, 3 * N, dtype=dt)
    Xd, Nd, Fd = (synth.uniform(s, 0, 3 * N, dtype=dt) for s in (2, 3, 4))
    Xs, Fs = (synth.uniform(s, 0, 3 * (N + 7), dtype=dt) for s in (5, 6))
    Xt[: 3 * 16] = Xd[: 3 * 16]  # coincident points
    return fmm, Xt, Xd, Nd, Fd, Xs, Fs


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_registry_matches_checker(checker, dt, eval_path):
    N = 5000
    fmm, Xt, Xd, Nd, Fd, Xs, Fs = _setup(dt, N)
    fmm.SetTrgCoord("Velocity", Xt)
    fmm.SetSrcCoord("DoubleLayer", Xd, Nd)
    fmm.SetSrcCoord("SingleLayer", Xs)
    fmm.SetSrcDensity("DoubleLayer", Fd)
    fmm.SetSrcDensity("SingleLayer", Fs)
    want = checker.eval(4, Xt, Xd, Nd, Fd) + checker.eval(3, Xt, Xs, None, Fs)
    U = fmm.Eval("Velocity")
    assert U.dtype == dt and rel_l2(U, want) <= TOL[np.dtype(dt)]
    # solver-loop pattern: only the density changes; stale U contents are overwritten (fmm-wrapper.txx:501-502)
    n0 = capi.launch_count()
    Fd2 = -0.5 * Fd
    fmm.SetSrcDensity("DoubleLayer", Fd2)
    U2 = fmm.EvalDirect("Velocity", np.full_like(U, 123.0))
    want2 = checker.eval(4, Xt, Xd, Nd, Fd2) + checker.eval(3, Xt, Xs, None, Fs)
    assert rel_l2(U2, want2) <= TOL[np.dtype(dt)]
    n1 = capi.launch_count()
    # nothing changed -> no re-pack at all: strictly fewer launches than the previous evaluation, identical result
    U3 = fmm.Eval("Velocity")
    assert np.array_equal(U3, U2)
    if eval_path == "tiled":
        assert capi.launch_count() - n1 < n1 - n0
    else:  # the one-launch kernel reads the resident AoS arrays directly: one launch per source type, never a pack
        assert capi.launch_count() - n1 == n1 - n0 == 2
    # DeleteSrc drops its S2T kernels (fmm-wrapper.txx:407-418)
    fmm.DeleteSrc("SingleLayer")
    assert rel_l2(fmm.Eval("Velocity"), checker.eval(4, Xt, Xd, Nd, Fd2)) <= TOL[np.dtype(dt)]


def test_registry_device_tensors(checker):
    import torch

    N = 4000
    fmm, Xt, Xd, Nd, Fd, Xs, Fs = _setup(np.float64, N)
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(a).to(dev)
    fmm.SetTrgCoord("Velocity", t(Xt))
    fmm.SetSrcCoord("DoubleLayer", t(Xd), t(Nd))
    fmm.SetSrcCoord("SingleLayer", t(Xs))
    fmm.SetSrcDensity("DoubleLayer", t(Fd))
    fmm.SetSrcDensity("SingleLayer", t(Fs))
    U = torch.full((3 * N,), 7.0, dtype=torch.float64, device=dev)
    fmm.Eval("Velocity", U)
    want = checker.eval(4, Xt, Xd, Nd, Fd) + checker.eval(3, Xt, Xs, None, Fs)
    assert rel_l2(U.cpu().numpy(), want) <= 1e-12


def test_registry_errors():
    fmm = ParticleFMM(np.float64)
    sl, lap = Stokes3D_FxU(), Laplace3D_FxU()
    fmm.AddSrc("S", sl, sl)
    fmm.AddTrg("T", sl, sl)
    with pytest.raises(capi.SctlB200Error):  # duplicate name (fmm-wrapper.txx:298)
        fmm.AddSrc("S", sl, sl)
    with pytest.raises(capi.SctlB200Error):  # unknown target (fmm-wrapper.txx:373)
        fmm.SetKernelS2T("S", "nope", sl)
    with pytest.raises(capi.SctlB200Error):  # dims disagree with the declared source (CheckKernelDims, fmm-wrapper.txx:596-598)
        fmm.SetKernelS2T("S", "T", lap)
    fmm.SetKernelS2T("S", "T", sl)
    fmm.SetTrgCoord("T", np.zeros(30))
    fmm.SetSrcCoord("S", np.ones(60))
    fmm.SetSrcDensity("S", np.ones(10))  # wrong length
    with pytest.raises(capi.SctlB200Error):  # F.Dim() != Ns*SrcDim (fmm-wrapper.txx:533)
        fmm.Eval("T")
    # a target without sources evaluates to zero
    fmm.AddTrg("T2", sl, sl)
    fmm.SetTrgCoord("T2", np.zeros(30))
    assert np.all(fmm.Eval("T2") == 0)


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_far_field_glue_weights_and_target_normal(checker, dt):
    """the host-side glue of BoundaryIntegralOp::ComputeFarField fused on the device: F_far *= wts_far (boundary_integral.txx:1036-1051)
    and the target-normal dot product after Eval (:1060-1071), here with Laplace3D_FxdU (TrgDim 3 -> 1 value per target)"""
    from sctl_b200 import Laplace3D_FxdU

    N, Ns = 3000, 4100
    ker = Laplace3D_FxdU()
    fmm = ParticleFMM(dt)
    fmm.AddSrc("Src", ker, ker)
    fmm.AddTrg("Trg", ker, ker)
    fmm.SetKernelS2T("Src", "Trg", ker)
    Xt, Xs = synth.uniform(1, 0, 3 * N, dtype=dt), synth.uniform(2, 0, 3 * Ns, dtype=dt)
    F, W, Nt = synth.uniform(4, 0, Ns, dtype=dt), synth.uniform(5, 0, Ns, dtype=dt) + dt(1), synth.uniform(7, 0, 3 * N, dtype=dt)
    fmm.SetTrgCoord("Trg", Xt)
    fmm.SetSrcCoord("Src", Xs)
    fmm.SetSrcDensity("Src", F)
    fmm.SetSrcWeights("Src", W)
    fmm.SetTrgNormal("Trg", Nt)
    U = fmm.Eval("Trg")
    full = checker.eval(2, Xt, Xs, None, (F * W).astype(dt)).astype(np.float64).reshape(N, 3)
    want = (full * Nt.astype(np.float64).reshape(N, 3)).sum(1)
    assert U.shape == (N,) and rel_l2(U, want) <= 20 * TOL[np.dtype(dt)]  # the dot product can cancel: looser than the per-component bound
    # removing both restores the plain result
    fmm.SetSrcWeights("Src", None)
    fmm.SetTrgNormal("Trg", None)
    assert rel_l2(fmm.Eval("Trg"), checker.eval(2, Xt, Xs, None, F)) <= TOL[np.dtype(dt)]


def test_sharded_registry_single_rank(checker):
    """ShardedParticleFMM without a process group (one rank): device tensors in, device tensor out, same numbers"""
    import torch

    from sctl_b200.distributed import ShardedParticleFMM

    N = 3000
    dt = np.float64
    sl, dl = Stokes3D_FxU(), Stokes3D_DxU()
    Xt, Xd, Nd, Fd = (synth.uniform(s, 0, 3 * N, dtype=dt) for s in (1, 2, 3, 4))
    fmm = ShardedParticleFMM(dt)
    fmm.SetKernels(sl, sl, sl)
    fmm.AddTrg("U", sl, sl)
    fmm.AddSrc("DL", dl, dl)
    fmm.SetKernelS2T("DL", "U", dl)
    fmm.SetTrgCoord("U", torch.from_numpy(Xt).cuda())
    fmm.SetSrcCoord("DL", Xd, Nd)
    fmm.SetSrcDensity("DL", Fd)
    U = fmm.Eval("U")
    assert U.is_cuda and U.numel() == 3 * N
    assert rel_l2(U.cpu().numpy(), checker.eval(4, Xt, Xd, Nd, Fd)) <= TOL[np.dtype(dt)]
