"""Edge cases of the drop-in semantics, on the GPU through the C ABI: empty inputs, self-interaction (targets ARE the
sources), aliasing of input arrays, far-apart magnitudes, non-finite inputs, and the error contract."""
import ctypes

import numpy as np
import pytest

from sctl_b200 import KERNELS, capi, synth
from util import TOL, rel_l2

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("eval_path")]


@pytest.mark.parametrize("kid", [0, 3, 4])
def test_empty_sources_and_targets(kid):
    d = capi.kernel_dims(kid)
    K1 = d["trg_dim"]
    ker = KERNELS[kid]()
    Xt = synth.uniform(1, 0, 3 * 10)
    e = np.zeros(0)
    # no sources: accumulate leaves v_trg alone, overwrite zeroes it (generic-kernel.txx:98-101 / fmm-wrapper.txx:501-502)
    U0 = np.arange(10 * K1, dtype=np.float64)
    assert np.array_equal(ker.Eval(U0.copy(), Xt, e, e if d["normal_dim"] else None, e), U0)
    assert np.all(ker.Eval(U0.copy(), Xt, e, e if d["normal_dim"] else None, e, overwrite=True) == 0)
    # no targets: nothing to do, a zero-length result
    Xs, Xn, F = synth.uniform(2, 0, 30), synth.uniform(3, 0, 30) if d["normal_dim"] else None, synth.uniform(4, 0, 10 * d["src_dim"])
    assert ker.Eval(None, e, Xs, Xn, F).size == 0
    assert ker.KernelMatrix(e, Xs, Xn).shape == (10 * d["src_dim"], 0)
    assert ker.KernelMatrix(Xt, e, e if d["normal_dim"] else None).shape == (0, 10 * K1)


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kid", range(8))
def test_self_interaction_targets_are_the_sources(checker, kid, dt):
    """the N-body use: every target coincides with exactly one source (the same array is passed twice) -> each thread repairs
    the tiles that hold its own points; results must match the reference's r2 > 0 mask exactly in structure"""
    d = capi.kernel_dims(kid)
    N = 3000
    _, X, Xn, F = synth.make_problem(d, 1, N, dt, coincident=0)
    want = checker.eval(kid, X, X, Xn, F)
    got = KERNELS[kid]().Eval(None, X, X, Xn, F)  # r_trg and r_src alias
    assert np.all(np.isfinite(got))
    assert rel_l2(got, want) <= TOL[np.dtype(dt)]


def test_duplicate_sources_and_clustered_points(checker):
    """many coincident pairs per tile (duplicated sources) and points clustered at 1e-9 spacing: no NaN leaks out of a repaired tile"""
    kid, dt = 3, np.float64
    d = capi.kernel_dims(kid)
    Xt, Xs, _, F = synth.make_problem(d, 1500, 2000, dt, coincident=0)
    Xs[3 * 500:3 * 1000] = Xs[:3 * 500]  # duplicates
    Xt[:3 * 700] = Xs[:3 * 700]  # 700 targets each hit 2 coincident sources
    Xt[3 * 700:3 * 800] = Xs[3 * 1200:3 * 1300] + 1e-9  # almost coincident: huge but finite contributions
    want = checker.eval(kid, Xt, Xs, None, F)
    got = KERNELS[kid]().Eval(None, Xt, Xs, None, F)
    assert np.all(np.isfinite(got))
    assert rel_l2(got, want) <= 1e-12


def test_wide_coordinate_range(checker):
    """coordinates O(1e6) with O(1) separations, and a far cluster at 1e12: dX must be formed in full precision"""
    kid, dt = 4, np.float64
    d = capi.kernel_dims(kid)
    Xt, Xs, Xn, F = synth.make_problem(d, 600, 900, dt, coincident=8)
    Xt, Xs = Xt + 1.0e6, Xs + 1.0e6
    Xs[3 * 800:] += 1.0e12
    want = checker.eval(kid, Xt, Xs, Xn, F)
    got = KERNELS[kid]().Eval(None, Xt, Xs, Xn, F)
    assert rel_l2(got, want) <= 1e-12


@pytest.mark.parametrize("dt,scale,tol", [(np.float32, 6e-20, 1e-2), (np.float64, 8e-155, 1e-8)], ids=["f32-6e-20", "f64-8e-155"])
def test_subnormal_r2_is_not_masked(reference, dt, scale, tol):
    """The mask of the reference is r2 > 0 (kernel_functions.hpp:28), not "r2 is a normal number": a point cloud scaled so that
    every |dX|^2 is SUBNORMAL (f32: |dX| < 1e-19, r2 < 1.18e-38; f64: |dX| < 1.4e-154, r2 < 2.2e-308) still interacts - MUFU.RSQ
    flushes subnormals, so rsqrt_masked rescales them, and the unmasked hot loop repairs the tile. Checked against the
    closed form in extended precision (tolerance = the precision a subnormal r2 still carries: its quantum is 1.4e-45 / 4.9e-324,
    so the closest pairs hold 3-4 digits in f32; a DROPPED contribution would be an O(1) error) and, on an AVX-512 host (whose
    seed instruction handles denormals, intrin-wrapper.hpp:3038-3049), against the reference itself at digits = 4 - the
    seed-only level: with Newton steps the reference squares the seed (y*y, intrin-wrapper.hpp:519-528), which overflows
    for r2 this small and turns its own result into -inf, so only its unrefined form can be compared. Laplace FxU only:
    rinv^2 of such a pair overflows in every other kernel, in the reference too."""
    kid = 0
    d = capi.kernel_dims(kid)
    Nt, Ns = 70, 300
    Xt, Xs, _, F = synth.make_problem(d, Nt, Ns, np.float64, coincident=5)
    Xt, Xs, F = (Xt * scale).astype(dt), (Xs * scale).astype(dt), F.astype(dt)
    dX = Xt.astype(np.longdouble).reshape(-1, 1, 3) - Xs.astype(np.longdouble).reshape(1, -1, 3)
    r2 = (dX**2).sum(-1)
    assert 0 < float(np.max(r2)) < np.finfo(dt).tiny and np.count_nonzero(r2 == 0) == 5
    rinv = np.where(r2 > 0, 1 / np.sqrt(np.where(r2 > 0, r2, 1)), 0)
    want = (rinv * F.astype(np.longdouble).reshape(1, Ns)).sum(1) / (4 * np.pi)
    unit = np.max(np.abs(want))  # potentials are O(1e19) / O(1e153): normalise before taking norms (their squares overflow f64)
    want = np.asarray(want / unit, dtype=np.float64)
    nrm = lambda u: np.asarray(u.astype(np.longdouble) / unit, dtype=np.float64)
    for digits in (-1, 8, 3):
        got = KERNELS[kid]().Eval(None, Xt, Xs, None, F, digits=digits)
        assert np.all(np.isfinite(got)) and np.all(got != 0)
        err = rel_l2(nrm(got), want)
        print(f"subnormal r2 {np.dtype(dt).name} digits {digits}: rel-L2 vs closed form {err:.3e}")
        assert err <= max(tol, 10.0**-digits if digits > 0 else 0)
    if "avx512" in reference.path:
        ref = reference.eval(kid, Xt, Xs, None, F, digits=4)
        assert np.all(np.isfinite(ref))
        assert rel_l2(nrm(KERNELS[kid]().Eval(None, Xt, Xs, None, F)), nrm(ref)) <= tol + 2e-4  # 14-bit seed of the reference


@pytest.mark.parametrize("Ns", [300, 256, 1, 2049])
def test_far_from_origin_f32_with_ragged_source_tile(checker, Ns):
    """f32 coordinates beyond sqrt(FLT_MAX) = 1.8e19 with a source count that is not a multiple of the 256-source tile: the tile
    padding must not invent geometry (a dummy source at the origin would overflow r2 and poison the tile with inf * 0)."""
    kid, dt = 0, np.float32
    d = capi.kernel_dims(kid)
    Xt, Xs, _, F = synth.make_problem(d, 500, Ns, np.float64, coincident=0)
    Xt, Xs = (3.0e19 + Xt * 1e16).astype(dt), (3.0e19 + Xs * 1e16).astype(dt)
    Xt[:3] = Xs[-3:]  # one target sits on the LAST source, the one the padding repeats
    F = F.astype(dt)
    f64 = lambda a: a.astype(np.float64)
    want = checker.eval(kid, f64(Xt), f64(Xs), None, f64(F))
    got = KERNELS[kid]().Eval(None, Xt, Xs, None, F)
    assert np.all(np.isfinite(got))
    assert rel_l2(got, want) <= 1e-5


def test_non_finite_inputs_do_not_hang_or_spread():
    """a NaN density or an Inf coordinate poisons every target it interacts with (as the reference's arithmetic does); the call
    returns, and later calls are unaffected (no sticky device error, no stale workspace)"""
    kid, dt = 0, np.float64
    d = capi.kernel_dims(kid)
    Xt, Xs, _, F = synth.make_problem(d, 300, 700, dt, coincident=4)
    ker = KERNELS[kid]()
    clean = ker.Eval(None, Xt, Xs, None, F)
    F2 = F.copy()
    F2[13] = np.nan
    assert np.all(np.isnan(ker.Eval(None, Xt, Xs, None, F2)))
    Xs2 = Xs.copy()
    Xs2[3 * 5] = np.inf  # r2 = inf: the Newton refinement evaluates inf * 0 = NaN, in the reference's SIMD path as here
    got = ker.Eval(None, Xt, Xs2, None, F)
    assert got.shape == clean.shape and np.all(np.isnan(got))
    assert np.array_equal(ker.Eval(None, Xt, Xs, None, F), clean)


def test_torch_outputs_are_validated():
    """device-pointer entries take raw data_ptr()s: a v_trg / out of the wrong dtype, device layout or size must raise, not write"""
    import torch

    k = KERNELS[3]()
    dev = torch.device("cuda:0")
    Xt = torch.zeros(30, dtype=torch.float64, device=dev)
    Xs = torch.ones(60, dtype=torch.float64, device=dev)
    F = torch.ones(60, dtype=torch.float64, device=dev)
    with pytest.raises(ValueError):
        k.Eval(torch.zeros(30, dtype=torch.float32, device=dev), Xt, Xs, None, F)  # would be an 8-byte write into 4-byte slots
    with pytest.raises(ValueError):
        k.Eval(torch.zeros(60, dtype=torch.float64, device=dev)[::2], Xt, Xs, None, F)  # strided view
    with pytest.raises(ValueError):
        k.Eval(None, Xt.reshape(10, 3).t(), Xs, None, F)  # non-contiguous r_trg
    with pytest.raises(ValueError):
        k.KernelMatrix(Xt, Xs, out=torch.zeros(60 * 30, dtype=torch.float32, device=dev))
    with pytest.raises(ValueError):
        k.KernelMatrix(Xt, Xs, out=torch.zeros(60 * 30 - 1, dtype=torch.float64, device=dev))
    from sctl_b200.fmm import ParticleFMM

    fmm = ParticleFMM(np.float64)
    fmm.AddSrc("s", k, k), fmm.AddTrg("t", k, k), fmm.SetKernelS2T("s", "t", k)
    fmm.SetSrcCoord("s", Xs), fmm.SetSrcDensity("s", F), fmm.SetTrgCoord("t", Xt)
    with pytest.raises(ValueError):  # a strided output would receive its result in a temporary copy
        fmm.EvalDirect("t", torch.zeros(60, dtype=torch.float64, device=dev)[::2])
    U = fmm.EvalDirect("t", torch.zeros(30, dtype=torch.float64, device=dev))
    assert torch.all(torch.isfinite(U)) and float(U.abs().sum()) > 0


def test_error_contract():
    L = capi.lib()
    p = ctypes.c_void_p(16)
    assert L.sctl_b200_eval(99, capi.F64, 1, 1, p, p, None, p, p, -1, 0) == capi.ERR_KERNEL
    assert L.sctl_b200_eval(0, 7, 1, 1, p, p, None, p, p, -1, 0) == capi.ERR_DTYPE
    assert L.sctl_b200_eval(0, capi.F64, -1, 1, p, p, None, p, p, -1, 0) == capi.ERR_ARG
    assert L.sctl_b200_eval(1, capi.F64, 1, 1, p, p, None, p, p, -1, 0) == capi.ERR_ARG  # Laplace DxU needs normals
    assert L.sctl_b200_eval(0, capi.F64, 1, 1, None, p, None, p, p, -1, 0) == capi.ERR_ARG
    assert b"null" in L.sctl_b200_last_error()
    with pytest.raises(ValueError):  # size mismatch aborts in the reference (generic-kernel.txx:94-97)
        KERNELS[3]().Eval(None, np.zeros(9), np.zeros(9), None, np.zeros(8))
