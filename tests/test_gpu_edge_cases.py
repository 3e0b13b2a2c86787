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
