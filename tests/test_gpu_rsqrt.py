"""The device reciprocal square root itself (sctl_b200/csrc/common.cuh: rsqrt_masked / rsqrt_unmasked, refinement levels
0 / 1 / 2), element by element through the test hook sctl_b200_test_rsqrt - not through sums.

It is the GPU counterpart of the one unit test the reference has for this arithmetic, VecTest::test_reals_rsqrt
(/root/reference/include/sctl/vec-test.hpp:384-418): approx_rsqrt<4> within 2^-11, approx_rsqrt<7> within 2^-22, masked
lanes exactly 0. On top of the reference's pins it checks the library's own claims (seed <= 2^-19, one Newton step <= 2^-37,
full precision <= 2 ulp), the digits -> level table (newton_level, capi.cu) against the measured bounds, the r2 > 0 mask
rule for zero and SUBNORMAL arguments (kernel_functions.hpp:28), and the property the unmasked hot loop relies on: a zero or
subnormal argument makes the unmasked form non-finite, so the tile is detected and repaired.
"""
import ctypes
import os

import numpy as np
import pytest

from sctl_b200 import capi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MASKED, UNMASKED, UNMASKED_WORST, MASKED_SEL = 1, 0, 2, 3
DT = {np.dtype(np.float32): capi.F32, np.dtype(np.float64): capi.F64}


def dev_rsqrt(x, newton, mode):
    x = np.ascontiguousarray(x)
    y = np.empty_like(x)
    capi.check(capi.lib().sctl_b200_test_rsqrt(DT[x.dtype], newton, mode, x.size, x.ctypes.data_as(ctypes.c_void_p), y.ctypes.data_as(ctypes.c_void_p)))
    return y


def exact(x):
    return 1.0 / np.sqrt(np.asarray(x, dtype=np.longdouble))


def max_rel(y, x):
    want = exact(x)
    return float(np.max(np.abs(np.asarray(y, dtype=np.longdouble) - want) / want))


def normal_inputs(dt, per_exp=8, seed=5):
    """2^k over the whole normal exponent range, each with the exact power of two, mantissas next to 1 and 2, and random mantissas"""
    fi = np.finfo(dt)
    rng = np.random.default_rng(seed)
    ks = np.arange(fi.minexp, fi.maxexp)  # 2^k normal for every k in here
    m = np.concatenate([[1.0, 1.0 + fi.eps, 2.0 - fi.eps, 1.5], 1.0 + rng.random(per_exp)])
    x = np.ldexp(np.repeat(m[None, :], ks.size, 0), ks[:, None]).astype(dt).ravel()
    assert np.all(np.isfinite(x)) and np.all(x >= fi.tiny)
    return x


# the library's claims (common.cuh / capi.cu newton_level): max relative error of rinv per refinement level
CLAIM = {(np.dtype(np.float64), 0): 2.0**-19, (np.dtype(np.float64), 1): 2.0**-37, (np.dtype(np.float64), 2): 2 * 2.0**-52,
         (np.dtype(np.float32), 0): 2.0**-21, (np.dtype(np.float32), 1): 2 * 2.0**-23, (np.dtype(np.float32), 2): 2 * 2.0**-23}


def test_newton_level_table_cpu():
    """digits -> level, no GPU needed: machine precision for -1 / >= 16 as the reference maps them (generic-kernel.txx:50-69)"""
    L = capi.lib()
    lv64 = [L.sctl_b200_newton_level(capi.F64, d) for d in range(0, 17)]
    lv32 = [L.sctl_b200_newton_level(capi.F32, d) for d in range(0, 17)]
    assert lv64 == [0] * 5 + [1] * 6 + [2] * 6
    assert lv32 == [0] * 6 + [1] * 11
    assert L.sctl_b200_newton_level(capi.F64, -1) == 2 and L.sctl_b200_newton_level(capi.F32, -1) == 1
    assert sorted(lv64) == lv64 and sorted(lv32) == lv32  # more digits never means less refinement
    assert L.sctl_b200_newton_level(5, 3) == -1


@pytest.mark.gpu
@pytest.mark.parametrize("dt", [np.float64, np.float32], ids=["f64", "f32"])
@pytest.mark.parametrize("mode", [MASKED, MASKED_SEL, UNMASKED, UNMASKED_WORST], ids=["masked", "masked-branch-free", "unmasked", "unmasked-worst-seed"])
def test_levels_meet_their_claims_over_the_whole_exponent_range(dt, mode):
    dt = np.dtype(dt)
    x = normal_inputs(dt)
    errs = {}
    for newton in (0, 1, 2):
        y = dev_rsqrt(x, newton, mode)
        assert np.all(np.isfinite(y)) and np.all(y > 0)
        errs[newton] = max_rel(y, x)
        assert errs[newton] <= CLAIM[(dt, newton)], f"{dt} level {newton} mode {mode}: max rel err {errs[newton]:.3e} = 2^{np.log2(errs[newton]):.2f}"
    print(f"{dt} mode {mode}: max rel err per level", {k: f"2^{np.log2(v):.2f}" for k, v in errs.items()})


@pytest.mark.gpu
@pytest.mark.parametrize("dt", [np.float64, np.float32], ids=["f64", "f32"])
def test_digits_table_never_delivers_fewer_digits(dt):
    """for every digits value the level picked by newton_level is accurate to 0.2 * 10^-digits (so that rinv^5 still holds
    `digits` digits), or to 4 eps where that is below machine precision - measured, unmasked worst-case seed"""
    dt = np.dtype(dt)
    x = normal_inputs(dt, per_exp=4, seed=6)
    err = {nw: max_rel(dev_rsqrt(x, nw, UNMASKED_WORST), x) for nw in (0, 1, 2)}
    err_m = {nw: max(max_rel(dev_rsqrt(x, nw, MASKED), x), max_rel(dev_rsqrt(x, nw, MASKED_SEL), x)) for nw in (0, 1, 2)}
    for digits in list(range(0, 17)) + [-1]:
        lv = capi.lib().sctl_b200_newton_level(DT[dt], digits)
        want_digits = digits if 0 <= digits < 16 else 17
        bound = max(0.2 * 10.0**-want_digits, 4 * np.finfo(dt).eps)
        assert max(err[lv], err_m[lv]) <= bound, (digits, lv, err[lv], err_m[lv], bound)


@pytest.mark.gpu
def test_reference_pins_vec_test_hpp(reference):
    """vec-test.hpp:384-418 restated for the device: positive random arguments, digits 4 -> 2^-11, digits 7 -> 2^-22 (judged in
    f64 like the reference does for float), masked lanes (here: x == 0) exactly 0; and agreement with the reference's own
    approx_rsqrt<digits>(x, x > 0) on the same inputs, and with the committed golden vectors."""
    L = capi.lib()
    rng = np.random.default_rng(11)
    for dt in (np.float64, np.float32):
        dt = np.dtype(dt)
        x = rng.integers(1, 2**31 - 1, 4096).astype(dt)  # the reference draws rand()
        x[::17] = 0  # masked lanes
        pos = x > 0
        for digits, tol in ((4, 2.0**-11), (7, 2.0**-22)):
            lv = L.sctl_b200_newton_level(DT[dt], digits)
            y = dev_rsqrt(x, lv, MASKED)
            assert np.all(y[~pos] == 0)  # exactly 0, sign included
            assert not np.any(np.signbit(y[~pos]))
            assert max_rel(y[pos], x[pos]) < tol
            yref = reference.rsqrt(x, digits)
            assert np.all(yref[~pos] == 0)
            assert np.max(np.abs(y[pos].astype(np.float64) - yref[pos]) / yref[pos]) < 2 * tol
        y = dev_rsqrt(x, 2, MASKED)
        yref = reference.rsqrt(x, -1)
        assert np.max(np.abs(y[pos].astype(np.float64) - yref[pos].astype(np.float64)) / yref[pos]) <= 4 * np.finfo(dt).eps
    g = np.load(os.path.join(GOLDEN, "rsqrt.npz"))
    for tag, dt in (("64", np.float64), ("32", np.float32)):
        x = g["x" + tag]
        for digits, tol in ((4, 2.0**-11), (7, 2.0**-22), (-1, 4 * np.finfo(dt).eps)):
            y = dev_rsqrt(x, L.sctl_b200_newton_level(DT[np.dtype(dt)], digits), MASKED)
            want = g[f"y{tag}_d{digits}"]
            assert y[0] == 0 and want[0] == 0
            assert np.max(np.abs(y[1:].astype(np.float64) - want[1:]) / want[1:]) < 2 * tol


@pytest.mark.gpu
@pytest.mark.parametrize("dt", [np.float64, np.float32], ids=["f64", "f32"])
def test_zero_and_subnormal_arguments(dt):
    """The r2 > 0 mask (kernel_functions.hpp:28): 0 -> exactly 0; every positive argument, SUBNORMAL ones included, a finite
    accurate rinv (the reference's seed instruction handles denormals, intrin-wrapper.hpp:3038-3049). The unmasked form must
    turn all of them into inf/NaN - that is what flags a tile for repair in eval_kernel."""
    dt = np.dtype(dt)
    fi = np.finfo(dt)
    rng = np.random.default_rng(3)
    sub = np.concatenate([[fi.smallest_subnormal, fi.smallest_subnormal * 3, fi.tiny / 2, fi.tiny * (1 - fi.eps)],
                          fi.tiny * rng.random(64)]).astype(dt)
    sub = sub[(sub > 0) & (sub < fi.tiny)]
    assert sub.size > 32
    for newton in (0, 1, 2):
        for masked in (MASKED, MASKED_SEL):
            y0 = dev_rsqrt(np.zeros(4, dt), newton, masked)
            assert np.all(y0 == 0) and not np.any(np.signbit(y0))
            y = dev_rsqrt(sub, newton, masked)
            assert np.all(np.isfinite(y))
            assert max_rel(y, sub) <= CLAIM[(dt, newton)]
            # smallest normal numbers: the boundary of the rescaling branch
            edge = np.array([fi.tiny, fi.tiny * (1 + fi.eps), fi.tiny * 2], dtype=dt)
            assert max_rel(dev_rsqrt(edge, newton, masked), edge) <= CLAIM[(dt, newton)]
        for mode in (UNMASKED, UNMASKED_WORST):
            bad = dev_rsqrt(np.concatenate([np.zeros(2, dt), sub]), newton, mode)
            assert not np.any(np.isfinite(bad)), "unmasked rsqrt of a zero/subnormal argument must be non-finite (tile repair trigger)"
