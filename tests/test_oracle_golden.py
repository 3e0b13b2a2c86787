"""CPU: pins the oracle (C port, oracle/oracle_port.c) — and the C-ABI metadata — to golden vectors produced by the
UNMODIFIED reference (tests/tools/make_golden.py -> tests/golden/). The reference ships no golden vectors for this path
(SURVEY.md section 4), so its own outputs, generated here, are the pin."""
import numpy as np
import pytest

from golden_util import TAGS, inputs, load
from sctl_b200 import capi, synth
from util import rel_l2

# the port uses 1/sqrt where the reference refines a 14-bit seed (intrin-wrapper.hpp:3025-3050): agreement to rounding
PIN_TOL = {"f64": 5e-15, "f32": 3e-6}


@pytest.mark.parametrize("kid", range(8))
def test_metadata_matches_reference(port, kid):
    g, d = load(kid)
    assert port.kernel_dims(kid) == d
    assert capi.kernel_dims(kid) == d  # C ABI metadata needs no GPU


@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("kid", range(8))
def test_port_eval_matches_golden(port, kid, tag):
    g, d = load(kid)
    for ci in range(len(g["cases"])):
        Xt, Xs, Xn, F = inputs(g, d, kid, ci, TAGS[tag])
        got = port.eval(kid, Xt, Xs, Xn, F)
        assert rel_l2(got, g[f"eval_{tag}_{ci}"]) <= PIN_TOL[tag], (d["name"], ci)
        U0 = synth.uniform(9, 0, got.size, seed=kid, dtype=TAGS[tag])
        got_acc = port.eval(kid, Xt, Xs, Xn, F, v_trg=U0.copy())
        assert rel_l2(got_acc, g[f"evalacc_{tag}_{ci}"]) <= PIN_TOL[tag]
        assert rel_l2(port.eval_direct(kid, Xt, Xs, Xn, F), g[f"direct_{tag}_{ci}"]) <= PIN_TOL[tag]


@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("kid", range(8))
def test_port_kernel_matrix_matches_golden(port, kid, tag):
    g, d = load(kid)
    n = 0
    for ci in range(len(g["cases"])):
        if f"kmat_{tag}_{ci}" not in g:
            continue
        Xt, Xs, Xn, _ = inputs(g, d, kid, ci, TAGS[tag])
        M = port.kernel_matrix(kid, Xt, Xs, Xn)
        want = g[f"kmat_{tag}_{ci}"]
        assert M.shape == want.shape == (Xs.size // 3 * d["src_dim"], Xt.size // 3 * d["trg_dim"])
        assert rel_l2(M, want) <= PIN_TOL[tag]
        n += 1
    assert n >= 3


@pytest.mark.parametrize("kid", range(8))
def test_golden_self_consistency(kid):
    """reference facts: EvalDirect == Eval from zero (fmm-wrapper.txx:501-557); U = F^T M (generic-kernel.txx:85-89,284);
    coincident source/target pairs contribute 0 (first `coincident` targets sit on sources and stay finite)"""
    g, d = load(kid)
    for ci in range(len(g["cases"])):
        assert rel_l2(g[f"direct_f64_{ci}"], g[f"eval_f64_{ci}"]) <= 1e-15
        assert np.all(np.isfinite(g[f"eval_f64_{ci}"]))
        if f"kmat_f64_{ci}" in g:
            _, _, _, F = inputs(g, d, kid, ci, np.float64)
            assert rel_l2(F @ g[f"kmat_f64_{ci}"], g[f"eval_f64_{ci}"]) <= 1e-13


def test_reference_still_matches_its_golden(reference):
    """only where oracle/_ref is built: the compiled reference reproduces the committed vectors (any ISA variant)"""
    for kid in (0, 3, 4):
        g, d = load(kid)
        Xt, Xs, Xn, F = inputs(g, d, kid, 3, np.float64)
        assert rel_l2(reference.eval(kid, Xt, Xs, Xn, F), g["eval_f64_3"]) <= 1e-14


def test_rsqrt_golden_accuracy():
    """approx_rsqrt pins of the reference's own test (vec-test.hpp:384-418): digits 4 -> 2^-11, digits 7 -> 2^-22,
    masked (x == 0) lanes exactly 0; recorded full-precision values are within a few ulp"""
    import os

    from golden_util import GOLDEN

    g = np.load(os.path.join(GOLDEN, "rsqrt.npz"))
    for x, pre, eps in ((g["x64"], "y64", 2.0 ** -52), (g["x32"], "y32", 2.0 ** -23)):
        exact = np.where(x > 0, 1.0 / np.sqrt(np.where(x > 0, x, 1).astype(np.float64)), 0.0)
        for dg, tol in ((4, 2.0 ** -11), (7, 2.0 ** -22), (-1, 8 * eps)):
            y = g[f"{pre}_d{dg}"].astype(np.float64)
            assert y[0] == 0.0
            assert np.max(np.abs(y[1:] - exact[1:]) / exact[1:]) <= tol
