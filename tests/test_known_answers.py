"""Analytic two-particle known answers — the reference has none, so they are added here.
CPU: against the oracle port. GPU (-m gpu): the same values through the C ABI."""
import numpy as np
import pytest

from sctl_b200 import KERNELS, capi

PI = np.pi
r = np.array([0.3, -0.2, 0.5])
n = np.array([0.7, 0.1, -0.4])
f3 = np.array([0.9, -0.6, 0.25])
f4 = np.array([0.9, -0.6, 0.25, -0.8])
R = np.linalg.norm(r)


def analytic(kid):
    """source at the origin, target at r: returns (n_src or None, v_src, expected v_trg)"""
    if kid == 0:
        return None, np.array([2.0]), np.array([2.0 / (4 * PI * R)])
    if kid == 1:
        return n, np.array([2.0]), np.array([2.0 * (r @ n) / (4 * PI * R**3)])
    if kid == 2:
        return None, np.array([2.0]), -2.0 * r / (4 * PI * R**3)
    if kid == 3:
        return None, f3, (f3 / R + r * (r @ f3) / R**3) / (8 * PI)
    if kid == 4:
        return n, f3, 3 * r * (r @ f3) * (r @ n) / (4 * PI * R**5)
    if kid == 5:
        return None, f3, (-3 / (4 * PI)) * np.outer(r, r).ravel() * (r @ f3) / R**5
    if kid == 6:
        return None, f4, (f3 / R + r * (r @ f3) / R**3 + f4[3] * r / R**3) / (8 * PI)
    if kid == 7:
        return None, f3, np.concatenate([(f3 / R + r * (r @ f3) / R**3), [(r @ f3) / R**3]]) / (8 * PI)


@pytest.mark.parametrize("kid", range(8))
def test_port_known_answer(port, kid):
    nn, f, want = analytic(kid)
    got = port.eval(kid, r.copy(), np.zeros(3), nn, f)
    assert np.allclose(got, want, rtol=1e-14, atol=0)
    # translation invariance and accumulate semantics: {10, 20} + value (probe in SURVEY.md 8c)
    sh = np.array([5.0, -3.0, 1.0])
    U = np.full(want.size, 10.0)
    got2 = port.eval(kid, r + sh, sh.copy(), nn, f, v_trg=U)
    assert np.allclose(got2 - 10.0, want, rtol=1e-12)


def test_port_laplace_gradient_is_fd_of_potential(port):
    h = 1e-5
    src = np.array([0.1, 0.2, -0.3])
    fs = np.array([1.7])
    g = port.eval(2, r.copy(), src, None, fs)
    # FxdU has scale -1/(4 pi) and u = r/|r|^3, i.e. it IS the gradient of FxU's f/(4 pi |r|)  (kernel_functions.hpp:61-70)
    for k in range(3):
        e = np.zeros(3)
        e[k] = h
        fd = (port.eval(0, r + e, src, None, fs) - port.eval(0, r - e, src, None, fs)) / (2 * h)
        assert abs(fd[0] - g[k]) <= 1e-8 * abs(g[k]) + 1e-10


@pytest.mark.gpu
@pytest.mark.parametrize("dt,rtol", [(np.float64, 1e-13), (np.float32, 2e-6)])
@pytest.mark.parametrize("kid", range(8))
def test_cuda_known_answer(kid, dt, rtol):
    nn, f, want = analytic(kid)
    c = lambda a: None if a is None else a.astype(dt)
    got = KERNELS[kid]().Eval(None, c(r), np.zeros(3, dtype=dt), c(nn), c(f))
    assert np.allclose(got, want, rtol=rtol, atol=rtol * np.max(np.abs(want)))
    M = KERNELS[kid]().KernelMatrix(c(r), np.zeros(3, dtype=dt), c(nn))
    assert np.allclose(f.astype(dt) @ M, want, rtol=10 * rtol, atol=10 * rtol * np.max(np.abs(want)))
    assert M.shape == (capi.kernel_dims(kid)["src_dim"], capi.kernel_dims(kid)["trg_dim"])
