"""GPU: closed-surface identities as oracle-independent known answers (tests/identities.py), on both device paths: a
3200-node sphere (one-launch kernel) and a 320k-node sphere against 20000 targets (tiled TMA path)."""
import numpy as np
import pytest

from identities import sphere_quadrature, targets_inside_outside
from sctl_b200 import Laplace3D_DxU, Stokes3D_DxU

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dt,tol", [(np.float64, 1e-11), (np.float32, 3e-5)])
@pytest.mark.parametrize("nth,nph,nt", [(40, 80, 500), (400, 800, 10000)], ids=["3200-nodes", "320k-nodes"])
def test_gauss_law_laplace_double_layer(dt, tol, nth, nph, nt):
    Xs, Nw = sphere_quadrature(nth, nph, dtype=dt)
    Xt = targets_inside_outside(nt, dtype=dt)
    F = np.full(Xs.size // 3, 0.75, dtype=dt)
    u = Laplace3D_DxU().Eval(None, Xt, Xs, Nw, F).astype(np.float64)
    assert np.abs(u[:nt] + 0.75).max() < tol  # inside: -f
    assert np.abs(u[nt:]).max() < tol         # outside: 0


@pytest.mark.parametrize("dt,tol", [(np.float64, 1e-11), (np.float32, 3e-5)])
@pytest.mark.parametrize("nth,nph,nt", [(40, 80, 500), (400, 800, 10000)], ids=["3200-nodes", "320k-nodes"])
def test_stokes_double_layer_rigid_body_identity(dt, tol, nth, nph, nt):
    Xs, Nw = sphere_quadrature(nth, nph, dtype=dt)
    Xt = targets_inside_outside(nt, dtype=dt)
    f = np.array([0.3, -1.1, 0.6])
    F = np.tile(f, Xs.size // 3).astype(dt)
    u = Stokes3D_DxU().Eval(None, Xt, Xs, Nw, F).astype(np.float64).reshape(-1, 3)
    assert np.abs(u[:nt] + f[None, :]).max() < tol
    assert np.abs(u[nt:]).max() < tol
