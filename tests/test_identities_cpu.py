"""CPU: the oracle port against closed-surface identities (Gauss' law, Stokes rigid-body identity) - known answers
that do not come from the reference's code."""
import numpy as np

from identities import sphere_quadrature, targets_inside_outside


def test_gauss_law_laplace_double_layer(port):
    Xs, Nw = sphere_quadrature(40, 80)
    n = 50
    Xt = targets_inside_outside(n)
    F = np.full(Xs.size // 3, 0.75)
    u = port.eval(1, Xt, Xs, Nw, F)
    assert np.abs(u[:n] + 0.75).max() < 1e-9   # inside: -f
    assert np.abs(u[n:]).max() < 1e-9          # outside: 0


def test_stokes_double_layer_rigid_body_identity(port):
    Xs, Nw = sphere_quadrature(40, 80)
    n = 50
    Xt = targets_inside_outside(n)
    f = np.array([0.3, -1.1, 0.6])
    F = np.tile(f, Xs.size // 3)
    u = port.eval(4, Xt, Xs, Nw, F).reshape(-1, 3)
    assert np.abs(u[:n] + f[None, :]).max() < 1e-9
    assert np.abs(u[n:]).max() < 1e-9
