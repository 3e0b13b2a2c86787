"""Closed-surface identities used as oracle-independent known answers (the reference's legacy Quadrature::test checks
Green's identity with Laplace3D_FxU/DxU the same way, boundary_quadrature.hpp:1584-1700).

For a closed surface S with outward normal n and a CONSTANT density f, with r = x - y:
    Laplace double layer   (1/4pi) int_S (r.n)/|r|^3 f dS_y            = -f inside S, 0 outside        (Gauss' law)
    Stokes  double layer   (3/4pi) int_S r_i r_j (r.n)/|r|^5 f_i dS_y  = -f_j inside S, 0 outside      (rigid-body identity)
The kernels take the normal un-normalised (kernel_functions.hpp:36-51, :100-120), so the quadrature weight is folded into it:
n_src = n * w. The sphere is discretised with Gauss-Legendre nodes in cos(theta) x uniform nodes in phi (spectrally accurate
for targets away from the surface)."""
import numpy as np


def sphere_quadrature(n_theta, n_phi, radius=1.0, centre=(0.1, -0.2, 0.05), dtype=np.float64):
    """-> (points [N*3], weighted outward normals [N*3]) of a sphere, N = n_theta*n_phi"""
    mu, w = np.polynomial.legendre.leggauss(n_theta)  # mu = cos(theta)
    phi = 2 * np.pi * (np.arange(n_phi) + 0.5) / n_phi
    st = np.sqrt(1 - mu**2)
    nx = (st[:, None] * np.cos(phi)[None, :]).reshape(-1)
    ny = (st[:, None] * np.sin(phi)[None, :]).reshape(-1)
    nz = np.repeat(mu, n_phi)
    n = np.stack([nx, ny, nz], axis=1)
    wq = np.repeat(w, n_phi) * (2 * np.pi / n_phi) * radius**2
    X = np.asarray(centre)[None, :] + radius * n
    return np.ascontiguousarray(X.reshape(-1), dtype=dtype), np.ascontiguousarray((n * wq[:, None]).reshape(-1), dtype=dtype)


def targets_inside_outside(n, radius=1.0, centre=(0.1, -0.2, 0.05), seed=3, dtype=np.float64):
    """n targets well inside (|x-c| < 0.5 R) followed by n well outside (1.6 R < |x-c| < 3 R)"""
    rng = np.random.default_rng(seed)
    d = rng.standard_normal((2 * n, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    rad = np.concatenate([0.5 * radius * rng.random(n), radius * (1.6 + 1.4 * rng.random(n))])
    return np.ascontiguousarray((np.asarray(centre)[None, :] + d * rad[:, None]).reshape(-1), dtype=dtype)
