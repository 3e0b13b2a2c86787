import numpy as np


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    nb = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / (nb if nb > 0 else 1.0))


# contract of BASELINE.json north_star: rel-L2 <= 1e-12 (f64) / 1e-5 (f32) against the reference's CPU Eval
TOL = {np.dtype(np.float64): 1e-12, np.dtype(np.float32): 1e-5}
