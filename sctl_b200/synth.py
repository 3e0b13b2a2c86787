"""Deterministic synthetic inputs shared by tests, bench.py and the CPU baseline.

Counter-based generator: u(stream, i) = splitmix64(seed ^ (stream << 56) + i) -> [0, 1) with 53 random bits, so any
slice of any array can be produced independently on any rank. Coordinates and (un-normalised) normals are
uniform in [-0.5, 0.5)^3 and densities uniform in [-0.5, 0.5), mirroring the reference's own test
(/root/reference/include/sctl/fmm-wrapper.txx:48-55). The first min(1024, N) targets coincide with the first
sources to exercise the r2 == 0 mask (kernel_functions.hpp:28).
"""
from __future__ import annotations

import numpy as np

STREAM_TRG, STREAM_SRC, STREAM_NRM, STREAM_DEN = 1, 2, 3, 4
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x += np.uint64(0x9E3779B97F4A7C15)
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def uniform(stream: int, start: int, count: int, seed: int = 0, dtype=np.float64) -> np.ndarray:
    """count values of stream `stream`, starting at element index `start`, uniform in [-0.5, 0.5)."""
    idx = np.arange(start, start + count, dtype=np.uint64)
    with np.errstate(over="ignore"):
        key = idx + np.uint64((seed & 0xFFFFFFFF) ^ (stream << 56))
    u = (splitmix64(key) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    return (u - 0.5).astype(dtype)


def make_problem(kernel_dims: dict, Nt: int, Ns: int, dtype=np.float64, seed: int = 0, t0: int = 0, t1: int | None = None,
                 s0: int = 0, s1: int | None = None, coincident: int = 1024):
    """Returns (r_trg[t0:t1], r_src[s0:s1], n_src[s0:s1] or None, v_src[s0:s1]) of the global (Nt, Ns) problem."""
    t1 = Nt if t1 is None else t1
    s1 = Ns if s1 is None else s1
    Nd, K0 = kernel_dims["normal_dim"], kernel_dims["src_dim"]
    r_src = uniform(STREAM_SRC, 3 * s0, 3 * (s1 - s0), seed, dtype)
    r_trg = uniform(STREAM_TRG, 3 * t0, 3 * (t1 - t0), seed, dtype)
    nco = min(coincident, Nt, Ns)
    if t0 < nco:  # first nco targets sit exactly on the first nco sources
        hi = min(nco, t1)
        r_trg[: 3 * (hi - t0)] = uniform(STREAM_SRC, 3 * t0, 3 * (hi - t0), seed, dtype)
    n_src = uniform(STREAM_NRM, Nd * s0, Nd * (s1 - s0), seed, dtype) if Nd else None
    v_src = uniform(STREAM_DEN, K0 * s0, K0 * (s1 - s0), seed, dtype)
    return r_trg, r_src, n_src, v_src
