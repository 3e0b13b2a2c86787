"""TEST INFRASTRUCTURE — ctypes loaders for the two CPU checkers. NOT part of the product path.

  port()      -> oracle/liboracle_port.so   plain-C restatement (oracle_port.c), builds anywhere
  reference() -> oracle/_ref/libsctl_ref_<isa>.so   the UNMODIFIED reference headers compiled by oracle/Makefile
                 (built where /root/reference exists; prebuilt files travel to the GPU box). The ISA variant
                 is chosen from /proc/cpuinfo because the reference selects its SIMD width at compile time.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline leg / --impl reference) may import this.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_P = ctypes.c_void_p
_I64 = ctypes.c_int64


def build_port() -> str:
    subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)
    return os.path.join(HERE, "liboracle_port.so")


def _cpu_flags() -> set:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def reference_path() -> str | None:
    flags = _cpu_flags()
    order = []
    if {"avx512f", "avx512dq", "avx512bw", "avx512vl"} <= flags:
        order.append("avx512")
    if {"avx2", "fma"} <= flags:
        order.append("avx2")
    if "sse4_2" in flags:
        order.append("sse42")
    for isa in order:
        p = os.path.join(HERE, "_ref", f"libsctl_ref_{isa}.so")
        if os.path.exists(p):
            return p
    return None


def _ptr(a):
    return None if a is None or a.size == 0 else a.ctypes.data_as(_P)


def _dt(a) -> int:
    return {np.dtype(np.float32): 0, np.dtype(np.float64): 1}[np.dtype(a.dtype)]


class _Checker:
    def __init__(self, lib, prefix, kind, path):
        self.lib, self.prefix, self.kind, self.path = lib, prefix, kind, path
        f = getattr(lib, prefix + "_kernel_dims")
        f.restype = ctypes.c_int
        f.argtypes = [ctypes.c_int] + [ctypes.POINTER(ctypes.c_int)] * 4 + [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int),
                                                                            ctypes.c_char_p, ctypes.c_int]

    def kernel_dims(self, kid: int) -> dict:
        c, n, s, t, fl = (ctypes.c_int() for _ in range(5))
        sc = ctypes.c_double()
        name = ctypes.create_string_buffer(64)
        rc = getattr(self.lib, self.prefix + "_kernel_dims")(kid, c, n, s, t, sc, fl, name, 64)
        assert rc == 0
        return dict(coord_dim=c.value, normal_dim=n.value, src_dim=s.value, trg_dim=t.value, scale=sc.value, flops=fl.value,
                    name=name.value.decode())


class Port(_Checker):
    """oracle_port.c"""

    def __init__(self):
        path = build_port()
        lib = ctypes.CDLL(path)
        super().__init__(lib, "oracle", "port", path)
        lib.oracle_eval.restype = ctypes.c_int
        lib.oracle_eval.argtypes = [ctypes.c_int, ctypes.c_int, _I64, _I64, _P, _P, _P, _P, _P]
        lib.oracle_eval_direct.restype = ctypes.c_int
        lib.oracle_eval_direct.argtypes = [ctypes.c_int, ctypes.c_int, _I64, _I64, _P, _P, _P, _P, _P]
        lib.oracle_kernel_matrix.restype = ctypes.c_int
        lib.oracle_kernel_matrix.argtypes = [ctypes.c_int, ctypes.c_int, _I64, _I64, _P, _P, _P, _P]

    def eval(self, kid, r_trg, r_src, n_src, v_src, v_trg=None, **_):
        d = self.kernel_dims(kid)
        Nt, Ns = r_trg.size // 3, r_src.size // 3
        if v_trg is None:
            v_trg = np.zeros(Nt * d["trg_dim"], dtype=r_trg.dtype)
        rc = self.lib.oracle_eval(kid, _dt(r_trg), Nt, Ns, _ptr(r_trg), _ptr(r_src), _ptr(n_src), _ptr(v_src), _ptr(v_trg))
        assert rc == 0
        return v_trg

    def eval_direct(self, kid, r_trg, r_src, n_src, v_src, **_):
        d = self.kernel_dims(kid)
        Nt, Ns = r_trg.size // 3, r_src.size // 3
        v_trg = np.full(Nt * d["trg_dim"], 7.0, dtype=r_trg.dtype)
        rc = self.lib.oracle_eval_direct(kid, _dt(r_trg), Nt, Ns, _ptr(r_trg), _ptr(r_src), _ptr(n_src), _ptr(v_src), _ptr(v_trg))
        assert rc == 0
        return v_trg

    def kernel_matrix(self, kid, r_trg, r_src, n_src, **_):
        d = self.kernel_dims(kid)
        Nt, Ns = r_trg.size // 3, r_src.size // 3
        M = np.zeros((Ns * d["src_dim"], Nt * d["trg_dim"]), dtype=r_trg.dtype)
        rc = self.lib.oracle_kernel_matrix(kid, _dt(r_trg), Nt, Ns, _ptr(r_trg), _ptr(r_src), _ptr(n_src), _ptr(M))
        assert rc == 0
        return M


class Reference(_Checker):
    """The reference's own CPU code (oracle/ref_oracle.cpp around /root/reference/include)."""

    def __init__(self, path=None):
        path = path or reference_path()
        if path is None:
            raise FileNotFoundError("oracle/_ref/libsctl_ref_*.so not built (needs /root/reference: make -C oracle ref)")
        lib = ctypes.CDLL(path)
        super().__init__(lib, "sctl_ref", "reference", path)
        lib.sctl_ref_eval.restype = ctypes.c_int
        lib.sctl_ref_eval.argtypes = [ctypes.c_int, ctypes.c_int, _I64, _I64, _P, _P, _P, _P, _P, ctypes.c_int, ctypes.c_int]
        lib.sctl_ref_eval_direct.restype = ctypes.c_int
        lib.sctl_ref_eval_direct.argtypes = [ctypes.c_int, ctypes.c_int, _I64, _I64, _P, _P, _P, _P, _P, ctypes.c_int]
        lib.sctl_ref_kernel_matrix.restype = ctypes.c_int
        lib.sctl_ref_kernel_matrix.argtypes = [ctypes.c_int, ctypes.c_int, _I64, _I64, _P, _P, _P, _P, ctypes.c_int]
        lib.sctl_ref_rsqrt.restype = ctypes.c_int
        lib.sctl_ref_rsqrt.argtypes = [ctypes.c_int, ctypes.c_int, _I64, _P, _P]
        lib.sctl_ref_set_threads.argtypes = [ctypes.c_int]

    def veclen(self, dtype) -> int:
        return self.lib.sctl_ref_veclen(_dt(np.zeros(0, dtype)))

    def max_threads(self) -> int:
        return self.lib.sctl_ref_max_threads()

    def set_threads(self, n: int) -> None:
        self.lib.sctl_ref_set_threads(n)

    def eval(self, kid, r_trg, r_src, n_src, v_src, v_trg=None, digits=-1, omp=True):
        d = self.kernel_dims(kid)
        Nt, Ns = r_trg.size // 3, r_src.size // 3
        if v_trg is None:
            v_trg = np.zeros(Nt * d["trg_dim"], dtype=r_trg.dtype)
        rc = self.lib.sctl_ref_eval(kid, _dt(r_trg), Nt, Ns, _ptr(r_trg), _ptr(r_src), _ptr(n_src), _ptr(v_src), _ptr(v_trg), digits, int(omp))
        assert rc == 0
        return v_trg

    def eval_direct(self, kid, r_trg, r_src, n_src, v_src, digits=-1):
        d = self.kernel_dims(kid)
        Nt, Ns = r_trg.size // 3, r_src.size // 3
        v_trg = np.full(Nt * d["trg_dim"], 7.0, dtype=r_trg.dtype)
        rc = self.lib.sctl_ref_eval_direct(kid, _dt(r_trg), Nt, Ns, _ptr(r_trg), _ptr(r_src), _ptr(n_src), _ptr(v_src), _ptr(v_trg), digits)
        assert rc == 0
        return v_trg

    def kernel_matrix(self, kid, r_trg, r_src, n_src, omp=True):
        d = self.kernel_dims(kid)
        Nt, Ns = r_trg.size // 3, r_src.size // 3
        M = np.zeros((Ns * d["src_dim"], Nt * d["trg_dim"]), dtype=r_trg.dtype)
        rc = self.lib.sctl_ref_kernel_matrix(kid, _dt(r_trg), Nt, Ns, _ptr(r_trg), _ptr(r_src), _ptr(n_src), _ptr(M), int(omp))
        assert rc == 0
        return M

    def rsqrt(self, x, digits=-1):
        y = np.empty_like(x)
        rc = self.lib.sctl_ref_rsqrt(_dt(x), digits, x.size, _ptr(x), _ptr(y))
        assert rc == 0
        return y


_port = None
_ref = None


def port() -> Port:
    global _port
    if _port is None:
        _port = Port()
    return _port


def reference() -> Reference | None:
    """The compiled reference, or None when oracle/_ref has not been built."""
    global _ref
    if _ref is None:
        try:
            _ref = Reference()
        except (FileNotFoundError, OSError):
            return None
    return _ref


def best() -> _Checker:
    """The strongest checker available: the compiled reference if present, else the C port."""
    return reference() or port()
