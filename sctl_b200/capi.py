"""ctypes binding of the C ABI declared in include/sctl_b200.h.

The library is built in-tree (sctl_b200/lib/libsctl_b200.so, see sctl_b200/csrc/Makefile). There is no
Python or CPU fallback for any compute entry point: if the shared library is missing, loading raises;
if no CUDA device is usable, every compute call raises SctlB200Error (code SCTL_B200_ERR_CUDA).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_int, c_int64, c_uint, c_void_p

# SCTL_B200_LIB selects another build of the SAME library (tuning variants under sctl_b200/lib/variants); there is no other backend
LIB_PATH = os.environ.get("SCTL_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libsctl_b200.so")

# enums of include/sctl_b200.h
F32, F64 = 0, 1
DEVICE_PTRS, OVERWRITE, ALL_GPUS = 1, 2, 4
ERR_KERNEL, ERR_DTYPE, ERR_ARG, ERR_CUDA, ERR_NOMEM, ERR_NCCL = 1, 2, 3, 4, 5, 6
NOALLOC_SCRATCH_BYTES, NOALLOC_RECORD_BYTES, NOALLOC_MAX_SRC = 32 << 20, 24 << 20, 314368

# every symbol include/sctl_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "sctl_b200_kernel_dims": (c_int, [c_int, POINTER(c_int), POINTER(c_int), POINTER(c_int), POINTER(c_int), POINTER(c_double), POINTER(c_int)]),
    "sctl_b200_kernel_name": (c_char_p, [c_int]),
    "sctl_b200_eval": (c_int, [c_int, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint]),
    "sctl_b200_eval_async": (c_int, [c_int, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint, c_void_p]),
    "sctl_b200_eval_noalloc": (c_int, [c_int, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint, c_void_p]),
    "sctl_b200_kernel_matrix": (c_int, [c_int, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint]),
    "sctl_b200_kernel_matrix_async": (c_int, [c_int, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint, c_void_p]),
    "sctl_b200_kernel_matrix_batch": (c_int, [c_int, c_int, c_int64, POINTER(c_int64), POINTER(c_int64), POINTER(c_int64), c_void_p, c_void_p, c_void_p,
                                              c_void_p, c_void_p, c_void_p, c_int64, c_int, c_uint]),
    "sctl_b200_buf_create": (c_int, [POINTER(c_void_p), c_int, c_int64]),
    "sctl_b200_buf_zero": (c_int, [c_void_p, c_int, c_int64]),
    "sctl_b200_buf_upload": (c_int, [c_void_p, c_void_p, c_int, c_int64]),
    "sctl_b200_buf_download": (c_int, [c_void_p, c_void_p, c_int, c_int64]),
    "sctl_b200_buf_destroy": (c_int, [c_void_p]),
    "sctl_b200_fmm_create": (c_int, [POINTER(c_void_p), c_int]),
    "sctl_b200_fmm_destroy": (c_int, [c_void_p]),
    "sctl_b200_fmm_set_accuracy": (c_int, [c_void_p, c_int]),
    "sctl_b200_fmm_add_src": (c_int, [c_void_p, c_char_p, c_int, c_int]),
    "sctl_b200_fmm_add_trg": (c_int, [c_void_p, c_char_p, c_int]),
    "sctl_b200_fmm_set_kernel_s2t": (c_int, [c_void_p, c_char_p, c_char_p, c_int]),
    "sctl_b200_fmm_delete_src": (c_int, [c_void_p, c_char_p]),
    "sctl_b200_fmm_delete_trg": (c_int, [c_void_p, c_char_p]),
    "sctl_b200_fmm_set_src_coord": (c_int, [c_void_p, c_char_p, c_int64, c_void_p, c_void_p, c_uint]),
    "sctl_b200_fmm_set_src_density": (c_int, [c_void_p, c_char_p, c_int64, c_void_p, c_uint]),
    "sctl_b200_fmm_set_trg_coord": (c_int, [c_void_p, c_char_p, c_int64, c_void_p, c_uint]),
    "sctl_b200_fmm_set_src_weights": (c_int, [c_void_p, c_char_p, c_int64, c_void_p, c_uint]),
    "sctl_b200_fmm_set_trg_normal": (c_int, [c_void_p, c_char_p, c_int64, c_void_p, c_uint]),
    "sctl_b200_fmm_trg_count": (c_int, [c_void_p, c_char_p, POINTER(c_int64), POINTER(c_int)]),
    "sctl_b200_fmm_eval_direct": (c_int, [c_void_p, c_char_p, c_void_p, c_int64, c_uint]),
    "sctl_b200_device_count": (c_int, [POINTER(c_int)]),
    "sctl_b200_set_device": (c_int, [c_int]),
    "sctl_b200_init": (c_int, [c_int, POINTER(c_int)]),
    "sctl_b200_finalize": (c_int, []),
    "sctl_b200_get_device": (c_int, [POINTER(c_int)]),
    "sctl_b200_sm_count": (c_int, [POINTER(c_int)]),
    "sctl_b200_partition": (c_int, [c_int64, c_int, c_int, POINTER(c_int64), POINTER(c_int64)]),
    "sctl_b200_comm_unique_id": (c_int, [c_void_p, c_int64]),
    "sctl_b200_comm_create": (c_int, [POINTER(c_void_p), c_int, c_int, c_void_p]),
    "sctl_b200_comm_destroy": (c_int, [c_void_p]),
    "sctl_b200_comm_info": (c_int, [c_void_p, POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "sctl_b200_nccl_version": (c_int, [POINTER(c_int)]),
    "sctl_b200_eval_sharded": (c_int, [c_void_p, c_int, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint,
                                       POINTER(c_int64)]),
    "sctl_b200_eval_sharded_async": (c_int, [c_void_p, c_int, c_int, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint,
                                             POINTER(c_int64), c_void_p]),
    "sctl_b200_fmm_set_comm": (c_int, [c_void_p, c_void_p]),
    "sctl_b200_launch_count": (c_int64, []),
    "sctl_b200_measure_peak": (c_int, [c_int, c_double, POINTER(c_double)]),
    "sctl_b200_last_timing": (c_int, [POINTER(c_double)]),
    "sctl_b200_newton_level": (c_int, [c_int, c_int]),
    "sctl_b200_test_rsqrt": (c_int, [c_int, c_int, c_int, c_int64, c_void_p, c_void_p]),
    "sctl_b200_last_error": (c_char_p, []),
    "sctl_b200_version": (c_char_p, []),
}


class SctlB200Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"sctl_b200 error {code}: {msg}")
        self.code = code


_lib = None


def lib() -> ctypes.CDLL:
    """Load libsctl_b200.so (once). Raises if it has not been built — there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(or make -C sctl_b200/csrc). sctl_b200 has no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise SctlB200Error(rc, lib().sctl_b200_last_error().decode())


def kernel_dims(kernel_id: int) -> dict:
    cd, nd, sd, td, fl = c_int(), c_int(), c_int(), c_int(), c_int()
    sc = c_double()
    check(lib().sctl_b200_kernel_dims(kernel_id, cd, nd, sd, td, sc, fl))
    return dict(coord_dim=cd.value, normal_dim=nd.value, src_dim=sd.value, trg_dim=td.value, scale=sc.value, flops=fl.value,
                name=lib().sctl_b200_kernel_name(kernel_id).decode())


def device_count() -> int:
    n = c_int()
    rc = lib().sctl_b200_device_count(n)
    return n.value if rc == 0 else 0


def launch_count() -> int:
    return int(lib().sctl_b200_launch_count())


def measure_peak(dtype: int, seconds: float = 0.5) -> float:
    tf = c_double()
    check(lib().sctl_b200_measure_peak(dtype, seconds, tf))
    return tf.value


def last_timing() -> dict:
    ms = (c_double * 5)()
    check(lib().sctl_b200_last_timing(ms))
    return dict(zip(("h2d_ms", "pack_ms", "eval_ms", "reduce_ms", "d2h_ms"), list(ms)))
