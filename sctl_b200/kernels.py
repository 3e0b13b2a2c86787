"""Host-side mirror of SCTL's GenericKernel interface for the direct-summation path.

Same names, argument order and semantics as the reference's C++ API
(/root/reference/include/sctl/generic-kernel.hpp:59-147, kernel_functions.hpp:207-214):

    ker = Stokes3D_FxU()
    ker.Eval(v_trg, r_trg, r_src, n_src, v_src)      # generic-kernel.txx:76-189
    M = ker.KernelMatrix(r_trg, r_src, n_src)        # generic-kernel.txx:191-307
    ker.CoordDim(), ker.NormalDim(), ker.SrcDim(), ker.TrgDim(), ker.Name(), ker.FLOPS()

Arrays are flat AoS exactly like the reference's Vector<Real>. numpy arrays go through the
host-pointer C-ABI entry (copies in/out); torch CUDA tensors go through the device-pointer entry on torch's
current stream (no copies, no synchronisation). Everything forwards to libsctl_b200.so — there is no Python
or CPU implementation here.

Semantics kept from the reference:
  * Eval ACCUMULATES into v_trg when v_trg already has Nt*TrgDim elements, otherwise a zero-initialised
    result is created (generic-kernel.txx:98-101). The (possibly new) array is returned.
  * size mismatches are errors (the reference aborts, generic-kernel.txx:94-97): ValueError here.
  * n_src may be None / empty when NormalDim() == 0.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import capi

_NP2DT = {np.dtype(np.float32): capi.F32, np.dtype(np.float64): capi.F64}


def _is_torch(x) -> bool:
    return type(x).__module__.startswith("torch")


class GenericKernel:
    """Base of the eight kernel objects; subclasses only set KERNEL_ID."""

    KERNEL_ID = -1

    def __init__(self):
        self._dims = capi.kernel_dims(self.KERNEL_ID)

    # --- static metadata (generic-kernel.txx:22-36, kernel_functions.hpp Name/FLOPS/uKerScaleFactor)
    def CoordDim(self) -> int:
        return self._dims["coord_dim"]

    def NormalDim(self) -> int:
        return self._dims["normal_dim"]

    def SrcDim(self) -> int:
        return self._dims["src_dim"]

    def TrgDim(self) -> int:
        return self._dims["trg_dim"]

    def Name(self) -> str:
        return self._dims["name"]

    def FLOPS(self) -> int:
        return self._dims["flops"]

    def uKerScaleFactor(self) -> float:
        return self._dims["scale"]

    # --- helpers
    def _sizes(self, r_trg, r_src, n_src, v_src):
        def n(x):
            return 0 if x is None else int(x.numel() if _is_torch(x) else x.size)

        D, Nd, K0 = 3, self.NormalDim(), self.SrcDim()
        Nt, Ns = n(r_trg) // D, n(r_src) // D
        if n(r_trg) != Nt * D or n(r_src) != Ns * D:
            raise ValueError("coordinate arrays must hold 3 values per point (generic-kernel.txx:94-95)")
        if v_src is not None and n(v_src) != Ns * K0:
            raise ValueError(f"v_src must hold Ns*SrcDim = {Ns * K0} values (generic-kernel.txx:96)")
        if Nd and n(n_src) != Ns * Nd:
            raise ValueError(f"n_src must hold Ns*NormalDim = {Ns * Nd} values (generic-kernel.txx:97)")
        return Nt, Ns

    @staticmethod
    def _check_torch(a, like, numel, what):
        """every tensor handed to the C ABI as a raw pointer: same dtype and device as r_trg, contiguous, exact size"""
        if a.dtype != like.dtype or a.device != like.device or not a.is_contiguous() or a.numel() != numel:
            raise ValueError(f"{what}: expected a contiguous {like.dtype} tensor of {numel} values on {like.device}, got {a.dtype} "
                             f"{tuple(a.shape)} on {a.device}{'' if a.is_contiguous() else ' (non-contiguous)'}")

    @staticmethod
    def _check_np(a, dt, numel, what):
        if not isinstance(a, np.ndarray) or a.dtype != dt or not a.flags.c_contiguous or a.size != numel:
            raise ValueError(f"{what}: expected a C-contiguous {dt} array of {numel} values")

    @staticmethod
    def _np_ptr(a):
        return None if a is None or a.size == 0 else a.ctypes.data_as(ctypes.c_void_p)

    # --- Eval
    def Eval(self, v_trg, r_trg, r_src, n_src, v_src, digits: int = -1, overwrite: bool = False, all_gpus: bool = False):
        if _is_torch(r_trg):
            return self._eval_torch(v_trg, r_trg, r_src, n_src, v_src, digits, overwrite)
        dt = np.dtype(r_trg.dtype)
        if dt not in _NP2DT:
            raise ValueError("Real must be float32 or float64")
        c = lambda a: None if a is None else np.ascontiguousarray(a, dtype=dt).reshape(-1)
        r_trg, r_src, n_src, v_src = c(r_trg), c(r_src), c(n_src), c(v_src)
        Nt, Ns = self._sizes(r_trg, r_src, n_src, v_src)
        if v_trg is None or v_trg.size != Nt * self.TrgDim():
            v_trg = np.zeros(Nt * self.TrgDim(), dtype=dt)  # ReInit + SetZero
        elif v_trg.dtype != dt or not v_trg.flags.c_contiguous:
            raise ValueError("v_trg must be a contiguous array of the same Real type")
        flags = (capi.OVERWRITE if overwrite else 0) | (capi.ALL_GPUS if all_gpus else 0)
        capi.check(capi.lib().sctl_b200_eval(self.KERNEL_ID, _NP2DT[dt], Nt, Ns, self._np_ptr(r_trg), self._np_ptr(r_src),
                                             self._np_ptr(n_src) if self.NormalDim() else None, self._np_ptr(v_src),
                                             self._np_ptr(v_trg), digits, flags))
        return v_trg

    def _eval_torch(self, v_trg, r_trg, r_src, n_src, v_src, digits, overwrite):
        import torch

        if not r_trg.is_cuda:
            raise ValueError("torch tensors must be CUDA tensors (use numpy arrays for host data)")
        tdt = {torch.float32: capi.F32, torch.float64: capi.F64}.get(r_trg.dtype)
        if tdt is None:
            raise ValueError("Real must be float32 or float64")
        Nt, Ns = self._sizes(r_trg, r_src, n_src, v_src)
        self._check_torch(r_trg, r_trg, 3 * Nt, "r_trg")
        self._check_torch(r_src, r_trg, 3 * Ns, "r_src")
        self._check_torch(v_src, r_trg, self.SrcDim() * Ns, "v_src")
        if self.NormalDim():
            self._check_torch(n_src, r_trg, self.NormalDim() * Ns, "n_src")
        if v_trg is None or v_trg.numel() != Nt * self.TrgDim():
            v_trg = torch.zeros(Nt * self.TrgDim(), dtype=r_trg.dtype, device=r_trg.device)
        else:  # accumulated into IN PLACE: a wrong dtype / device / stride would be an out-of-bounds device write
            self._check_torch(v_trg, r_trg, Nt * self.TrgDim(), "v_trg")
        p = lambda a: None if a is None or a.numel() == 0 else ctypes.c_void_p(a.data_ptr())
        with torch.cuda.device(r_trg.device):
            stream = torch.cuda.current_stream().cuda_stream
            capi.check(capi.lib().sctl_b200_eval_async(self.KERNEL_ID, tdt, Nt, Ns, p(r_trg), p(r_src), p(n_src) if self.NormalDim() else None,
                                                       p(v_src), p(v_trg), digits, capi.OVERWRITE if overwrite else 0,
                                                       ctypes.c_void_p(stream)))
        return v_trg

    # --- KernelMatrix: returns the (Ns*SrcDim) x (Nt*TrgDim) row-major matrix
    def KernelMatrix(self, r_trg, r_src, n_src=None, digits: int = -1, out=None):
        K0, K1 = self.SrcDim(), self.TrgDim()
        if _is_torch(r_trg):
            import torch

            tdt = {torch.float32: capi.F32, torch.float64: capi.F64}.get(r_trg.dtype)
            if tdt is None or not r_trg.is_cuda:
                raise ValueError("torch tensors must be float32 / float64 CUDA tensors")
            Nt, Ns = self._sizes(r_trg, r_src, n_src, None)
            self._check_torch(r_trg, r_trg, 3 * Nt, "r_trg")
            self._check_torch(r_src, r_trg, 3 * Ns, "r_src")
            if self.NormalDim():
                self._check_torch(n_src, r_trg, self.NormalDim() * Ns, "n_src")
            if out is not None:
                self._check_torch(out, r_trg, Ns * K0 * Nt * K1, "out")
            M = out if out is not None else torch.empty((Ns * K0, Nt * K1), dtype=r_trg.dtype, device=r_trg.device)
            p = lambda a: None if a is None or a.numel() == 0 else ctypes.c_void_p(a.data_ptr())
            with torch.cuda.device(r_trg.device):
                stream = torch.cuda.current_stream().cuda_stream
                capi.check(capi.lib().sctl_b200_kernel_matrix_async(self.KERNEL_ID, tdt, Nt, Ns, p(r_trg), p(r_src),
                                                                    p(n_src) if self.NormalDim() else None, p(M), digits, 0,
                                                                    ctypes.c_void_p(stream)))
            return M
        dt = np.dtype(r_trg.dtype)
        c = lambda a: None if a is None else np.ascontiguousarray(a, dtype=dt).reshape(-1)
        r_trg, r_src, n_src = c(r_trg), c(r_src), c(n_src)
        if dt not in _NP2DT:
            raise ValueError("Real must be float32 or float64")
        Nt, Ns = self._sizes(r_trg, r_src, n_src, None)
        if out is not None:
            self._check_np(out, dt, Ns * K0 * Nt * K1, "out")
        M = out if out is not None else np.zeros((Ns * K0, Nt * K1), dtype=dt)
        capi.check(capi.lib().sctl_b200_kernel_matrix(self.KERNEL_ID, _NP2DT[dt], Nt, Ns, self._np_ptr(r_trg), self._np_ptr(r_src),
                                                      self._np_ptr(n_src) if self.NormalDim() else None, self._np_ptr(M), digits, 0))
        return M


    # --- batched small-block KernelMatrix (the per-element calls of BoundaryIntegralOp::SetupNear, boundary_integral.txx:946-1008)
    def KernelMatrixBatch(self, trg_dsp, src_dsp, r_trg, r_src, n_src=None, src_wts=None, n_trg=None, digits: int = -1, out=None, M_dsp=None):
        """All blocks in ONE launch. Block b = targets [trg_dsp[b], trg_dsp[b+1]) x sources [src_dsp[b], src_dsp[b+1]) (point
        displacements, like the reference's near_elem_dsp / elem_nds_dsp_far). Returns (M, M_dsp): the blocks, each row-major
        (Ns_b*SrcDim) x (Nt_b*K1out), concatenated in a flat array at element offsets M_dsp. src_wts scales row (s, k0) by
        src_wts[s]; n_trg contracts every triplet of target components with the target normal (K1out = TrgDim/3)."""
        K0, K1 = self.SrcDim(), self.TrgDim()
        trg_dsp = np.ascontiguousarray(trg_dsp, dtype=np.int64)
        src_dsp = np.ascontiguousarray(src_dsp, dtype=np.int64)
        nb = trg_dsp.size - 1
        if nb < 0 or src_dsp.size != nb + 1:
            raise ValueError("trg_dsp and src_dsp must both hold nbatch+1 displacements")
        k1out = K1 // 3 if n_trg is not None else K1
        if n_trg is not None and K1 % 3:
            raise ValueError("target-normal contraction needs TrgDim % 3 == 0")
        sizes = np.diff(src_dsp) * K0 * np.diff(trg_dsp) * k1out
        if M_dsp is None:
            M_dsp = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64) if nb else np.zeros(0, np.int64)
        M_dsp = np.ascontiguousarray(M_dsp, dtype=np.int64)
        M_len = int((M_dsp + sizes).max()) if nb else 0
        i64p = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))
        if _is_torch(r_trg):
            import torch

            tdt = {torch.float32: capi.F32, torch.float64: capi.F64}.get(r_trg.dtype)
            if tdt is None or not r_trg.is_cuda:
                raise ValueError("torch tensors must be float32 / float64 CUDA tensors")
            Nt_all, Ns_all = int(trg_dsp[-1]) if nb else 0, int(src_dsp[-1]) if nb else 0
            for a, n, what in ((r_trg, 3 * Nt_all, "r_trg"), (r_src, 3 * Ns_all, "r_src"), (n_src, self.NormalDim() * Ns_all, "n_src"), (src_wts, Ns_all, "src_wts"),
                               (n_trg, 3 * Nt_all, "n_trg")):
                if a is not None and (what != "n_src" or self.NormalDim()):
                    if a.dtype != r_trg.dtype or a.device != r_trg.device or not a.is_contiguous() or a.numel() < n:
                        raise ValueError(f"{what}: expected a contiguous {r_trg.dtype} tensor of at least {n} values on {r_trg.device}")
            if out is not None and (out.dtype != r_trg.dtype or out.device != r_trg.device or not out.is_contiguous() or out.numel() < M_len):
                raise ValueError(f"out: expected a contiguous {r_trg.dtype} tensor of at least {M_len} values on {r_trg.device}")
            M = out if out is not None else torch.zeros(M_len, dtype=r_trg.dtype, device=r_trg.device)
            p = lambda a: None if a is None or a.numel() == 0 else ctypes.c_void_p(a.data_ptr())
            with torch.cuda.device(r_trg.device):
                torch.cuda.current_stream().synchronize()  # the library runs the batch on its own stream
                capi.check(capi.lib().sctl_b200_kernel_matrix_batch(self.KERNEL_ID, tdt, nb, i64p(trg_dsp), i64p(src_dsp), i64p(M_dsp), p(r_trg), p(r_src),
                                                                    p(n_src) if self.NormalDim() else None, p(src_wts), p(n_trg), p(M), M.numel(), digits,
                                                                    capi.DEVICE_PTRS))
            return M, M_dsp
        dt = np.dtype(r_trg.dtype)
        c = lambda a: None if a is None else np.ascontiguousarray(a, dtype=dt).reshape(-1)
        r_trg, r_src, n_src, src_wts, n_trg = c(r_trg), c(r_src), c(n_src), c(src_wts), c(n_trg)
        Nt_all, Ns_all = int(trg_dsp[-1]) if nb else 0, int(src_dsp[-1]) if nb else 0
        for a, n, what in ((r_trg, 3 * Nt_all, "r_trg"), (r_src, 3 * Ns_all, "r_src"), (n_src if self.NormalDim() else None, self.NormalDim() * Ns_all, "n_src"),
                           (src_wts, Ns_all, "src_wts"), (n_trg, 3 * Nt_all, "n_trg")):
            if a is not None and a.size < n:
                raise ValueError(f"{what}: expected at least {n} values, got {a.size}")
        if out is not None and (not isinstance(out, np.ndarray) or out.dtype != dt or not out.flags.c_contiguous or out.size < M_len):
            raise ValueError(f"out: expected a C-contiguous {dt} array of at least {M_len} values")
        M = out if out is not None else np.zeros(M_len, dtype=dt)
        capi.check(capi.lib().sctl_b200_kernel_matrix_batch(self.KERNEL_ID, _NP2DT[dt], nb, i64p(trg_dsp), i64p(src_dsp), i64p(M_dsp), self._np_ptr(r_trg),
                                                            self._np_ptr(r_src), self._np_ptr(n_src) if self.NormalDim() else None, self._np_ptr(src_wts),
                                                            self._np_ptr(n_trg), self._np_ptr(M), M.size, digits, 0))
        return M, M_dsp


class Laplace3D_FxU(GenericKernel):
    KERNEL_ID = 0


class Laplace3D_DxU(GenericKernel):
    KERNEL_ID = 1


class Laplace3D_FxdU(GenericKernel):
    KERNEL_ID = 2


class Stokes3D_FxU(GenericKernel):
    KERNEL_ID = 3


class Stokes3D_DxU(GenericKernel):
    KERNEL_ID = 4


class Stokes3D_FxT(GenericKernel):
    KERNEL_ID = 5


class Stokes3D_FSxU(GenericKernel):
    KERNEL_ID = 6


class Stokes3D_FxUP(GenericKernel):
    KERNEL_ID = 7


KERNELS = [Laplace3D_FxU, Laplace3D_DxU, Laplace3D_FxdU, Stokes3D_FxU, Stokes3D_DxU, Stokes3D_FxT, Stokes3D_FSxU, Stokes3D_FxUP]
KERNEL_BY_NAME = {k.__name__: k for k in KERNELS}
