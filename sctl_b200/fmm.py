"""Host-side mirror of SCTL's ParticleFMM interface for the direct-summation path, backed by the device-resident
particle context of libsctl_b200.so (sctl_b200_fmm_* in include/sctl_b200.h).

Same method names, argument order and semantics as the reference
(/root/reference/include/sctl/fmm-wrapper.hpp:33-160, fmm-wrapper.txx:240-562):

    fmm = ParticleFMM(np.float64)
    fmm.SetKernels(ker_m2m, ker_m2l, ker_l2l)            # accepted and dim-checked; the far field (PVFMM) is not part of this library
    fmm.AddTrg("Velocity", ker_m2t, ker_l2t)
    fmm.AddSrc("DoubleLayer", ker_s2m, ker_s2l)
    fmm.SetKernelS2T("DoubleLayer", "Velocity", Stokes3D_DxU())
    fmm.SetTrgCoord("Velocity", Xt); fmm.SetSrcCoord("DoubleLayer", Xs, Xn); fmm.SetSrcDensity("DoubleLayer", F)
    U = fmm.Eval("Velocity")                              # == EvalDirect, as in a reference build without PVFMM (fmm-wrapper.txx:481-489)

What is different from the reference: coordinates, normals, densities and the packed source records live in HBM
between calls (the reference keeps host Vectors and re-sends them around an MPI ring on every call); only what changed
is re-uploaded / re-packed. numpy arrays are copied from the host, torch CUDA tensors device-to-device. There is no
CPU implementation here.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import capi
from .kernels import GenericKernel, _is_torch

_NP2DT = {np.dtype(np.float32): capi.F32, np.dtype(np.float64): capi.F64}


class ParticleFMM:
    def __init__(self, dtype=np.float64):
        self.dtype = np.dtype(dtype)
        if self.dtype not in _NP2DT:
            raise ValueError("Real must be float32 or float64")
        h = ctypes.c_void_p()
        capi.check(capi.lib().sctl_b200_fmm_create(ctypes.byref(h), _NP2DT[self.dtype]))
        self._h = h
        self._fmm_ker = None

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                capi.lib().sctl_b200_fmm_destroy(h)
            except Exception:
                pass

    def SetComm(self, comm) -> None:
        """multi-rank registry (the communicator handed to ParticleFMM(comm), fmm-wrapper.txx:204): `comm` is a
        sctl_b200.distributed.LibComm (or None to detach). Set* calls become collective; see sctl_b200_fmm_set_comm."""
        self._comm = comm  # keep the handle alive
        capi.check(capi.lib().sctl_b200_fmm_set_comm(self._h, comm.handle if comm is not None else None))

    def TrgCount(self, trg_name: str):
        """(targets this rank owns, values per target that Eval returns)"""
        nt, kd = ctypes.c_int64(), ctypes.c_int()
        capi.check(capi.lib().sctl_b200_fmm_trg_count(self._h, trg_name.encode(), ctypes.byref(nt), ctypes.byref(kd)))
        return nt.value, kd.value

    # -- registry (fmm-wrapper.txx:240-443) ---------------------------------------------------------------------
    def SetAccuracy(self, digits: int) -> None:
        capi.check(capi.lib().sctl_b200_fmm_set_accuracy(self._h, int(digits)))

    def SetKernels(self, ker_m2m: GenericKernel, ker_m2l: GenericKernel, ker_l2l: GenericKernel) -> None:
        # dimension checks of fmm-wrapper.txx:265-270; the translation kernels themselves are only used by the (external) FMM
        if ker_m2l.SrcDim() != ker_m2m.SrcDim() or ker_m2l.TrgDim() != ker_l2l.TrgDim():
            raise ValueError("M2L kernel dims must match M2M SrcDim and L2L TrgDim (fmm-wrapper.txx:269-270)")
        self._fmm_ker = (ker_m2m, ker_m2l, ker_l2l)

    def AddSrc(self, name: str, ker_s2m: GenericKernel, ker_s2l: GenericKernel) -> None:
        if ker_s2l.SrcDim() != ker_s2m.SrcDim() or ker_s2l.NormalDim() != ker_s2m.NormalDim():
            raise ValueError("S2M and S2L kernels must agree in SrcDim and NormalDim (fmm-wrapper.txx:312-313)")
        capi.check(capi.lib().sctl_b200_fmm_add_src(self._h, name.encode(), ker_s2m.SrcDim(), ker_s2m.NormalDim()))

    def AddTrg(self, name: str, ker_m2t: GenericKernel, ker_l2t: GenericKernel) -> None:
        if ker_m2t.TrgDim() != ker_l2t.TrgDim():
            raise ValueError("M2T and L2T kernels must agree in TrgDim (fmm-wrapper.txx:351)")
        capi.check(capi.lib().sctl_b200_fmm_add_trg(self._h, name.encode(), ker_l2t.TrgDim()))

    def SetKernelS2T(self, src_name: str, trg_name: str, ker_s2t: GenericKernel) -> None:
        capi.check(capi.lib().sctl_b200_fmm_set_kernel_s2t(self._h, src_name.encode(), trg_name.encode(), ker_s2t.KERNEL_ID))

    def DeleteSrc(self, name: str) -> None:
        capi.check(capi.lib().sctl_b200_fmm_delete_src(self._h, name.encode()))

    def DeleteTrg(self, name: str) -> None:
        capi.check(capi.lib().sctl_b200_fmm_delete_trg(self._h, name.encode()))

    # -- particle data (fmm-wrapper.txx:444-479) -----------------------------------------------------------------
    def _arg(self, a, out: bool = False):
        """-> (pointer, element count, flags, keep-alive); out=True: the array is written by the library"""
        if a is None:
            return None, 0, 0, None
        if _is_torch(a):
            import torch

            want = {np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64}[self.dtype]
            if not a.is_cuda or a.dtype != want:
                raise ValueError("torch tensors must be CUDA tensors of the context's Real type")
            if out and not a.is_contiguous():
                raise ValueError("the output tensor must be contiguous (a strided view would receive the result in a temporary copy)")
            a = a.contiguous()
            torch.cuda.current_stream(a.device).synchronize()  # the library copies on its own stream
            return (ctypes.c_void_p(a.data_ptr()) if a.numel() else None), a.numel(), capi.DEVICE_PTRS, a
        a = np.ascontiguousarray(a, dtype=self.dtype).reshape(-1)
        return (a.ctypes.data_as(ctypes.c_void_p) if a.size else None), a.size, 0, a

    def SetSrcCoord(self, name: str, src_coord, src_normal=None) -> None:
        px, nx, fl, _k1 = self._arg(src_coord)
        pn, nn, fl2, _k2 = self._arg(src_normal)
        if nx % 3:
            raise ValueError("coordinate arrays must hold 3 values per point")
        if nn and (fl2 != fl or nn != nx):
            raise ValueError("normals must match the coordinates (same count, same memory space)")
        capi.check(capi.lib().sctl_b200_fmm_set_src_coord(self._h, name.encode(), nx // 3, px, pn, fl))

    def SetSrcDensity(self, name: str, src_density) -> None:
        p, n, fl, _k = self._arg(src_density)
        capi.check(capi.lib().sctl_b200_fmm_set_src_density(self._h, name.encode(), n, p, fl))

    def SetTrgCoord(self, name: str, trg_coord) -> None:
        p, n, fl, _k = self._arg(trg_coord)
        if n % 3:
            raise ValueError("coordinate arrays must hold 3 values per point")
        capi.check(capi.lib().sctl_b200_fmm_set_trg_coord(self._h, name.encode(), n // 3, p, fl))

    # -- fused glue of BoundaryIntegralOp::ComputeFarField (extensions; not methods of the reference's ParticleFMM) ---------
    def SetSrcWeights(self, name: str, wts) -> None:
        """far-field quadrature weights: density * wts[s] is applied on the device when records are packed
        (the reference scales F_far on the host before SetSrcDensity, boundary_integral.txx:1036-1051). None removes them."""
        p, n, fl, _k = self._arg(wts)
        capi.check(capi.lib().sctl_b200_fmm_set_src_weights(self._h, name.encode(), n, p, fl))

    def SetTrgNormal(self, name: str, trg_normal) -> None:
        """Eval then returns the potential's component triplets dotted with the target normal, TrgDim/3 values per target
        (trg_normal_dot_prod_ branch, boundary_integral.txx:1060-1071). None removes them."""
        p, n, fl, _k = self._arg(trg_normal)
        capi.check(capi.lib().sctl_b200_fmm_set_trg_normal(self._h, name.encode(), n, p, fl))

    # -- evaluation (fmm-wrapper.txx:481-562) --------------------------------------------------------------------
    def EvalDirect(self, trg_name: str, U=None):
        """U (overwritten; created when None or wrongly sized — fmm-wrapper.txx:500-502) = sum over all sources with an S2T
        kernel for `trg_name`. A torch CUDA tensor U is filled on the device; otherwise a numpy array is returned."""
        nt, kd = ctypes.c_int64(), ctypes.c_int()
        capi.check(capi.lib().sctl_b200_fmm_trg_count(self._h, trg_name.encode(), ctypes.byref(nt), ctypes.byref(kd)))
        n = nt.value * kd.value
        if U is not None and _is_torch(U):
            if U.numel() != n:
                raise ValueError(f"U must hold Nt*TrgDim = {n} values")
            p, _, fl, _k = self._arg(U, out=True)
            capi.check(capi.lib().sctl_b200_fmm_eval_direct(self._h, trg_name.encode(), p, n, fl))
            return U
        if U is None or U.size != n or U.dtype != self.dtype or not U.flags.c_contiguous:
            U = np.zeros(n, dtype=self.dtype)
        capi.check(capi.lib().sctl_b200_fmm_eval_direct(self._h, trg_name.encode(), U.ctypes.data_as(ctypes.c_void_p) if n else None, n, 0))
        return U

    def Eval(self, trg_name: str, U=None):
        return self.EvalDirect(trg_name, U)
