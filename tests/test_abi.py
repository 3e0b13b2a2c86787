"""CPU: the C-ABI library loads, exports EVERY symbol include/sctl_b200.h declares, answers metadata queries without
a GPU, validates arguments, and FAILS LOUDLY (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest

from sctl_b200 import KERNELS, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "sctl_b200.h")).read()
    return sorted(set(re.findall(r"SCTL_B200_API\s+[\w\s\*]+?\b(sctl_b200_\w+)\s*\(", txt)))


def test_header_and_binding_agree():
    decl = declared_symbols()
    assert len(decl) >= 19
    assert sorted(capi.SYMBOLS) == decl, "capi.SYMBOLS must bind exactly what the header declares"


def test_library_exports_every_declared_symbol():
    L = ctypes.CDLL(capi.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(L, name), name
    assert capi.lib().sctl_b200_version().decode().startswith("sctl_b200")


def test_kernel_metadata_no_gpu_needed():
    names = ["Laplace3D-FxU", "Laplace3D-DxU", "Laplace3D-FxdU", "Stokes3D-FxU", "Stokes3D-DxU", "Stokes3D-FxT", "Stokes3D-FSxU", "Stokes3D-FxUP"]
    dims = [(0, 1, 1, 6), (3, 1, 1, 14), (0, 1, 3, 11), (0, 3, 3, 23), (3, 3, 3, 26), (0, 3, 9, 39), (0, 4, 3, 26), (0, 3, 4, 26)]
    scales = [1 / (4 * np.pi), 1 / (4 * np.pi), -1 / (4 * np.pi), 1 / (8 * np.pi), 3 / (4 * np.pi), -3 / (4 * np.pi), 1 / (8 * np.pi), 1 / (8 * np.pi)]
    for kid in range(8):
        k = KERNELS[kid]()
        assert k.Name() == names[kid] and k.CoordDim() == 3
        assert (k.NormalDim(), k.SrcDim(), k.TrgDim(), k.FLOPS()) == dims[kid]
        assert k.uKerScaleFactor() == pytest.approx(scales[kid], rel=1e-15)
    assert capi.lib().sctl_b200_kernel_name(99) is None
    assert capi.lib().sctl_b200_kernel_dims(8, None, None, None, None, None, None) == capi.ERR_KERNEL


def test_argument_errors():
    L = capi.lib()
    x = np.zeros(3)
    p = x.ctypes.data_as(ctypes.c_void_p)
    assert L.sctl_b200_eval(42, 1, 1, 1, p, p, None, p, p, -1, 0) == capi.ERR_KERNEL
    assert L.sctl_b200_eval(0, 7, 1, 1, p, p, None, p, p, -1, 0) == capi.ERR_DTYPE
    assert L.sctl_b200_eval(0, 1, -1, 1, p, p, None, p, p, -1, 0) == capi.ERR_ARG
    assert L.sctl_b200_eval(0, 1, 1, 1, None, p, None, p, p, -1, 0) == capi.ERR_ARG
    assert L.sctl_b200_eval(1, 1, 1, 1, p, p, None, p, p, -1, 0) == capi.ERR_ARG  # DxU needs normals
    assert L.sctl_b200_eval(0, 1, 1, 1, p, p, None, p, p, -1, capi.DEVICE_PTRS | capi.ALL_GPUS) == capi.ERR_ARG
    assert b"" != L.sctl_b200_last_error()
    # host-side mirror: size mismatches are errors like the reference's asserts (generic-kernel.txx:94-97)
    with pytest.raises(ValueError):
        KERNELS[3]().Eval(None, np.zeros(6), np.zeros(6), None, np.zeros(5))
    with pytest.raises(ValueError):
        KERNELS[4]().Eval(None, np.zeros(6), np.zeros(6), None, np.zeros(6))
    with pytest.raises(ValueError):
        KERNELS[0]().Eval(None, np.zeros(4), np.zeros(6), None, np.zeros(2))


def test_output_arrays_are_validated_before_the_abi_sees_them():
    """an `out=` / v_trg handed in by the caller is written through a raw pointer: wrong dtype, size or stride must be a ValueError,
    never an out-of-bounds write (numpy side here; the torch side is tests/test_gpu_edge_cases.py::test_torch_outputs_are_validated)"""
    k = KERNELS[3]()
    Xt, Xs = np.zeros(6), np.ones(9)
    for bad in (np.zeros((9, 6), np.float32), np.zeros(53), np.zeros((9, 12))[:, ::2]):
        with pytest.raises(ValueError):
            k.KernelMatrix(Xt, Xs, out=bad)
    with pytest.raises(ValueError):
        k.Eval(np.zeros(6, np.float32), Xt, Xs, None, np.zeros(9))  # right size, wrong dtype
    with pytest.raises(ValueError):
        k.Eval(np.zeros(12)[::2], Xt, Xs, None, np.zeros(9))  # right size, strided
    dsp = np.array([0, 2], np.int64)
    with pytest.raises(ValueError):
        k.KernelMatrixBatch(dsp, np.array([0, 3], np.int64), Xt, Xs, out=np.zeros(10))  # needs 3*3*2*3 = 54 values


def test_no_cpu_fallback_without_device():
    if capi.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.SctlB200Error) as e:
        KERNELS[0]().Eval(None, np.zeros(3), np.ones(3), None, np.ones(1))
    assert e.value.code == capi.ERR_CUDA and "no CPU fallback" in str(e.value)
    with pytest.raises(capi.SctlB200Error):
        KERNELS[0]().KernelMatrix(np.zeros(3), np.ones(3))


def test_product_never_imports_oracle():
    """the product path must not route through the oracle: no file under sctl_b200/ or include/ mentions it"""
    for base in ("sctl_b200", "include"):
        for dp, _, fs in os.walk(os.path.join(ROOT, base)):
            for f in fs:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                    txt = open(os.path.join(dp, f)).read()
                    for needle in ("import oracle", "from oracle", "liboracle", "oracle/", "oracle_port", "sctl_ref"):
                        assert needle not in txt, (os.path.join(dp, f), needle)
