"""sctl_b200 — B200-native direct particle-kernel summation behind SCTL's GenericKernel interface.

Only the hot path of the reference is here: GenericKernel::Eval / KernelMatrix for the
Laplace3D / Stokes3D kernels and the per-rank body of ParticleFMM::EvalDirect. The compute lives in
sctl_b200/lib/libsctl_b200.so (hand-written sm_100a CUDA behind the C ABI of include/sctl_b200.h); this
package is the Python mirror of the reference's C++ interface used by tests, bench.py and torchrun drivers.
"""
from . import capi  # noqa: F401
from .fmm import ParticleFMM  # noqa: F401
from .kernels import (  # noqa: F401
    KERNEL_BY_NAME,
    KERNELS,
    GenericKernel,
    Laplace3D_DxU,
    Laplace3D_FxdU,
    Laplace3D_FxU,
    Stokes3D_DxU,
    Stokes3D_FSxU,
    Stokes3D_FxT,
    Stokes3D_FxU,
    Stokes3D_FxUP,
)

__all__ = ["capi", "ParticleFMM", "GenericKernel", "KERNELS", "KERNEL_BY_NAME", "Laplace3D_FxU", "Laplace3D_DxU", "Laplace3D_FxdU", "Stokes3D_FxU",
           "Stokes3D_DxU", "Stokes3D_FxT", "Stokes3D_FSxU", "Stokes3D_FxUP"]
