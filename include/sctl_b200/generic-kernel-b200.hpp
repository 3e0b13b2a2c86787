// generic-kernel-b200.hpp — reference-side binding of libsctl_b200.so (C ABI: include/sctl_b200.h).
//
// Include AFTER <sctl.hpp> (or after sctl/kernel_functions.hpp + generic-kernel.txx) and BEFORE the first use
// of a kernel object in the translation unit:
//
//     #include <sctl.hpp>
//     #include <sctl_b200/generic-kernel-b200.hpp>      // link with -lsctl_b200
//
// It adds explicit specialisations of the reference's member templates
//
//     GenericKernel<uKernel>::Eval<Real,enable_openmp>(…, Integer digits, ConstIterator<char> self)   [static]
//                                                        /root/reference/include/sctl/generic-kernel.txx:46-74
//     GenericKernel<uKernel>::Eval<Real,enable_openmp,digits>(v_trg,r_trg,r_src,n_src,v_src) const
//                                                        generic-kernel.txx:76-189
//     GenericKernel<uKernel>::KernelMatrix<Real,enable_openmp,digits>(M,Xt,Xs,Xn) const
//                                                        generic-kernel.txx:191-307
//
// for uKernel = the eight kernel_impl structs (kernel_functions.hpp:15-198), Real = float | double, both values
// of enable_openmp (ignored by the GPU path) and digits = -1, 0..15. Every caller of those templates — user
// code, ParticleFMM (which stores the ADDRESS of the static Eval, fmm-wrapper.txx:388-389, and calls it at
// :557,:623) and BoundaryIntegralOp (boundary_integral.txx:971,986) — then runs on the B200 without a source
// change in the reference. User-defined micro-kernels keep the reference's CPU templates (a host template
// cannot run on the device).
//
// Semantics preserved (each has a test in tests/cpp/shim_test.cpp):
//   * size checks abort through SCTL_ASSERT                                   generic-kernel.txx:94-97
//   * v_trg is ReInit'ed + zeroed if its size differs, else ACCUMULATED into    generic-kernel.txx:98-101
//   * KernelMatrix: M ReInit'ed if dims differ, every entry overwritten         generic-kernel.txx:198-201
//   * the FLOP profile counter still advances by Ns*Nt*FLOPS()                  generic-kernel.txx:188,202
//   * a failing library call (no GPU, CUDA error, …) is SCTL_ERROR -> abort, the reference's only error
//     convention (common.hpp:53-70). There is no CPU fallback.
//   * raw pointers are taken with &v[0], so Iterator<T> may be the checked class of SCTL_MEMDEBUG builds.
// Optional: define SCTL_B200_SHIM_FLAGS (e.g. SCTL_B200_ALL_GPUS) before including to shard the targets of
// every host-pointer Eval over all GPUs visible to the process.
#ifndef SCTL_B200_GENERIC_KERNEL_B200_HPP_
#define SCTL_B200_GENERIC_KERNEL_B200_HPP_

#include <string>

#include "../sctl_b200.h"
#include "sctl/common.hpp"
#include "sctl/generic-kernel.hpp"
#include "sctl/generic-kernel.txx"
#include "sctl/kernel_functions.hpp"
#include "sctl/matrix.hpp"
#include "sctl/profile.hpp"
#include "sctl/vector.hpp"

#ifndef SCTL_B200_SHIM_FLAGS
#define SCTL_B200_SHIM_FLAGS 0u
#endif

namespace sctl_b200_shim {

template <class uKernel> struct KernelId { static constexpr bool known = false; };  // kernel_functions.hpp struct -> id of sctl_b200.h
template <> struct KernelId<sctl::kernel_impl::Laplace3D_FxU> { static constexpr bool known = true; static constexpr int value = SCTL_B200_LAPLACE3D_FXU; };
template <> struct KernelId<sctl::kernel_impl::Laplace3D_DxU> { static constexpr bool known = true; static constexpr int value = SCTL_B200_LAPLACE3D_DXU; };
template <> struct KernelId<sctl::kernel_impl::Laplace3D_FxdU> { static constexpr bool known = true; static constexpr int value = SCTL_B200_LAPLACE3D_FXDU; };
template <> struct KernelId<sctl::kernel_impl::Stokes3D_FxU> { static constexpr bool known = true; static constexpr int value = SCTL_B200_STOKES3D_FXU; };
template <> struct KernelId<sctl::kernel_impl::Stokes3D_DxU> { static constexpr bool known = true; static constexpr int value = SCTL_B200_STOKES3D_DXU; };
template <> struct KernelId<sctl::kernel_impl::Stokes3D_FxT> { static constexpr bool known = true; static constexpr int value = SCTL_B200_STOKES3D_FXT; };
template <> struct KernelId<sctl::kernel_impl::Stokes3D_FSxU> { static constexpr bool known = true; static constexpr int value = SCTL_B200_STOKES3D_FSXU; };
template <> struct KernelId<sctl::kernel_impl::Stokes3D_FxUP> { static constexpr bool known = true; static constexpr int value = SCTL_B200_STOKES3D_FXUP; };

template <class Real> struct DType;
template <> struct DType<float> { static constexpr bool known = true; static constexpr int value = SCTL_B200_F32; };
template <> struct DType<double> { static constexpr bool known = true; static constexpr int value = SCTL_B200_F64; };

template <class Real> inline const Real* cptr(const sctl::Vector<Real>& v) { return v.Dim() ? &v[0] : nullptr; }
template <class Real> inline Real* mptr(sctl::Vector<Real>& v) { return v.Dim() ? &v[0] : nullptr; }

inline void check(int rc, const char* what) {
  if (rc != 0) {
    const std::string msg = std::string(what) + " failed (code " + std::to_string(rc) + "): " + sctl_b200_last_error();
    SCTL_ERROR(msg);
  }
}

template <class uKernel, class Real>
inline void Eval(sctl::Vector<Real>& v_trg, const sctl::Vector<Real>& r_trg, const sctl::Vector<Real>& r_src, const sctl::Vector<Real>& n_src,
                 const sctl::Vector<Real>& v_src, sctl::Integer digits) {
  using Ker = sctl::GenericKernel<uKernel>;
  constexpr sctl::Integer DIM = Ker::CoordDim(), K0 = Ker::SrcDim(), K1 = Ker::TrgDim(), ND = Ker::NormalDim();
  const sctl::Long Ns = r_src.Dim() / DIM;
  const sctl::Long Nt = r_trg.Dim() / DIM;
  SCTL_ASSERT(r_trg.Dim() == Nt * DIM);
  SCTL_ASSERT(r_src.Dim() == Ns * DIM);
  SCTL_ASSERT(v_src.Dim() == Ns * K0);
  SCTL_ASSERT(n_src.Dim() == Ns * ND || !ND);
  unsigned flags = SCTL_B200_SHIM_FLAGS;
  if (v_trg.Dim() != Nt * K1) {
    v_trg.ReInit(Nt * K1);
    v_trg.SetZero();
    flags |= SCTL_B200_OVERWRITE;  // freshly zeroed: no need to upload it
  }
  check(sctl_b200_eval(KernelId<uKernel>::value, DType<Real>::value, (int64_t)Nt, (int64_t)Ns, cptr(r_trg), cptr(r_src), ND ? cptr(n_src) : nullptr,
                       cptr(v_src), mptr(v_trg), (int)digits, flags),
        "sctl_b200_eval");
  sctl::Profile::IncrementCounter(sctl::ProfileCounter::FLOP, Ns * Nt * uKernel::FLOPS());
}

template <class uKernel, class Real>
inline void KernelMatrix(sctl::Matrix<Real>& M, const sctl::Vector<Real>& Xt, const sctl::Vector<Real>& Xs, const sctl::Vector<Real>& Xn,
                         sctl::Integer digits) {
  using Ker = sctl::GenericKernel<uKernel>;
  constexpr sctl::Integer DIM = Ker::CoordDim(), K0 = Ker::SrcDim(), K1 = Ker::TrgDim(), ND = Ker::NormalDim();
  const sctl::Long Ns = Xs.Dim() / DIM;
  const sctl::Long Nt = Xt.Dim() / DIM;
  if (M.Dim(0) != Ns * K0 || M.Dim(1) != Nt * K1) {
    M.ReInit(Ns * K0, Nt * K1);
    M.SetZero();
  }
  sctl::Profile::IncrementCounter(sctl::ProfileCounter::FLOP, Ns * Nt * uKernel::FLOPS());
  if (Ns == 0 || Nt == 0) return;
  check(sctl_b200_kernel_matrix(KernelId<uKernel>::value, DType<Real>::value, (int64_t)Nt, (int64_t)Ns, cptr(Xt), cptr(Xs), ND ? cptr(Xn) : nullptr,
                                &M[0][0], (int)digits, 0u),
        "sctl_b200_kernel_matrix");
}

}  // namespace sctl_b200_shim

namespace sctl {

// clang-format off
#define SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, DIG)                                                                                         \
  template <> template <> inline void GenericKernel<kernel_impl::UKER>::Eval<REAL, OMP, DIG>(                                                \
      Vector<REAL>& v_trg, const Vector<REAL>& r_trg, const Vector<REAL>& r_src, const Vector<REAL>& n_src, const Vector<REAL>& v_src) const { \
    sctl_b200_shim::Eval<kernel_impl::UKER, REAL>(v_trg, r_trg, r_src, n_src, v_src, DIG);                                                   \
  }                                                                                                                                         \
  template <> template <> inline void GenericKernel<kernel_impl::UKER>::KernelMatrix<REAL, OMP, DIG>(                                        \
      Matrix<REAL>& M, const Vector<REAL>& Xt, const Vector<REAL>& Xs, const Vector<REAL>& Xn) const {                                       \
    sctl_b200_shim::KernelMatrix<kernel_impl::UKER, REAL>(M, Xt, Xs, Xn, DIG);                                                               \
  }

#define SCTL_B200_SPEC_OMP(UKER, REAL, OMP)                                                                                                 \
  template <> template <> inline void GenericKernel<kernel_impl::UKER>::Eval<REAL, OMP>(                                                     \
      Vector<REAL>& v_trg, const Vector<REAL>& r_trg, const Vector<REAL>& r_src, const Vector<REAL>& n_src, const Vector<REAL>& v_src,       \
      Integer digits, ConstIterator<char> /*self: the kernels carry no state (ctx_ptr unused, generic-kernel.txx:20)*/) {                    \
    sctl_b200_shim::Eval<kernel_impl::UKER, REAL>(v_trg, r_trg, r_src, n_src, v_src, digits);                                                \
  }                                                                                                                                         \
  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, -1) SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 0)  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 1)              \
  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 2)  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 3)  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 4)              \
  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 5)  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 6)  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 7)              \
  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 8)  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 9)  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 10)             \
  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 11) SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 12) SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 13)             \
  SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 14) SCTL_B200_SPEC_DIGITS(UKER, REAL, OMP, 15)

#define SCTL_B200_SPEC_KERNEL(UKER)                                                \
  SCTL_B200_SPEC_OMP(UKER, double, true) SCTL_B200_SPEC_OMP(UKER, double, false)   \
  SCTL_B200_SPEC_OMP(UKER, float, true)  SCTL_B200_SPEC_OMP(UKER, float, false)

SCTL_B200_SPEC_KERNEL(Laplace3D_FxU)
SCTL_B200_SPEC_KERNEL(Laplace3D_DxU)
SCTL_B200_SPEC_KERNEL(Laplace3D_FxdU)
SCTL_B200_SPEC_KERNEL(Stokes3D_FxU)
SCTL_B200_SPEC_KERNEL(Stokes3D_DxU)
SCTL_B200_SPEC_KERNEL(Stokes3D_FxT)
SCTL_B200_SPEC_KERNEL(Stokes3D_FSxU)
SCTL_B200_SPEC_KERNEL(Stokes3D_FxUP)

#undef SCTL_B200_SPEC_KERNEL
#undef SCTL_B200_SPEC_OMP
#undef SCTL_B200_SPEC_DIGITS
// clang-format on

}  // namespace sctl

#endif  // SCTL_B200_GENERIC_KERNEL_B200_HPP_
