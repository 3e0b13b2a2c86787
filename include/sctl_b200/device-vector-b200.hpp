// device-vector-b200.hpp — device-resident Vector storage for the direct-summation path (north_star: "device-resident
// Vector buffers"), over the sctl_b200_buf_* C ABI (include/sctl_b200.h).
//
//     #include <sctl.hpp>
//     #include <sctl_b200/generic-kernel-b200.hpp>
//     #include <sctl_b200/device-vector-b200.hpp>
//
//     sctl_b200::DeviceVector<double> Xt(Xt_host), Xs(Xs_host), Xn, F(F_host), U;   // uploaded once
//     sctl_b200::Eval(ker, U, Xt, Xs, Xn, F);         // same argument order and semantics as GenericKernel::Eval
//     F = F_next;                                      // only the density crosses the bus
//     sctl_b200::Eval(ker, U, Xt, Xs, Xn, F);         // accumulates into U like the reference (generic-kernel.txx:98-101)
//     U.CopyTo(U_host);
//
// DeviceVector mirrors the part of sctl::Vector<Real> the path uses (vector.hpp:18-150: Dim, ReInit, SetZero, assignment
// from a Vector); it owns HBM on the device that was current when it was created and is move-only. Errors follow the
// reference's convention: SCTL_ERROR -> abort (common.hpp:53-70); there is no CPU fallback.
#ifndef SCTL_B200_DEVICE_VECTOR_B200_HPP_
#define SCTL_B200_DEVICE_VECTOR_B200_HPP_

#include <string>
#include <utility>

#include "generic-kernel-b200.hpp"

namespace sctl_b200 {

template <class Real> class DeviceVector {
 public:
  typedef Real ValueType;

  DeviceVector() = default;
  explicit DeviceVector(sctl::Long n) { ReInit(n); }
  explicit DeviceVector(const sctl::Vector<Real>& v) { *this = v; }
  DeviceVector(const DeviceVector&) = delete;
  DeviceVector& operator=(const DeviceVector&) = delete;
  DeviceVector(DeviceVector&& o) noexcept : p_(o.p_), n_(o.n_), cap_(o.cap_) { o.p_ = nullptr; o.n_ = o.cap_ = 0; }
  DeviceVector& operator=(DeviceVector&& o) noexcept {
    if (this != &o) {
      Free();
      p_ = o.p_; n_ = o.n_; cap_ = o.cap_;
      o.p_ = nullptr; o.n_ = o.cap_ = 0;
    }
    return *this;
  }
  ~DeviceVector() { Free(); }

  sctl::Long Dim() const { return n_; }

  // Like Vector::ReInit(n) without a data pointer: the contents are undefined afterwards (vector.txx:90-105).
  void ReInit(sctl::Long n) {
    if (n > cap_) {
      Free();
      void* p = nullptr;
      sctl_b200_shim::check(sctl_b200_buf_create(&p, sctl_b200_shim::DType<Real>::value, n), "sctl_b200_buf_create");
      p_ = static_cast<Real*>(p);
      cap_ = n;
    }
    n_ = n;
  }
  void SetZero() { sctl_b200_shim::check(sctl_b200_buf_zero(p_, sctl_b200_shim::DType<Real>::value, n_), "sctl_b200_buf_zero"); }

  // upload (resizes) / download (resizes the host Vector)
  DeviceVector& operator=(const sctl::Vector<Real>& v) {
    ReInit(v.Dim());
    if (n_) sctl_b200_shim::check(sctl_b200_buf_upload(p_, &v[0], sctl_b200_shim::DType<Real>::value, n_), "sctl_b200_buf_upload");
    return *this;
  }
  void CopyTo(sctl::Vector<Real>& v) const {
    if (v.Dim() != n_) v.ReInit(n_);
    if (n_) sctl_b200_shim::check(sctl_b200_buf_download(p_, &v[0], sctl_b200_shim::DType<Real>::value, n_), "sctl_b200_buf_download");
  }

  Real* data() { return p_; }              // DEVICE pointer
  const Real* data() const { return p_; }  // DEVICE pointer

 private:
  void Free() {
    if (p_) sctl_b200_buf_destroy(p_);
    p_ = nullptr;
    n_ = cap_ = 0;
  }
  Real* p_ = nullptr;
  sctl::Long n_ = 0, cap_ = 0;
};

// GenericKernel<uKernel>::Eval on device-resident vectors (generic-kernel.txx:76-189): same size checks, the same
// resize-and-zero / accumulate rule for v_trg, runtime `digits` (-1 = machine precision).
template <class uKernel, class Real>
void Eval(const sctl::GenericKernel<uKernel>&, DeviceVector<Real>& v_trg, const DeviceVector<Real>& r_trg, const DeviceVector<Real>& r_src,
          const DeviceVector<Real>& n_src, const DeviceVector<Real>& v_src, sctl::Integer digits = -1) {
  static_assert(sctl_b200_shim::KernelId<uKernel>::known, "only the kernel_impl micro-kernels of kernel_functions.hpp run on the device");
  constexpr sctl::Integer DIM = 3;
  constexpr sctl::Integer K0 = sctl::GenericKernel<uKernel>::SrcDim(), K1 = sctl::GenericKernel<uKernel>::TrgDim(), ND = sctl::GenericKernel<uKernel>::NormalDim();
  const sctl::Long Ns = r_src.Dim() / DIM, Nt = r_trg.Dim() / DIM;
  SCTL_ASSERT(r_trg.Dim() == Nt * DIM);  // generic-kernel.txx:94-97
  SCTL_ASSERT(r_src.Dim() == Ns * DIM);
  SCTL_ASSERT(v_src.Dim() == Ns * K0);
  SCTL_ASSERT(n_src.Dim() == Ns * ND || !ND);
  unsigned flags = SCTL_B200_DEVICE_PTRS;
  if (v_trg.Dim() != Nt * K1) {  // :98-101
    v_trg.ReInit(Nt * K1);
    flags |= SCTL_B200_OVERWRITE;
  }
  sctl_b200_shim::check(sctl_b200_eval(sctl_b200_shim::KernelId<uKernel>::value, sctl_b200_shim::DType<Real>::value, Nt, Ns, r_trg.data(), r_src.data(),
                                       ND ? n_src.data() : nullptr, v_src.data(), v_trg.data(), digits, flags), "sctl_b200_eval");
  sctl::Profile::IncrementCounter(sctl::ProfileCounter::FLOP, Ns * Nt * uKernel::FLOPS());  // :188
}

// GenericKernel<uKernel>::KernelMatrix on device-resident vectors (generic-kernel.txx:191-307): M holds the row-major
// (Ns*SrcDim) x (Nt*TrgDim) matrix, every entry overwritten.
template <class uKernel, class Real>
void KernelMatrix(const sctl::GenericKernel<uKernel>&, DeviceVector<Real>& M, const DeviceVector<Real>& r_trg, const DeviceVector<Real>& r_src,
                  const DeviceVector<Real>& n_src, sctl::Integer digits = -1) {
  static_assert(sctl_b200_shim::KernelId<uKernel>::known, "only the kernel_impl micro-kernels of kernel_functions.hpp run on the device");
  constexpr sctl::Integer DIM = 3;
  constexpr sctl::Integer K0 = sctl::GenericKernel<uKernel>::SrcDim(), K1 = sctl::GenericKernel<uKernel>::TrgDim(), ND = sctl::GenericKernel<uKernel>::NormalDim();
  const sctl::Long Ns = r_src.Dim() / DIM, Nt = r_trg.Dim() / DIM;
  SCTL_ASSERT(r_trg.Dim() == Nt * DIM);
  SCTL_ASSERT(r_src.Dim() == Ns * DIM);
  SCTL_ASSERT(n_src.Dim() == Ns * ND || !ND);
  if (M.Dim() != Ns * K0 * Nt * K1) M.ReInit(Ns * K0 * Nt * K1);
  sctl_b200_shim::check(sctl_b200_kernel_matrix(sctl_b200_shim::KernelId<uKernel>::value, sctl_b200_shim::DType<Real>::value, Nt, Ns, r_trg.data(), r_src.data(),
                                                ND ? n_src.data() : nullptr, M.data(), digits, SCTL_B200_DEVICE_PTRS), "sctl_b200_kernel_matrix");
  sctl::Profile::IncrementCounter(sctl::ProfileCounter::FLOP, Ns * Nt * uKernel::FLOPS());
}

}  // namespace sctl_b200

#endif  // SCTL_B200_DEVICE_VECTOR_B200_HPP_
