// TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT PATH.
//
// Thin extern "C" wrapper around the UNMODIFIED reference (SCTL) headers under
// /root/reference/include.  It is compiled by oracle/Makefile into
// oracle/_ref/libsctl_ref_<isa>.so (git-ignored, travels to the GPU box as a built
// file).  It contains none of the reference's source: it only #includes it where
// it lies and instantiates the reference's own templates:
//
//   GenericKernel<uKernel>::Eval (static, runtime digits)   generic-kernel.txx:46-74
//   GenericKernel<uKernel>::Eval<Real,omp,digits>            generic-kernel.txx:76-189
//   GenericKernel<uKernel>::KernelMatrix                     generic-kernel.txx:191-307
//   ParticleFMM<Real,3>::EvalDirect                          fmm-wrapper.txx:490-562
//
// Used by: tests/ (parity checker), scripts/make_golden.py (fixture generator),
// bench.py's cpu_baseline leg and `bench.py --impl reference`.
#include <sctl.hpp>

#include <cstdint>
#include <cstring>
#include <string>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

using sctl::Integer;
using sctl::Long;

template <class Real> sctl::Vector<Real> view(const void* p, Long n) {
  // non-owning view (vector.txx:26-40); n == 0 gives an empty Vector like the
  // reference's callers pass for kernels without normals.
  if (n == 0 || p == nullptr) return sctl::Vector<Real>();
  return sctl::Vector<Real>(n, sctl::Ptr2Itr<Real>((Real*)p, n), false);
}

template <class Ker, class Real>
int eval_t(Long Nt, Long Ns, const void* r_trg, const void* r_src, const void* n_src, const void* v_src,
           void* v_trg, int digits, int use_omp) {
  Ker ker;
  const sctl::Vector<Real> Xt = view<Real>(r_trg, Nt * Ker::CoordDim());
  const sctl::Vector<Real> Xs = view<Real>(r_src, Ns * Ker::CoordDim());
  const sctl::Vector<Real> Xn = view<Real>(n_src, Ns * Ker::NormalDim());
  const sctl::Vector<Real> F = view<Real>(v_src, Ns * Ker::SrcDim());
  sctl::Vector<Real> U = view<Real>(v_trg, Nt * Ker::TrgDim());
  if (Nt == 0) return 0;
  // the type-erased static entry point stored by ParticleFMM (fmm-wrapper.txx:388-389)
  const auto self = (sctl::ConstIterator<char>)sctl::Ptr2ConstItr<Ker>(&ker, 1);
  if (use_omp)
    Ker::template Eval<Real, true>(U, Xt, Xs, Xn, F, (Integer)digits, self);
  else
    Ker::template Eval<Real, false>(U, Xt, Xs, Xn, F, (Integer)digits, self);
  return 0;
}

template <class Ker, class Real>
int kmat_t(Long Nt, Long Ns, const void* r_trg, const void* r_src, const void* n_src, void* M_, int use_omp) {
  Ker ker;
  const sctl::Vector<Real> Xt = view<Real>(r_trg, Nt * Ker::CoordDim());
  const sctl::Vector<Real> Xs = view<Real>(r_src, Ns * Ker::CoordDim());
  const sctl::Vector<Real> Xn = view<Real>(n_src, Ns * Ker::NormalDim());
  if (Nt == 0 || Ns == 0) return 0;
  sctl::Matrix<Real> M(Ns * Ker::SrcDim(), Nt * Ker::TrgDim(), sctl::Ptr2Itr<Real>((Real*)M_, Ns * Ker::SrcDim() * Nt * Ker::TrgDim()), false);
  if (use_omp)
    ker.template KernelMatrix<Real, true>(M, Xt, Xs, Xn);
  else
    ker.template KernelMatrix<Real, false>(M, Xt, Xs, Xn);
  return 0;
}

template <class Ker> int dims_t(int* coord, int* normal, int* src, int* trg, double* scale, int* flops, char* name, int name_len) {
  if (coord) *coord = (int)Ker::CoordDim();
  if (normal) *normal = (int)Ker::NormalDim();
  if (src) *src = (int)Ker::SrcDim();
  if (trg) *trg = (int)Ker::TrgDim();
  if (scale) *scale = Ker::template uKerScaleFactor<double>();
  if (flops) *flops = (int)Ker::FLOPS();
  if (name && name_len > 0) {
    std::strncpy(name, Ker::Name().c_str(), name_len - 1);
    name[name_len - 1] = 0;
  }
  return 0;
}

// ParticleFMM::EvalDirect with one source type and one target type (np = 1 loopback Comm).
template <class Ker, class Real>
int eval_direct_t(Long Nt, Long Ns, const void* r_trg, const void* r_src, const void* n_src, const void* v_src, void* v_trg, int digits) {
  Ker ker;
  sctl::ParticleFMM<Real, 3> fmm(sctl::Comm::Self());
  fmm.SetAccuracy(digits);
  fmm.AddSrc("S", ker, ker);
  fmm.AddTrg("T", ker, ker);
  fmm.SetKernelS2T("S", "T", ker);
  fmm.SetTrgCoord("T", view<Real>(r_trg, Nt * 3));
  fmm.SetSrcCoord("S", view<Real>(r_src, Ns * 3), view<Real>(n_src, Ns * Ker::NormalDim()));
  fmm.SetSrcDensity("S", view<Real>(v_src, Ns * Ker::SrcDim()));
  sctl::Vector<Real> U;
  fmm.EvalDirect(U, "T");
  if (U.Dim() != Nt * Ker::TrgDim()) return 3;
  std::memcpy(v_trg, &U[0], sizeof(Real) * U.Dim());
  return 0;
}

#define DISPATCH_KERNEL(FN, ...)                                        \
  switch (kernel_id) {                                                  \
    case 0: return FN<sctl::Laplace3D_FxU, Real>(__VA_ARGS__);          \
    case 1: return FN<sctl::Laplace3D_DxU, Real>(__VA_ARGS__);          \
    case 2: return FN<sctl::Laplace3D_FxdU, Real>(__VA_ARGS__);         \
    case 3: return FN<sctl::Stokes3D_FxU, Real>(__VA_ARGS__);           \
    case 4: return FN<sctl::Stokes3D_DxU, Real>(__VA_ARGS__);           \
    case 5: return FN<sctl::Stokes3D_FxT, Real>(__VA_ARGS__);           \
    case 6: return FN<sctl::Stokes3D_FSxU, Real>(__VA_ARGS__);          \
    case 7: return FN<sctl::Stokes3D_FxUP, Real>(__VA_ARGS__);          \
    default: return 1;                                                  \
  }

template <class Real> int eval_k(int kernel_id, Long Nt, Long Ns, const void* a, const void* b, const void* c, const void* d, void* e, int digits, int omp) {
  DISPATCH_KERNEL(eval_t, Nt, Ns, a, b, c, d, e, digits, omp)
}
template <class Real> int kmat_k(int kernel_id, Long Nt, Long Ns, const void* a, const void* b, const void* c, void* M, int omp) {
  DISPATCH_KERNEL(kmat_t, Nt, Ns, a, b, c, M, omp)
}
template <class Real> int eval_direct_k(int kernel_id, Long Nt, Long Ns, const void* a, const void* b, const void* c, const void* d, void* e, int digits) {
  DISPATCH_KERNEL(eval_direct_t, Nt, Ns, a, b, c, d, e, digits)
}

}  // namespace

extern "C" {

// dtype: 0 = float, 1 = double.  v_trg is accumulated into (generic-kernel.txx:98-101,182-186).
int sctl_ref_eval(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                  const void* v_src, void* v_trg, int digits, int use_omp) {
  if (dtype == 0) return eval_k<float>(kernel_id, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg, digits, use_omp);
  if (dtype == 1) return eval_k<double>(kernel_id, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg, digits, use_omp);
  return 2;
}

int sctl_ref_kernel_matrix(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                           void* M, int use_omp) {
  if (dtype == 0) return kmat_k<float>(kernel_id, Nt, Ns, r_trg, r_src, n_src, M, use_omp);
  if (dtype == 1) return kmat_k<double>(kernel_id, Nt, Ns, r_trg, r_src, n_src, M, use_omp);
  return 2;
}

int sctl_ref_eval_direct(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                         const void* v_src, void* v_trg, int digits) {
  if (dtype == 0) return eval_direct_k<float>(kernel_id, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg, digits);
  if (dtype == 1) return eval_direct_k<double>(kernel_id, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg, digits);
  return 2;
}

int sctl_ref_kernel_dims(int kernel_id, int* coord, int* normal, int* src, int* trg, double* scale, int* flops, char* name, int name_len) {
  switch (kernel_id) {
    case 0: return dims_t<sctl::Laplace3D_FxU>(coord, normal, src, trg, scale, flops, name, name_len);
    case 1: return dims_t<sctl::Laplace3D_DxU>(coord, normal, src, trg, scale, flops, name, name_len);
    case 2: return dims_t<sctl::Laplace3D_FxdU>(coord, normal, src, trg, scale, flops, name, name_len);
    case 3: return dims_t<sctl::Stokes3D_FxU>(coord, normal, src, trg, scale, flops, name, name_len);
    case 4: return dims_t<sctl::Stokes3D_DxU>(coord, normal, src, trg, scale, flops, name, name_len);
    case 5: return dims_t<sctl::Stokes3D_FxT>(coord, normal, src, trg, scale, flops, name, name_len);
    case 6: return dims_t<sctl::Stokes3D_FSxU>(coord, normal, src, trg, scale, flops, name, name_len);
    case 7: return dims_t<sctl::Stokes3D_FxUP>(coord, normal, src, trg, scale, flops, name, name_len);
    default: return 1;
  }
}

// SIMD width the reference selected for this build (vec.txx:26-40)
int sctl_ref_veclen(int dtype) { return dtype == 0 ? (int)sctl::DefaultVecLen<float>() : (int)sctl::DefaultVecLen<double>(); }

int sctl_ref_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void sctl_ref_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

// approx_rsqrt<digits>(x, x>0) on VecLen lanes (vec.txx:361-364): pins the seed+Newton path for tests.
int sctl_ref_rsqrt(int dtype, int digits, int64_t n, const void* x, void* y) {
  if (dtype == 1) {
    using V = sctl::Vec<double, sctl::DefaultVecLen<double>()>;
    constexpr Long L = sctl::DefaultVecLen<double>();
    const double* xi = (const double*)x;
    double* yo = (double*)y;
    for (Long i = 0; i < n; i += L) {
      alignas(64) double a[L] = {0}, b[L];
      for (Long j = 0; j < L && i + j < n; j++) a[j] = xi[i + j];
      V va = V::LoadAligned(a), vb;
      if (digits == 4) vb = sctl::approx_rsqrt<4>(va, va > V::Zero());
      else if (digits == 7) vb = sctl::approx_rsqrt<7>(va, va > V::Zero());
      else vb = sctl::approx_rsqrt<-1>(va, va > V::Zero());
      vb.StoreAligned(b);
      for (Long j = 0; j < L && i + j < n; j++) yo[i + j] = b[j];
    }
    return 0;
  }
  if (dtype == 0) {
    using V = sctl::Vec<float, sctl::DefaultVecLen<float>()>;
    constexpr Long L = sctl::DefaultVecLen<float>();
    const float* xi = (const float*)x;
    float* yo = (float*)y;
    for (Long i = 0; i < n; i += L) {
      alignas(64) float a[L] = {0}, b[L];
      for (Long j = 0; j < L && i + j < n; j++) a[j] = xi[i + j];
      V va = V::LoadAligned(a), vb;
      if (digits == 4) vb = sctl::approx_rsqrt<4>(va, va > V::Zero());
      else if (digits == 7) vb = sctl::approx_rsqrt<7>(va, va > V::Zero());
      else vb = sctl::approx_rsqrt<-1>(va, va > V::Zero());
      vb.StoreAligned(b);
      for (Long j = 0; j < L && i + j < n; j++) yo[i + j] = b[j];
    }
    return 0;
  }
  return 2;
}

}  // extern "C"
