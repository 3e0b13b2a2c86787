// TEST-ONLY element list for the reference's BoundaryIntegralOp (boundary_integral.hpp:64-220): the reference ships the
// abstract ElementListBase but no implementation (its users bring one, e.g. CSBQ's SlenderElemList), so the call sites
// BoundaryIntegralOp::SetupNear / ComputeFarField / ComputePotential (boundary_integral.txx:608-614, 816-1077) cannot run
// from the reference tree alone. This is the smallest honest implementation: a sphere cut into n_theta x n_phi patches in
// spherical coordinates, q x q tensor Gauss-Legendre nodes per patch.
//   * surface nodes == far-field quadrature nodes (GetFarFieldDensity / FarFieldDensityOperatorTranspose keep their
//     defaults: identity), weights = GL weights x Jacobian R^2 sin(theta), normals = outward radial;
//   * dist_far = near_factor x patch diameter: targets closer than that to a node get a near correction;
//   * NearInterac / SelfInterac: an UPSAMPLED tensor GL rule (m = upsample*q per direction) on the same patch with the density
//     interpolated from the q x q nodes (tensor Lagrange) - smooth-quadrature accuracy only, i.e. crude for the singular self
//     interaction. That is deliberate and harmless here: both arms of the test (reference CPU templates / the B200 library)
//     share it and the test compares ARMS. Like a real element list (which brings its own special quadrature), it evaluates the
//     kernel on the HOST by default (HostKernelOf<Kernel>: the same micro-kernel under a type the shim does not match), so the
//     library calls of the GPU arm are exactly the reference's own call sites. near_on_device() = true routes these one-target
//     KernelMatrix calls through `ker` as well: ~2e5 tiny concurrent calls from the reference's omp loop, a stress test of the
//     shim's small-call path.
#ifndef SCTL_B200_TESTS_SPHERE_PATCH_LIST_HPP_
#define SCTL_B200_TESTS_SPHERE_PATCH_LIST_HPP_

#include <cmath>
#include <vector>

namespace sphere_patch_detail {
template <class K> struct HostCopy : K {};  // same arithmetic, distinct type: never matched by the shim's specialisations
template <class Kernel> struct HostKernelOf;
template <class uKernel> struct HostKernelOf<sctl::GenericKernel<uKernel>> { using type = sctl::GenericKernel<HostCopy<uKernel>>; };
inline bool& near_on_device() {
  static bool flag = false;
  return flag;
}
}  // namespace sphere_patch_detail

template <class Real> class SpherePatchList : public sctl::ElementListBase<Real> {
  using Long = sctl::Long;

 public:
  SpherePatchList() : R_(1), nth_(0), nph_(0), q_(0), m_(0), near_factor_(1) { c_[0] = c_[1] = c_[2] = 0; }
  SpherePatchList(Real R, Real cx, Real cy, Real cz, Long n_theta, Long n_phi, int q, int upsample, Real near_factor)
      : R_(R), nth_(n_theta), nph_(n_phi), q_(q), m_(q * upsample), near_factor_(near_factor) {
    c_[0] = cx, c_[1] = cy, c_[2] = cz;
    gauss_legendre(q_, xq_, wq_);
    gauss_legendre(m_, xm_, wm_);
    P_.assign((size_t)m_ * q_, 0);  // P[c][a] = Lagrange basis a (on the q nodes) at upsampled node c
    for (int c = 0; c < m_; c++)
      for (int a = 0; a < q_; a++) {
        Real l = 1;
        for (int b = 0; b < q_; b++)
          if (b != a) l *= (xm_[c] - xq_[b]) / (xq_[a] - xq_[b]);
        P_[(size_t)c * q_ + a] = l;
      }
  }

  Long Size() const override { return nth_ * nph_; }
  Long NodesPerElem() const { return (Long)q_ * q_; }

  void GetNodeCoord(sctl::Vector<Real>* X, sctl::Vector<Real>* Xn, sctl::Vector<Long>* element_wise_node_cnt) const override {
    sctl::Vector<Real> x, n, w;
    nodes(x, n, w, xq_, wq_, 0, Size());
    if (X) *X = x;
    if (Xn) *Xn = n;
    if (element_wise_node_cnt) {
      element_wise_node_cnt->ReInit(Size());
      *element_wise_node_cnt = NodesPerElem();
    }
  }

  void GetFarFieldNodes(sctl::Vector<Real>& X, sctl::Vector<Real>& Xn, sctl::Vector<Real>& wts, sctl::Vector<Real>& dist_far,
                        sctl::Vector<Long>& element_wise_node_cnt, const Real /*tol*/) const override {
    nodes(X, Xn, wts, xq_, wq_, 0, Size());
    element_wise_node_cnt.ReInit(Size());
    element_wise_node_cnt = NodesPerElem();
    dist_far.ReInit(Size() * NodesPerElem());
    const Real pi = (Real)3.14159265358979323846264338327950288L;
    for (Long e = 0; e < Size(); e++) {
      const Long i = e / nph_;
      const Real th_mid = pi * (i + (Real)0.5) / nth_;
      const Real diam = R_ * std::sqrt((pi / nth_) * (pi / nth_) + (std::sin(th_mid) * 2 * pi / nph_) * (std::sin(th_mid) * 2 * pi / nph_));
      for (Long k = 0; k < NodesPerElem(); k++) dist_far[e * NodesPerElem() + k] = near_factor_ * diam;
    }
  }

  // M: (nodes*K0) x K1_ for ONE target (boundary_integral.txx:937-939 hands in a pre-sized view)
  template <class Kernel>
  static void NearInterac(sctl::Matrix<Real>& M, const sctl::Vector<Real>& Xt, const sctl::Vector<Real>& normal_trg, const Kernel& ker, Real /*tol*/,
                          const Long elem_idx, const sctl::ElementListBase<Real>* self) {
    const auto* L = dynamic_cast<const SpherePatchList<Real>*>(self);
    SCTL_ASSERT(L != nullptr);
    L->template panel_to_targets<Kernel>(M, Xt, normal_trg, ker, elem_idx);
  }

  // M_lst[e]: (nodes*K0) x (nodes*K1_), the element's own nodes as targets (boundary_integral.txx:791-804)
  template <class Kernel>
  static void SelfInterac(sctl::Vector<sctl::Matrix<Real>>& M_lst, const Kernel& ker, Real /*tol*/, bool trg_dot_prod, const sctl::ElementListBase<Real>* self) {
    const auto* L = dynamic_cast<const SpherePatchList<Real>*>(self);
    SCTL_ASSERT(L != nullptr);
    if (M_lst.Dim() != L->Size()) M_lst.ReInit(L->Size());
#pragma omp parallel for schedule(dynamic)
    for (Long e = 0; e < L->Size(); e++) {
      sctl::Vector<Real> x, n, w;
      L->nodes(x, n, w, L->xq_, L->wq_, e, e + 1);
      L->template panel_to_targets<Kernel>(M_lst[e], x, trg_dot_prod ? n : sctl::Vector<Real>(), ker, e);
    }
  }

 private:
  // Gauss-Legendre nodes / weights on [-1, 1] by Newton iteration on P_n
  static void gauss_legendre(int n, std::vector<Real>& x, std::vector<Real>& w) {
    x.assign(n, 0);
    w.assign(n, 0);
    const long double pi = 3.14159265358979323846264338327950288L;
    for (int i = 0; i < n; i++) {
      long double z = std::cos(pi * (i + 0.75L) / (n + 0.5L)), pp = 1;
      for (int it = 0; it < 100; it++) {
        long double p0 = 1, p1 = z;
        for (int k = 2; k <= n; k++) {
          const long double p2 = ((2 * k - 1) * z * p1 - (k - 1) * p0) / k;
          p0 = p1, p1 = p2;
        }
        if (n == 1) p0 = 1, p1 = z;
        pp = n * (z * p1 - p0) / (z * z - 1);
        const long double dz = p1 / pp;
        z -= dz;
        if (std::fabs((double)dz) < 1e-18) break;
      }
      x[n - 1 - i] = (Real)z;
      w[n - 1 - i] = (Real)(2 / ((1 - z * z) * pp * pp));
    }
  }

  // nodes (AoS), outward normals and weights of elements [e0, e1) for a tensor rule (xr, wr)
  void nodes(sctl::Vector<Real>& X, sctl::Vector<Real>& Xn, sctl::Vector<Real>& W, const std::vector<Real>& xr, const std::vector<Real>& wr, Long e0, Long e1) const {
    const int p = (int)xr.size();
    const Long per = (Long)p * p;
    X.ReInit((e1 - e0) * per * 3);
    Xn.ReInit((e1 - e0) * per * 3);
    W.ReInit((e1 - e0) * per);
    const Real pi = (Real)3.14159265358979323846264338327950288L;
    const Real hth = pi / nth_ / 2, hph = 2 * pi / nph_ / 2;
    for (Long e = e0; e < e1; e++) {
      const Long i = e / nph_, j = e % nph_;
      const Real th_mid = pi * (i + (Real)0.5) / nth_, ph_mid = 2 * pi * (j + (Real)0.5) / nph_;
      for (int a = 0; a < p; a++)
        for (int b = 0; b < p; b++) {
          const Real th = th_mid + hth * xr[a], ph = ph_mid + hph * xr[b];
          const Real nx = std::sin(th) * std::cos(ph), ny = std::sin(th) * std::sin(ph), nz = std::cos(th);
          const Long k = (e - e0) * per + (Long)a * p + b;
          X[k * 3 + 0] = c_[0] + R_ * nx, X[k * 3 + 1] = c_[1] + R_ * ny, X[k * 3 + 2] = c_[2] + R_ * nz;
          Xn[k * 3 + 0] = nx, Xn[k * 3 + 1] = ny, Xn[k * 3 + 2] = nz;
          W[k] = wr[a] * wr[b] * hth * hph * R_ * R_ * std::sin(th);
        }
    }
  }

  // M[(j*K0+k0)][t*K1_+k1] = sum_k P[k][j] w_k K[(k*K0+k0)][t*K1+...]  over the upsampled nodes k of element e: the potential at the
  // targets Xt of a density given by its values at the element's q x q nodes
  template <class Kernel>
  void panel_to_targets(sctl::Matrix<Real>& M, const sctl::Vector<Real>& Xt, const sctl::Vector<Real>& normal_trg, const Kernel& ker, Long e) const {
    constexpr sctl::Integer K0 = Kernel::SrcDim(), K1 = Kernel::TrgDim();
    const bool dot = normal_trg.Dim() != 0;
    const sctl::Integer K1_ = dot ? K1 / 3 : K1;
    const Long Nt = Xt.Dim() / 3, nq = NodesPerElem(), nm = (Long)m_ * m_;
    sctl::Vector<Real> Y, Yn, Yw;
    nodes(Y, Yn, Yw, xm_, wm_, e, e + 1);
    sctl::Matrix<Real> Mk;
    if (sphere_patch_detail::near_on_device()) {
      ker.template KernelMatrix<Real, false>(Mk, Xt, Y, Yn);  // (nm*K0) x (Nt*K1): through the library in the GPU arm
    } else {
      const typename sphere_patch_detail::HostKernelOf<Kernel>::type host_ker;
      host_ker.template KernelMatrix<Real, false>(Mk, Xt, Y, Yn);
    }
    if (M.Dim(0) != nq * K0 || M.Dim(1) != Nt * K1_) M.ReInit(nq * K0, Nt * K1_);
    M.SetZero();
    for (int a = 0; a < q_; a++)
      for (int b = 0; b < q_; b++) {
        const Long j = (Long)a * q_ + b;
        for (int c = 0; c < m_; c++)
          for (int d = 0; d < m_; d++) {
            const Long k = (Long)c * m_ + d;
            const Real coef = P_[(size_t)c * q_ + a] * P_[(size_t)d * q_ + b] * Yw[k];
            for (sctl::Integer k0 = 0; k0 < K0; k0++)
              for (Long t = 0; t < Nt; t++)
                for (sctl::Integer k1 = 0; k1 < K1_; k1++) {
                  Real v;
                  if (dot) {
                    v = 0;
                    for (int l = 0; l < 3; l++) v += Mk[k * K0 + k0][(t * K1_ + k1) * 3 + l] * normal_trg[t * 3 + l];
                  } else {
                    v = Mk[k * K0 + k0][t * K1 + k1];
                  }
                  M[j * K0 + k0][t * K1_ + k1] += coef * v;
                }
          }
      }
  }

  Real R_, c_[3];
  Long nth_, nph_;
  int q_, m_;
  Real near_factor_;
  std::vector<Real> xq_, wq_, xm_, wm_, P_;
};

#endif  // SCTL_B200_TESTS_SPHERE_PATCH_LIST_HPP_
