// kernel-matrix-batch-b200.hpp — all the per-element KernelMatrix blocks of BoundaryIntegralOp::SetupNear in ONE launch
// (reference: include/sctl/boundary_integral.txx:946-1008), over sctl_b200_kernel_matrix_batch (include/sctl_b200.h).
//
// The reference loops over elements and, per element, calls ker_.KernelMatrix on (that element's far-field nodes) x (its
// near targets), scales row (s,k0) by wts_far[s] (:990-997) or contracts the target components with the target normal
// (:968-984). Here the whole sweep is one call:
//
//     sctl::Vector<Real> Mker_all;  sctl::Vector<sctl::Long> Mker_dsp;
//     sctl_b200::KernelMatrixBatch(ker_, Mker_all, Mker_dsp, near_dsp, far_dsp, Xtrg_near, X_far, Xn_far, wts_far, Xn_trg_near);
//     // inside the element loop, instead of the KernelMatrix call + scaling loops:
//     const sctl::Matrix<Real> Mker(far_src_cnt*KDIM0, trg_cnt*KDIM1_, Mker_all.begin() + Mker_dsp[elem_idx], false);
//
// near_dsp / far_dsp are the reference's near_elem_dsp / elem_nds_dsp_far displacement arrays (in POINTS) with the total
// appended, i.e. nbatch+1 entries: block b couples targets [near_dsp[b], near_dsp[b+1]) with sources [far_dsp[b], far_dsp[b+1]).
// wts may be empty (no row scaling); Xn_trg may be empty (no contraction; otherwise TrgDim must be a multiple of 3 and each
// block has TrgDim/3 columns per target). Block b is row-major (Ns_b*SrcDim) x (Nt_b*K1out) at Mker_all[Mker_dsp[b]].
#ifndef SCTL_B200_KERNEL_MATRIX_BATCH_B200_HPP_
#define SCTL_B200_KERNEL_MATRIX_BATCH_B200_HPP_

#include <vector>

#include "generic-kernel-b200.hpp"

namespace sctl_b200 {

template <class uKernel, class Real>
void KernelMatrixBatch(const sctl::GenericKernel<uKernel>&, sctl::Vector<Real>& M, sctl::Vector<sctl::Long>& M_dsp, const sctl::Vector<sctl::Long>& trg_dsp,
                       const sctl::Vector<sctl::Long>& src_dsp, const sctl::Vector<Real>& Xtrg, const sctl::Vector<Real>& Xsrc,
                       const sctl::Vector<Real>& Xn_src, const sctl::Vector<Real>& wts, const sctl::Vector<Real>& Xn_trg, sctl::Integer digits = -1) {
  static_assert(sctl_b200_shim::KernelId<uKernel>::known, "only the kernel_impl micro-kernels of kernel_functions.hpp run on the device");
  using Ker = sctl::GenericKernel<uKernel>;
  constexpr sctl::Integer DIM = Ker::CoordDim(), K0 = Ker::SrcDim(), K1 = Ker::TrgDim(), ND = Ker::NormalDim();
  const sctl::Long nb = trg_dsp.Dim() - 1;
  SCTL_ASSERT(nb >= 0 && src_dsp.Dim() == nb + 1);
  const bool dotn = Xn_trg.Dim() != 0;
  SCTL_ASSERT(!dotn || K1 % DIM == 0);
  const sctl::Integer K1out = dotn ? K1 / DIM : K1;
  const sctl::Long Nt = nb ? trg_dsp[nb] : 0, Ns = nb ? src_dsp[nb] : 0;
  SCTL_ASSERT(Xtrg.Dim() >= Nt * DIM && Xsrc.Dim() >= Ns * DIM);
  SCTL_ASSERT(!ND || Xn_src.Dim() >= Ns * ND);
  SCTL_ASSERT(!wts.Dim() || wts.Dim() >= Ns);
  SCTL_ASSERT(!dotn || Xn_trg.Dim() >= Nt * DIM);
  std::vector<int64_t> td(nb + 1), sd(nb + 1), md(nb > 0 ? nb : 1);
  if (M_dsp.Dim() != nb) M_dsp.ReInit(nb);
  int64_t len = 0;
  for (sctl::Long b = 0; b <= nb; b++) {
    td[b] = trg_dsp[b];
    sd[b] = src_dsp[b];
  }
  for (sctl::Long b = 0; b < nb; b++) {
    md[b] = len;
    M_dsp[b] = len;
    len += (sd[b + 1] - sd[b]) * K0 * (td[b + 1] - td[b]) * K1out;
  }
  if (M.Dim() != len) M.ReInit(len);
  if (!nb || !len) return;
  sctl_b200_shim::check(sctl_b200_kernel_matrix_batch(sctl_b200_shim::KernelId<uKernel>::value, sctl_b200_shim::DType<Real>::value, nb, td.data(), sd.data(),
                                                      md.data(), &Xtrg[0], &Xsrc[0], ND ? &Xn_src[0] : nullptr, wts.Dim() ? &wts[0] : nullptr,
                                                      dotn ? &Xn_trg[0] : nullptr, &M[0], len, digits, 0u),
                        "sctl_b200_kernel_matrix_batch");
  sctl::Long pairs = 0;
  for (sctl::Long b = 0; b < nb; b++) pairs += (sd[b + 1] - sd[b]) * (td[b + 1] - td[b]);
  sctl::Profile::IncrementCounter(sctl::ProfileCounter::FLOP, pairs * uKernel::FLOPS());
}

}  // namespace sctl_b200

#endif  // SCTL_B200_KERNEL_MATRIX_BATCH_B200_HPP_
