// particle-fmm-b200.hpp — GPU-aware replacement of the registry + EvalDirect of sctl::ParticleFMM<Real,3>
// (reference: include/sctl/fmm-wrapper.hpp:33-160, fmm-wrapper.txx:240-562), backed by the device-resident particle
// context of libsctl_b200.so (sctl_b200_fmm_*, include/sctl_b200.h).
//
//     #include <sctl.hpp>
//     #include <sctl_b200/particle-fmm-b200.hpp>
//     sctl_b200::ParticleFMM<double,3> fmm;          // same calls as sctl::ParticleFMM<double,3>
//
// Same public methods, argument order and error convention (SCTL_ASSERT_MSG / SCTL_ERROR -> abort) as the reference.
// Differences: (1) coordinates, normals, densities and packed source records stay in HBM between calls — a GMRES loop
// that only calls SetSrcDensity + Eval re-uploads just the density; (2) Eval == EvalDirect, exactly like a reference
// build without PVFMM (fmm-wrapper.txx:481-489): the FMM far field is an external dependency of the reference and not
// part of this library; SetKernels / the M2M..L2T kernels are accepted and dimension-checked only; (3) single rank: the
// MPI ring of EvalDirect (fmm-wrapper.txx:537-558) is replaced by the NCCL driver of sctl_b200/distributed.py for
// one-process-per-GPU runs. Only the eight kernel_impl micro-kernels are supported as S2T kernels.
#ifndef SCTL_B200_PARTICLE_FMM_B200_HPP_
#define SCTL_B200_PARTICLE_FMM_B200_HPP_

#include <string>

#include "generic-kernel-b200.hpp"
#include "sctl/comm.hpp"

namespace sctl_b200 {

template <class uKernel> struct UKernelOf;
template <class uKernel> struct UKernelOf<sctl::GenericKernel<uKernel>> { using type = uKernel; };

template <class Real, sctl::Integer DIM = 3> class ParticleFMM {
  static_assert(DIM == 3, "sctl_b200::ParticleFMM supports DIM == 3 (the Laplace3D / Stokes3D kernels)");

 public:
  ParticleFMM(const ParticleFMM&) = delete;
  ParticleFMM& operator=(const ParticleFMM&) = delete;

  ParticleFMM(const sctl::Comm& comm = sctl::Comm::Self()) : h_(nullptr) {
    SCTL_ASSERT_MSG(comm.Size() == 1, "sctl_b200::ParticleFMM is single-rank; use the NCCL driver for multi-GPU runs.");
    sctl_b200_shim::check(sctl_b200_fmm_create(&h_, sctl_b200_shim::DType<Real>::value), "sctl_b200_fmm_create");
  }
  ~ParticleFMM() { sctl_b200_fmm_destroy(h_); }

  void SetComm(const sctl::Comm& comm) { SCTL_ASSERT_MSG(comm.Size() == 1, "sctl_b200::ParticleFMM is single-rank."); }
  void SetAccuracy(sctl::Integer digits) { sctl_b200_shim::check(sctl_b200_fmm_set_accuracy(h_, (int)digits), "sctl_b200_fmm_set_accuracy"); }

  template <class KerM2M, class KerM2L, class KerL2L> void SetKernels(const KerM2M& ker_m2m, const KerM2L& ker_m2l, const KerL2L& ker_l2l) {
    SCTL_ASSERT(ker_m2m.CoordDim() == DIM && ker_m2l.CoordDim() == DIM && ker_l2l.CoordDim() == DIM);  // fmm-wrapper.txx:265-267
    SCTL_ASSERT(ker_m2l.SrcDim() == ker_m2m.SrcDim());                                                  // :268
    SCTL_ASSERT(ker_m2l.TrgDim() == ker_l2l.TrgDim());                                                  // :269
  }
  template <class KerS2M, class KerS2L> void AddSrc(const std::string& name, const KerS2M& ker_s2m, const KerS2L& ker_s2l) {
    SCTL_ASSERT(ker_s2m.CoordDim() == DIM && ker_s2l.CoordDim() == DIM);  // fmm-wrapper.txx:310-313
    SCTL_ASSERT(ker_s2l.SrcDim() == ker_s2m.SrcDim());
    SCTL_ASSERT(ker_s2l.NormalDim() == ker_s2m.NormalDim());
    sctl_b200_shim::check(sctl_b200_fmm_add_src(h_, name.c_str(), (int)ker_s2m.SrcDim(), (int)ker_s2m.NormalDim()), "sctl_b200_fmm_add_src");
  }
  template <class KerM2T, class KerL2T> void AddTrg(const std::string& name, const KerM2T& ker_m2t, const KerL2T& ker_l2t) {
    SCTL_ASSERT(ker_m2t.CoordDim() == DIM && ker_l2t.CoordDim() == DIM);  // fmm-wrapper.txx:349-351
    SCTL_ASSERT(ker_m2t.TrgDim() == ker_l2t.TrgDim());
    sctl_b200_shim::check(sctl_b200_fmm_add_trg(h_, name.c_str(), (int)ker_l2t.TrgDim()), "sctl_b200_fmm_add_trg");
  }
  template <class KerS2T> void SetKernelS2T(const std::string& src_name, const std::string& trg_name, const KerS2T&) {
    using uKernel = typename UKernelOf<KerS2T>::type;
    static_assert(sctl_b200_shim::KernelId<uKernel>::known, "S2T kernel must be one of the kernel_impl micro-kernels of kernel_functions.hpp");
    sctl_b200_shim::check(sctl_b200_fmm_set_kernel_s2t(h_, src_name.c_str(), trg_name.c_str(), sctl_b200_shim::KernelId<uKernel>::value),
                          "sctl_b200_fmm_set_kernel_s2t");
  }
  void DeleteSrc(const std::string& name) { sctl_b200_shim::check(sctl_b200_fmm_delete_src(h_, name.c_str()), "sctl_b200_fmm_delete_src"); }
  void DeleteTrg(const std::string& name) { sctl_b200_shim::check(sctl_b200_fmm_delete_trg(h_, name.c_str()), "sctl_b200_fmm_delete_trg"); }

  void SetSrcCoord(const std::string& name, const sctl::Vector<Real>& src_coord, const sctl::Vector<Real>& src_normal = sctl::Vector<Real>()) {
    const sctl::Long Ns = src_coord.Dim() / DIM;
    SCTL_ASSERT(src_coord.Dim() == Ns * DIM);
    sctl_b200_shim::check(sctl_b200_fmm_set_src_coord(h_, name.c_str(), (int64_t)Ns, sctl_b200_shim::cptr(src_coord), sctl_b200_shim::cptr(src_normal), 0u),
                          "sctl_b200_fmm_set_src_coord");
  }
  void SetSrcDensity(const std::string& name, const sctl::Vector<Real>& src_density) {
    sctl_b200_shim::check(sctl_b200_fmm_set_src_density(h_, name.c_str(), (int64_t)src_density.Dim(), sctl_b200_shim::cptr(src_density), 0u),
                          "sctl_b200_fmm_set_src_density");
  }
  void SetTrgCoord(const std::string& name, const sctl::Vector<Real>& trg_coord) {
    const sctl::Long Nt = trg_coord.Dim() / DIM;
    SCTL_ASSERT(trg_coord.Dim() == Nt * DIM);
    sctl_b200_shim::check(sctl_b200_fmm_set_trg_coord(h_, name.c_str(), (int64_t)Nt, sctl_b200_shim::cptr(trg_coord), 0u), "sctl_b200_fmm_set_trg_coord");
  }

  // Extensions (not in the reference's ParticleFMM): the glue BoundaryIntegralOp::ComputeFarField does on the host around
  // fmm.Eval, fused on the device. Weights scale the density when records are packed (boundary_integral.txx:1036-1051);
  // target normals make Eval return TrgDim/3 values per target, the component triplets dotted with the normal (:1060-1071).
  void SetSrcWeights(const std::string& name, const sctl::Vector<Real>& wts) {
    sctl_b200_shim::check(sctl_b200_fmm_set_src_weights(h_, name.c_str(), (int64_t)wts.Dim(), sctl_b200_shim::cptr(wts), 0u), "sctl_b200_fmm_set_src_weights");
  }
  void SetTrgNormal(const std::string& name, const sctl::Vector<Real>& trg_normal) {
    sctl_b200_shim::check(sctl_b200_fmm_set_trg_normal(h_, name.c_str(), (int64_t)trg_normal.Dim(), sctl_b200_shim::cptr(trg_normal), 0u),
                          "sctl_b200_fmm_set_trg_normal");
  }

  void Eval(sctl::Vector<Real>& U, const std::string& trg_name) const { EvalDirect(U, trg_name); }
  void EvalDirect(sctl::Vector<Real>& U, const std::string& trg_name) const {
    int64_t Nt = 0;
    int trg_dim = 0;
    sctl_b200_shim::check(sctl_b200_fmm_trg_count(h_, trg_name.c_str(), &Nt, &trg_dim), "sctl_b200_fmm_trg_count");
    if (U.Dim() != Nt * trg_dim) U.ReInit(Nt * trg_dim);  // fmm-wrapper.txx:500-502 (contents are overwritten)
    sctl_b200_shim::check(sctl_b200_fmm_eval_direct(h_, trg_name.c_str(), sctl_b200_shim::mptr(U), (int64_t)U.Dim(), 0u), "sctl_b200_fmm_eval_direct");
  }

 private:
  void* h_;
};

}  // namespace sctl_b200

#endif  // SCTL_B200_PARTICLE_FMM_B200_HPP_
