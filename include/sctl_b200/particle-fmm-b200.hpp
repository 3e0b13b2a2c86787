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
// part of this library; SetKernels / the M2M..L2T kernels are accepted and dimension-checked only; (3) multi-rank: the
// MPI ring of EvalDirect (fmm-wrapper.txx:537-558) is replaced by the NCCL driver inside libsctl_b200.so. With a
// sctl::Comm of Size() > 1 (an MPI build of SCTL, one rank per GPU, sctl_b200_set_device(local_rank) called first) the
// constructor / SetComm bootstrap an NCCL communicator through the reference's own Comm::Bcast (comm.hpp:167); programs
// that bring their own ranks (threads, another launcher) pass a handle from sctl_b200_comm_create to SetComm(void*).
// Every rank then hands ITS particles to the Set* calls (collective) and Eval returns the potentials of its own targets,
// exactly the reference's calling convention. Only the eight kernel_impl micro-kernels are supported as S2T kernels.
#ifndef SCTL_B200_PARTICLE_FMM_B200_HPP_
#define SCTL_B200_PARTICLE_FMM_B200_HPP_

#include <string>

#include "generic-kernel-b200.hpp"
#include "sctl/comm.hpp"

namespace sctl_b200 {

template <class uKernel> struct UKernelOf;
template <class uKernel> struct UKernelOf<sctl::GenericKernel<uKernel>> { using type = uKernel; };

// true for GenericKernel<K> with K one of the kernel_impl micro-kernels the library implements: lets reference code choose
// between this class and sctl::ParticleFMM at compile time (tests/cpp/boundary_integral_b200.patch)
template <class Kernel> struct IsDeviceKernel { static constexpr bool value = false; };
template <class uKernel> struct IsDeviceKernel<sctl::GenericKernel<uKernel>> { static constexpr bool value = sctl_b200_shim::KernelId<uKernel>::known; };

template <class Real, sctl::Integer DIM = 3> class ParticleFMM {
  static_assert(DIM == 3, "sctl_b200::ParticleFMM supports DIM == 3 (the Laplace3D / Stokes3D kernels)");

 public:
  ParticleFMM(const ParticleFMM&) = delete;
  ParticleFMM& operator=(const ParticleFMM&) = delete;

  ParticleFMM(const sctl::Comm& comm = sctl::Comm::Self()) : h_(nullptr), owned_comm_(nullptr) {
    sctl_b200_shim::check(sctl_b200_fmm_create(&h_, sctl_b200_shim::DType<Real>::value), "sctl_b200_fmm_create");
    SetComm(comm);
  }
  ~ParticleFMM() {
    sctl_b200_fmm_destroy(h_);
    sctl_b200_comm_destroy(owned_comm_);
  }

  // fmm-wrapper.txx:204 / fmm-wrapper.hpp:45. Collective over `comm` when comm.Size() > 1: rank 0 creates the NCCL id, the
  // reference's Comm::Bcast carries it, every rank joins on its current CUDA device.
  void SetComm(const sctl::Comm& comm) {
    sctl_b200_shim::check(sctl_b200_fmm_set_comm(h_, nullptr), "sctl_b200_fmm_set_comm");
    sctl_b200_comm_destroy(owned_comm_);
    owned_comm_ = nullptr;
    if (comm.Size() == 1) return;
    sctl::Vector<char> id(128);
    id.SetZero();
    if (comm.Rank() == 0) sctl_b200_shim::check(sctl_b200_comm_unique_id(&id[0], 128), "sctl_b200_comm_unique_id");
    comm.Bcast(id.begin(), 128, 0);
    sctl_b200_shim::check(sctl_b200_comm_create(&owned_comm_, (int)comm.Size(), (int)comm.Rank(), &id[0]), "sctl_b200_comm_create");
    sctl_b200_shim::check(sctl_b200_fmm_set_comm(h_, owned_comm_), "sctl_b200_fmm_set_comm");
  }
  // a communicator the caller created with sctl_b200_comm_create (not owned; nullptr detaches)
  void SetComm(void* sctl_b200_comm) { sctl_b200_shim::check(sctl_b200_fmm_set_comm(h_, sctl_b200_comm), "sctl_b200_fmm_set_comm"); }
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
    SCTL_ASSERT(src_normal.Dim() == 0 || src_normal.Dim() == Ns * DIM);  // the C ABI reads Ns*NormalDim values (fmm-wrapper.txx:535)
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
  void* owned_comm_;
};

}  // namespace sctl_b200

#endif  // SCTL_B200_PARTICLE_FMM_B200_HPP_
