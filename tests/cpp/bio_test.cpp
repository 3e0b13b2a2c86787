// The reference's BoundaryIntegralOp (boundary_integral.hpp / .txx) running on libsctl_b200.so - the call sites SURVEY 8 f3 / f4
// name, exercised for real instead of being restated:
//   SetupNear     : ker_.KernelMatrix<Real,true> on (far nodes of an element) x (its near targets) inside the reference's own
//                   `omp parallel for` (boundary_integral.txx:949-986), row scaling by wts_far / target-normal contraction (:968-997)
//   ComputeFarField: F_far *= wts_far, fmm.SetSrcDensity, fmm.Eval -> ParticleFMM::EvalDirect -> the stored static Eval pointer
//                   (:1016-1077, fmm-wrapper.txx:557), target-normal dot product (:1060-1071)
//   ComputePotential = far field + near correction (:608-614)
//
// Two arms with the SAME element list (tests/cpp/sphere_patch_list.hpp, test-only: the reference ships none):
//   GPU arm : BoundaryIntegralOp<Real, GenericKernel<kernel_impl::K>>           - the shim's specialisations forward to the library
//   CPU arm : BoundaryIntegralOp<Real, GenericKernel<CpuCopy<kernel_impl::K>>>  - a distinct type: the reference's own templates
// Built twice by tests/cpp/Makefile:
//   bio_test          against the UNMODIFIED reference headers (drop-in through the shim only)
//   bio_test_patched  against a copy of boundary_integral.{hpp,txx} with tests/cpp/boundary_integral_b200.patch applied:
//                     batched SetupNear (sctl_b200::KernelMatrixBatch) and the device-resident registry with fused weights /
//                     target normals (sctl_b200::ParticleFMM) for the kernel_impl kernels; CpuCopy kernels keep the reference path.
// Usage: bio_test [--cpu-only | --near-on-device] [n_theta=12] [reps=3] [case-filter]
//        elements per sphere = n_theta * 2 n_theta, 36 nodes each, two spheres; case-filter = substring of a kernel name, e.g.
//        "DxU", to run only those cases (used for the large timing runs)
#ifdef BIO_PATCHED
// the patched copies carry the reference's include guards, so they must come first; <sctl.hpp> then skips the originals
#include <sctl/boundary_integral.hpp>
#include <sctl/boundary_integral.txx>
#endif
#include <sctl.hpp>
#include <sctl_b200/generic-kernel-b200.hpp>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>

#include "sphere_patch_list.hpp"

using sctl::Long;
template <class K> struct CpuCopy : K {};

#ifndef BIO_ARM_NAME
#ifdef BIO_PATCHED
#define BIO_ARM_NAME "B200-patched"
#else
#define BIO_ARM_NAME "B200-shim"
#endif
#endif

static int g_fail = 0;
static std::string g_filter;
static bool g_cpu_only = false;  // --cpu-only: run the sanity check and the reference arm alone (no GPU needed; used where the test is authored)
static void report(const std::string& what, double err, double tol) {
  const bool ok = err <= tol;
  std::printf("%-78s err = %9.3e  tol = %7.1e  %s\n", what.c_str(), err, tol, ok ? "ok" : "FAIL");
  if (!ok) g_fail++;
}
template <class Real> static double rel_l2(const sctl::Vector<Real>& a, const sctl::Vector<Real>& b) {
  if (a.Dim() != b.Dim() || a.Dim() == 0) return 1e300;
  long double num = 0, den = 0;
  for (Long i = 0; i < a.Dim(); i++) {
    const long double d = (long double)a[i] - (long double)b[i];
    num += d * d;
    den += (long double)b[i] * (long double)b[i];
  }
  return den > 0 ? (double)sqrtl(num / den) : (double)sqrtl(num);
}

struct Timing {
  double setup = 0, apply = 0;
  Long nodes = 0, targets = 0;
};

// builds the operator on two spheres (two element lists), optional off-surface targets, and applies it `reps` times
template <class Real, class Ker>
static sctl::Vector<Real> run_arm(Long nth, bool dot_normal, bool off_surface, int reps, Timing& tm, Real tol = (Real)1e-10) {
  const Ker ker;
  sctl::BoundaryIntegralOp<Real, Ker> op(ker, dot_normal, sctl::Comm::Self());
  op.SetAccuracy(tol);
  // q = 6 nodes per direction, 2x upsampled near rule, near region = 1.2 patch diameters
  const SpherePatchList<Real> a((Real)1.0, 0, 0, 0, nth, 2 * nth, 6, 2, (Real)1.2);
  const SpherePatchList<Real> b((Real)0.6, (Real)2.1, (Real)0.3, (Real)-0.2, nth, 2 * nth, 6, 2, (Real)1.2);
  op.template AddElemList<SpherePatchList<Real>>(a, "sphere-a");
  op.template AddElemList<SpherePatchList<Real>>(b, "sphere-b");
  if (off_surface) {  // targets on a shell just outside sphere a (near for some elements, far for most) and far away
    const Long Nt = 4000;
    sctl::Vector<Real> Xt(Nt * 3);
    unsigned long long st = 99;
    auto u01 = [&st]() {
      st = st * 6364136223846793005ull + 1442695040888963407ull;
      return (double)(st >> 11) * (1.0 / 9007199254740992.0);
    };
    for (Long i = 0; i < Nt; i++) {
      const double z = 2 * u01() - 1, ph = 6.283185307179586 * u01(), r = (i % 4 == 0) ? 3.5 : 1.02 + 0.1 * u01();
      const double s = std::sqrt(1 - z * z);
      Xt[i * 3 + 0] = (Real)(r * s * std::cos(ph)), Xt[i * 3 + 1] = (Real)(r * s * std::sin(ph)), Xt[i * 3 + 2] = (Real)(r * z);
    }
    op.SetTargetCoord(Xt);
    if (dot_normal) {  // any unit vectors will do
      sctl::Vector<Real> Xn(Nt * 3);
      for (Long i = 0; i < Nt; i++) {
        const Real x = Xt[i * 3], y = Xt[i * 3 + 1], z = Xt[i * 3 + 2], r = std::sqrt(x * x + y * y + z * z);
        Xn[i * 3] = x / r, Xn[i * 3 + 1] = y / r, Xn[i * 3 + 2] = z / r;
      }
      op.SetTargetNormal(Xn);
    }
  }
  double t0 = omp_get_wtime();
  op.Setup();
  tm.setup = omp_get_wtime() - t0;
  const Long N0 = op.Dim(0);
  sctl::Vector<Real> F(N0), U;
  for (Long i = 0; i < N0; i++) F[i] = (Real)(0.3 + std::sin(0.37 * (double)i));  // smooth-ish, deterministic
  op.ComputePotential(U, F);  // warm
  t0 = omp_get_wtime();
  for (int r = 0; r < reps; r++) {
    for (Long i = 0; i < N0; i++) F[i] = (Real)(0.3 + std::sin(0.37 * (double)i + 0.01 * r));  // a new density per solver iteration
    op.ComputePotential(U, F);
  }
  tm.apply = (omp_get_wtime() - t0) / reps;
  tm.nodes = N0 / Ker::SrcDim();
  tm.targets = op.Dim(1) / (dot_normal ? Ker::TrgDim() / 3 : Ker::TrgDim());
  return U;
}

template <class Real, class UK> static void compare(const char* name, Long nth, bool dot_normal, bool off_surface, int reps, double tol) {
  using Gpu = sctl::GenericKernel<UK>;
  using Cpu = sctl::GenericKernel<CpuCopy<UK>>;
  if (!g_filter.empty() && std::string(name).find(g_filter) == std::string::npos) return;
  Timing tg, tc;
  if (g_cpu_only) {
    const sctl::Vector<Real> Uc = run_arm<Real, Cpu>(nth, dot_normal, off_surface, reps, tc);
    double nrm = 0;
    for (Long i = 0; i < Uc.Dim(); i++) nrm += (double)Uc[i] * (double)Uc[i];
    std::printf("%-20s reference-CPU only: %lld nodes, %lld targets, |U| = %.6e, Setup %.3f s, apply %.4f s\n", name, (long long)tc.nodes, (long long)tc.targets,
                std::sqrt(nrm), tc.setup, tc.apply);
    if (!(nrm > 0) || !std::isfinite(nrm)) g_fail++;
    return;
  }
  const long long l0 = sctl_b200_launch_count();
  const sctl::Vector<Real> Ug = run_arm<Real, Gpu>(nth, dot_normal, off_surface, reps, tg);
  const long long launches = sctl_b200_launch_count() - l0;
  const sctl::Vector<Real> Uc = run_arm<Real, Cpu>(nth, dot_normal, off_surface, reps, tc);
  const std::string tag = std::string(name) + (sizeof(Real) == 8 ? " f64" : " f32") + (dot_normal ? " n.grad" : "") + (off_surface ? " off-surface" : " on-surface") +
                          " (" + std::to_string(tg.nodes) + " nodes, " + std::to_string(tg.targets) + " targets)";
  report(tag + " ComputePotential", rel_l2(Ug, Uc), tol);
  report(tag + " ran on the GPU (kernel launches > 0)", launches > 0 ? 0.0 : 1.0, 0.0);
  std::printf("    timing  %-14s Setup %9.3f s   ComputePotential %9.4f s/apply   (%lld kernel launches)\n", BIO_ARM_NAME, tg.setup, tg.apply, launches);
  std::printf("    timing  %-14s Setup %9.3f s   ComputePotential %9.4f s/apply   (%d OpenMP threads)\n", "reference-CPU", tc.setup, tc.apply, omp_get_max_threads());
}

int main(int argc, char** argv) {
  int arg0 = 1;
  if (argc > 1 && std::string(argv[1]) == "--cpu-only") g_cpu_only = true, arg0 = 2;
  if (argc > 1 && std::string(argv[1]) == "--near-on-device") sphere_patch_detail::near_on_device() = true, arg0 = 2;
  const Long nth = argc > arg0 ? std::atol(argv[arg0]) : 12;
  const int reps = argc > arg0 + 1 ? std::atoi(argv[arg0 + 1]) : 3;
  if (argc > arg0 + 2) g_filter = argv[arg0 + 2];
  std::printf("element-list near quadrature evaluated %s\n", sphere_patch_detail::near_on_device()
                                                                 ? "through the kernel object (GPU arm: ~1e5 one-target KernelMatrix calls through the shim)"
                                                                 : "on the host (the library sees only the reference's own call sites)");
  {  // the operator means something: exterior Laplace single-layer potential of a unit density on the unit sphere is R^2/|x|
    using Ker = sctl::GenericKernel<CpuCopy<sctl::kernel_impl::Laplace3D_FxU>>;
    const Ker ker;
    sctl::BoundaryIntegralOp<double, Ker> op(ker);
    const SpherePatchList<double> a(1.0, 0, 0, 0, 6, 12, 6, 2, 1.2);
    op.AddElemList(a, "a");
    sctl::Vector<double> Xt(3);
    Xt[0] = 1.7, Xt[1] = -0.4, Xt[2] = 0.9;
    op.SetTargetCoord(Xt);
    sctl::Vector<double> F(op.Dim(0)), U;
    F = 1;
    op.ComputePotential(U, F);
    const double want = 1.0 / std::sqrt(1.7 * 1.7 + 0.4 * 0.4 + 0.9 * 0.9);
    report("sanity: Laplace single layer of unit density on the unit sphere = R^2/|x| outside", std::fabs(U[0] - want) / want, 1e-8);
  }
  compare<double, sctl::kernel_impl::Laplace3D_FxU>("Laplace3D_FxU", nth, false, false, reps, 1e-11);
  compare<double, sctl::kernel_impl::Stokes3D_DxU>("Stokes3D_DxU", nth, false, false, reps, 1e-11);
  compare<double, sctl::kernel_impl::Stokes3D_DxU>("Stokes3D_DxU", nth, false, true, reps, 1e-11);
  compare<double, sctl::kernel_impl::Laplace3D_FxdU>("Laplace3D_FxdU", nth, true, false, reps, 1e-11);
  compare<double, sctl::kernel_impl::Stokes3D_FxT>("Stokes3D_FxT", nth, true, true, reps, 1e-11);
  compare<float, sctl::kernel_impl::Stokes3D_FxU>("Stokes3D_FxU", nth, false, false, reps, 2e-5);
  std::printf(g_fail ? "BIO TEST FAILED (%d)\n" : "BIO TEST PASSED\n", g_fail);
  return g_fail ? 1 : 0;
}
