// The NCCL target-sharding driver of libsctl_b200.so (sctl_b200_comm_* / sctl_b200_eval_sharded / sctl_b200_fmm_set_comm)
// against the reference's own ParticleFMM::EvalDirect (fmm-wrapper.txx:490-562) on the GLOBAL problem.
//
// The reference's multi-rank form needs MPI (absent here), so its single-rank form is the oracle: EvalDirect over all
// particles with CPU kernels (CpuCopy<K>, which the shim's specialisations do not match). The GPU side runs G ranks as G host
// THREADS of this process, one per GPU, each holding its block of the targets and sources - NCCL supports one rank per thread
// (ncclCommInitRank from a shared unique id). G = argv[1] or every visible GPU; with one GPU the communicator degenerates
// (no NCCL traffic) and the test still checks the calling convention.
//
// Built where /root/reference exists (tests/cpp/Makefile) and run on the GPU box by tests/test_cpp_shim.py.
#include <sctl.hpp>
#include <sctl_b200/generic-kernel-b200.hpp>
#include <sctl_b200/particle-fmm-b200.hpp>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

using sctl::Long;
template <class K> struct CpuCopy : K {};

static unsigned long long g_state = 77;
static double urand() {
  unsigned long long z = (g_state += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return (double)(z >> 11) * (1.0 / 9007199254740992.0) - 0.5;
}
template <class Real> static void fill(sctl::Vector<Real>& v) {
  for (auto& a : v) a = (Real)urand();
}
template <class Real> static double rel_l2(const sctl::Vector<Real>& a, const sctl::Vector<Real>& b) {
  if (a.Dim() != b.Dim()) return 1e300;
  long double num = 0, den = 0;
  for (Long i = 0; i < a.Dim(); i++) {
    const long double d = (long double)a[i] - (long double)b[i];
    num += d * d;
    den += (long double)b[i] * (long double)b[i];
  }
  return den > 0 ? (double)sqrtl(num / den) : (double)sqrtl(num);
}
static int g_fail = 0;
static void report(const std::string& what, double err, double tol) {
  const bool ok = err <= tol;
  std::printf("%-86s err = %9.3e  tol = %7.1e  %s\n", what.c_str(), err, tol, ok ? "ok" : "FAIL");
  if (!ok) g_fail++;
}
#define CHECK_RC(call)                                                                       \
  do {                                                                                       \
    int rc__ = (call);                                                                       \
    if (rc__) {                                                                              \
      std::printf("rank %d: %s failed (%d): %s\n", rank, #call, rc__, sctl_b200_last_error()); \
      fails[rank]++;                                                                         \
      return;                                                                                \
    }                                                                                        \
  } while (0)

// block edges: even = the reference's PartitionN formula (fmm-wrapper.txx:507); skewed = rank r owns a share ~ (r+1)
static std::vector<Long> edges(Long n, int G, bool skewed) {
  std::vector<Long> e(G + 1, 0);
  double tot = 0;
  for (int r = 0; r < G; r++) tot += r + 1;
  double acc = 0;
  for (int r = 0; r < G; r++) {
    acc += r + 1;
    e[r + 1] = skewed ? (Long)std::llround(acc / tot * n) : n * (r + 1) / G;
  }
  e[G] = n;
  return e;
}
template <class Real> static sctl::Vector<Real> slice(const sctl::Vector<Real>& v, Long b, Long e, Long w) {
  sctl::Vector<Real> out((e - b) * w);
  for (Long i = 0; i < out.Dim(); i++) out[i] = v[b * w + i];
  return out;
}

// reference: single-rank EvalDirect of the global problem, DL + SL summed into one target (the set-up of fmm-wrapper.txx:35-93)
template <class Real> static sctl::Vector<Real> reference_eval(const sctl::Vector<Real>& Xt, const sctl::Vector<Real>& Xs, const sctl::Vector<Real>& Xn,
                                                               const sctl::Vector<Real>& Fsl, const sctl::Vector<Real>& Fdl, bool with_dl) {
  using SL = sctl::GenericKernel<CpuCopy<sctl::kernel_impl::Stokes3D_FxU>>;
  using DL = sctl::GenericKernel<CpuCopy<sctl::kernel_impl::Stokes3D_DxU>>;
  SL sl;
  DL dl;
  sctl::ParticleFMM<Real, 3> fmm(sctl::Comm::Self());
  fmm.SetKernels(sl, sl, sl);
  fmm.AddTrg("U", sl, sl);
  fmm.AddSrc("SL", sl, sl);
  fmm.SetKernelS2T("SL", "U", sl);
  fmm.SetTrgCoord("U", Xt);
  fmm.SetSrcCoord("SL", Xs);
  fmm.SetSrcDensity("SL", Fsl);
  if (with_dl) {
    fmm.AddSrc("DL", dl, dl);
    fmm.SetKernelS2T("DL", "U", dl);
    fmm.SetSrcCoord("DL", Xs, Xn);
    fmm.SetSrcDensity("DL", Fdl);
  }
  sctl::Vector<Real> U;
  fmm.EvalDirect(U, "U");
  return U;
}

template <class Real> static void run(int G, Long Nt, Long Ns, bool skewed, double tol) {
  const int dtype = sizeof(Real) == 8 ? SCTL_B200_F64 : SCTL_B200_F32;
  const std::string tag = std::string(sizeof(Real) == 8 ? "f64" : "f32") + " G=" + std::to_string(G) + (skewed ? " skewed" : " even") + " " + std::to_string(Nt) + "x" + std::to_string(Ns);
  sctl::Vector<Real> Xt(Nt * 3), Xs(Ns * 3), Xn(Ns * 3), Fsl(Ns * 3), Fdl(Ns * 3);
  fill(Xt), fill(Xs), fill(Xn), fill(Fsl), fill(Fdl);
  for (Long i = 0; i < 3 * std::min<Long>(50, std::min(Nt, Ns)); i++) Xt[i] = Xs[i];  // coincident points
  const sctl::Vector<Real> Uref_sl = reference_eval(Xt, Xs, Xn, Fsl, Fdl, false);
  const sctl::Vector<Real> Uref_both = reference_eval(Xt, Xs, Xn, Fsl, Fdl, true);

  const std::vector<Long> te = edges(Nt, G, skewed), se = edges(Ns, G, skewed);
  char id[128] = {0};
  if (G > 1 && sctl_b200_comm_unique_id(id, 128)) {
    std::printf("sctl_b200_comm_unique_id: %s\n", sctl_b200_last_error());
    g_fail++;
    return;
  }
  sctl::Vector<Real> U_host(Nt * 3), U_dev(Nt * 3), U_fmm(Nt * 3), U_fmm2(Nt * 3), U_glue(Nt * 3);
  // reference for (4): ParticleFMM::EvalDirect with Stokes FxT on densities pre-scaled by the weights (boundary_integral.txx:1036-1051),
  // then the target-normal dot product on the host (:1060-1071)
  sctl::Vector<Real> Wq(Ns), Xnt(Nt * 3), Uref_glue(Nt * 3);
  fill(Wq), fill(Xnt);
  {
    using FxT = sctl::GenericKernel<CpuCopy<sctl::kernel_impl::Stokes3D_FxT>>;
    FxT fxt;
    sctl::Vector<Real> Fw = Fsl, U9;
    for (Long i = 0; i < Ns; i++)
      for (int k = 0; k < 3; k++) Fw[i * 3 + k] *= Wq[i];
    sctl::ParticleFMM<Real, 3> fmm(sctl::Comm::Self());
    fmm.SetKernels(fxt, fxt, fxt);
    fmm.AddTrg("T", fxt, fxt);
    fmm.AddSrc("S", fxt, fxt);
    fmm.SetKernelS2T("S", "T", fxt);
    fmm.SetTrgCoord("T", Xt);
    fmm.SetSrcCoord("S", Xs);
    fmm.SetSrcDensity("S", Fw);
    fmm.EvalDirect(U9, "T");
    for (Long t = 0; t < Nt; t++)
      for (int k = 0; k < 3; k++) {
        Real v = 0;
        for (int l = 0; l < 3; l++) v += U9[(t * 3 + k) * 3 + l] * Xnt[t * 3 + l];
        Uref_glue[t * 3 + k] = v;
      }
  }
  std::vector<int> fails(G, 0);
  std::vector<std::thread> th;
  for (int rank = 0; rank < G; rank++) {
    th.emplace_back([&, rank]() {
      CHECK_RC(sctl_b200_set_device(rank));
      void* comm = nullptr;
      CHECK_RC(sctl_b200_comm_create(&comm, G, rank, id));
      const Long t0 = te[rank], t1 = te[rank + 1], s0 = se[rank], s1 = se[rank + 1];
      const sctl::Vector<Real> xt = slice(Xt, t0, t1, 3), xs = slice(Xs, s0, s1, 3), xn = slice(Xn, s0, s1, 3), fsl = slice(Fsl, s0, s1, 3), fdl = slice(Fdl, s0, s1, 3);
      using sctl_b200_shim::cptr;
      {  // (1) host pointers, source counts exchanged by the library, overwrite
        sctl::Vector<Real> u((t1 - t0) * 3);
        u = 123;
        CHECK_RC(sctl_b200_eval_sharded(comm, SCTL_B200_STOKES3D_FXU, dtype, t1 - t0, s1 - s0, cptr(xt), cptr(xs), nullptr, cptr(fsl), u.Dim() ? &u[0] : nullptr, -1,
                                        SCTL_B200_OVERWRITE, nullptr));
        for (Long i = 0; i < u.Dim(); i++) U_host[t0 * 3 + i] = u[i];
      }
      {  // (2) device pointers, counts given, SL overwritten then DL ACCUMULATED on top (generic-kernel.txx:98-101)
        std::vector<int64_t> counts(G);
        for (int r = 0; r < G; r++) counts[r] = se[r + 1] - se[r];
        void *dxt, *dxs, *dxn, *dfs, *dfd, *du;
        CHECK_RC(sctl_b200_buf_create(&dxt, dtype, xt.Dim()));
        CHECK_RC(sctl_b200_buf_create(&dxs, dtype, xs.Dim()));
        CHECK_RC(sctl_b200_buf_create(&dxn, dtype, xn.Dim()));
        CHECK_RC(sctl_b200_buf_create(&dfs, dtype, fsl.Dim()));
        CHECK_RC(sctl_b200_buf_create(&dfd, dtype, fdl.Dim()));
        CHECK_RC(sctl_b200_buf_create(&du, dtype, (t1 - t0) * 3));
        CHECK_RC(sctl_b200_buf_upload(dxt, cptr(xt), dtype, xt.Dim()));
        CHECK_RC(sctl_b200_buf_upload(dxs, cptr(xs), dtype, xs.Dim()));
        CHECK_RC(sctl_b200_buf_upload(dxn, cptr(xn), dtype, xn.Dim()));
        CHECK_RC(sctl_b200_buf_upload(dfs, cptr(fsl), dtype, fsl.Dim()));
        CHECK_RC(sctl_b200_buf_upload(dfd, cptr(fdl), dtype, fdl.Dim()));
        CHECK_RC(sctl_b200_eval_sharded(comm, SCTL_B200_STOKES3D_FXU, dtype, t1 - t0, s1 - s0, dxt, dxs, nullptr, dfs, du, -1, SCTL_B200_DEVICE_PTRS | SCTL_B200_OVERWRITE, counts.data()));
        CHECK_RC(sctl_b200_eval_sharded(comm, SCTL_B200_STOKES3D_DXU, dtype, t1 - t0, s1 - s0, dxt, dxs, dxn, dfd, du, -1, SCTL_B200_DEVICE_PTRS, counts.data()));
        sctl::Vector<Real> u((t1 - t0) * 3);
        CHECK_RC(sctl_b200_buf_download(du, u.Dim() ? &u[0] : nullptr, dtype, u.Dim()));
        for (Long i = 0; i < u.Dim(); i++) U_dev[t0 * 3 + i] = u[i];
        for (void* p : {dxt, dxs, dxn, dfs, dfd, du}) sctl_b200_buf_destroy(p);
      }
      {  // (3) the multi-rank registry: every rank hands in ITS particles, Eval returns ITS potentials (fmm-wrapper.txx:490-562)
        sctl::Stokes3D_FxU sl;
        sctl::Stokes3D_DxU dl;
        sctl_b200::ParticleFMM<Real, 3> fmm;
        fmm.SetComm(comm);
        fmm.SetKernels(sl, sl, sl);
        fmm.AddTrg("U", sl, sl);
        fmm.AddSrc("SL", sl, sl);
        fmm.AddSrc("DL", dl, dl);
        fmm.SetKernelS2T("SL", "U", sl);
        fmm.SetKernelS2T("DL", "U", dl);
        fmm.SetTrgCoord("U", xt);
        fmm.SetSrcCoord("SL", xs);
        fmm.SetSrcCoord("DL", xs, xn);
        fmm.SetSrcDensity("SL", fsl);
        fmm.SetSrcDensity("DL", fdl);
        sctl::Vector<Real> u;
        fmm.Eval(u, "U");
        if (u.Dim() != (t1 - t0) * 3) fails[rank]++;
        for (Long i = 0; i < u.Dim(); i++) U_fmm[t0 * 3 + i] = u[i];
        // solver-loop pattern: only the density changes (boundary_integral.txx:1054); 2*F_dl then back to check against U_both
        sctl::Vector<Real> f2 = fdl;
        for (auto& a : f2) a *= 2;
        fmm.SetSrcDensity("DL", f2);
        fmm.Eval(u, "U");
        fmm.SetSrcDensity("DL", fdl);
        u = 7;
        fmm.Eval(u, "U");
        for (Long i = 0; i < u.Dim(); i++) U_fmm2[t0 * 3 + i] = u[i];
      }
      {  // (4) the fused ComputeFarField glue in the multi-rank registry: quadrature weights on the sources (gathered like the
         //     density) and target normals (re-partitioned with the targets when ownership is uneven), Stokes FxT: 9 -> 3 values
        sctl::Stokes3D_FxT fxt;
        sctl_b200::ParticleFMM<Real, 3> fmm;
        fmm.SetComm(comm);
        fmm.SetKernels(fxt, fxt, fxt);
        fmm.AddTrg("T", fxt, fxt);
        fmm.AddSrc("S", fxt, fxt);
        fmm.SetKernelS2T("S", "T", fxt);
        fmm.SetTrgCoord("T", xt);
        fmm.SetSrcCoord("S", xs);
        fmm.SetSrcDensity("S", fsl);
        const sctl::Vector<Real> w = slice(Wq, s0, s1, 1), nt = slice(Xnt, t0, t1, 3);
        fmm.SetSrcWeights("S", w);
        fmm.SetTrgNormal("T", nt);
        sctl::Vector<Real> u;
        fmm.Eval(u, "T");
        if (u.Dim() != (t1 - t0) * 3) fails[rank]++;
        for (Long i = 0; i < u.Dim(); i++) U_glue[t0 * 3 + i] = u[i];
      }
      sctl_b200_comm_destroy(comm);
      // no sctl_b200_finalize here: the thread's streams, events and staging blocks are released when the thread exits
    });
  }
  for (auto& t : th) t.join();
  int nf = 0;
  for (int f : fails) nf += f;
  report(tag + " every rank completed", (double)nf, 0.0);
  if (nf) return;
  report(tag + " eval_sharded host pointers, counts exchanged (Stokes FxU)", rel_l2(U_host, Uref_sl), tol);
  report(tag + " eval_sharded device pointers, FxU then DxU accumulated", rel_l2(U_dev, Uref_both), tol);
  report(tag + " sctl_b200::ParticleFMM + SetComm: DL + SL -> U", rel_l2(U_fmm, Uref_both), tol);
  report(tag + " sctl_b200::ParticleFMM + SetComm after two density updates", rel_l2(U_fmm2, Uref_both), tol);
  report(tag + " sctl_b200::ParticleFMM + SetComm: Stokes FxT with source weights and target normals", rel_l2(U_glue, Uref_glue), 4 * tol);
  {  // SCTL_B200_ALL_GPUS through the plain host-pointer entry (what -DSCTL_B200_SHIM_FLAGS=SCTL_B200_ALL_GPUS makes the shim pass)
    sctl::Vector<Real> u(Nt * 3);
    u = 0;
    const int rc = sctl_b200_eval(SCTL_B200_STOKES3D_FXU, dtype, Nt, Ns, &Xt[0], &Xs[0], nullptr, &Fsl[0], &u[0], -1, SCTL_B200_ALL_GPUS);
    if (rc) std::printf("ALL_GPUS: %s\n", sctl_b200_last_error());
    report(tag + " sctl_b200_eval(SCTL_B200_ALL_GPUS), in-process sharding over every GPU", rc ? 1e300 : rel_l2(u, Uref_sl), tol);
  }
}

int main(int argc, char** argv) {
  int ndev = 0;
  if (sctl_b200_device_count(&ndev) || ndev < 1) {
    std::printf("no CUDA device\n");
    return 2;
  }
  int G = argc > 1 ? std::atoi(argv[1]) : ndev;
  if (G < 1 || G > ndev) G = ndev;
  int ver = 0;
  if (G > 1 && sctl_b200_nccl_version(&ver) == 0) std::printf("NCCL version %d, %d ranks as threads of one process\n", ver, G);
  {
    int64_t b, e;  // the partition formula of fmm-wrapper.txx:507
    sctl_b200_partition(10, 1, 3, &b, &e);
    report("sctl_b200_partition(10, 1, 3) = [3, 6)", (b == 3 && e == 6) ? 0.0 : 1.0, 0.0);
  }
  run<double>(G, 6007, 9001, false, 1e-12);
  run<double>(G, 5000, 7003, true, 1e-12);   // uneven ownership: ragged source blocks, re-partitioned targets
  run<float>(G, 3001, 4000, true, 1e-5);
  run<double>(G, 3, 5, false, 1e-12);        // fewer particles than ranks on some ranks: empty blocks
  std::printf(g_fail ? "SHARDED TEST FAILED (%d)\n" : "SHARDED TEST PASSED\n", g_fail);
  return g_fail ? 1 : 0;
}
