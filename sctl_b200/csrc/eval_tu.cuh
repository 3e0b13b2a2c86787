// Included by eval_f64.cu / eval_f32.cu with SCTL_B200_REAL and SCTL_B200_TABLE_FN defined:
// instantiates pack/eval/reduce/kmat for all eight micro-kernels and fills the launch table.
#include <cstdlib>

#include "kmat_impl.cuh"
#include "launch.h"
#include "small_impl.cuh"

namespace sctl_b200 {
namespace {

using Real = SCTL_B200_REAL;
#ifndef SCTL_B200_RBIG
#define SCTL_B200_RBIG 6
#endif
#if !defined(SCTL_B200_RBIG_F32) && SCTL_B200_RBIG == 6
#define SCTL_B200_RBIG_F32 10
#endif
#ifndef SCTL_B200_MINB_BIG
#define SCTL_B200_MINB_BIG 2
#endif
#ifndef SCTL_B200_ONLY_KERNEL
#define SCTL_B200_ONLY_KERNEL -1  // >= 0: instantiate only that kernel id (fast variant builds for tuning sweeps)
#endif
constexpr int RBIG = SCTL_B200_RBIG;  // throughput variant (profiles/r01d_variant_sweep.txt: R 6, unroll 1, 256-source tiles)
constexpr int RSMALL = 1;             // small-Nt variant
constexpr bool IS_F32 = sizeof(Real) == 4;

template <class Ker>
cudaError_t pack_t(const void* r_src, const void* n_src, const void* v_src, const void* src_wts, void* rec, long long Ns, long long Ns_pad,
                   cudaStream_t st) {
  const unsigned blocks = (unsigned)((Ns_pad + 255) / 256);
  pack_kernel<Ker, Real><<<blocks, 256, 0, st>>>((const Real*)r_src, (const Real*)n_src, (const Real*)v_src, (const Real*)src_wts, (Real*)rec, Ns,
                                                 Ns_pad);
  count_launch();
  return cudaGetLastError();
}

template <int R> constexpr int minb_for() { return R > 1 ? SCTL_B200_MINB_BIG : 8; }
// targets per thread of the throughput variant: kernels with more than 4 accumulators per target (Stokes FxT: 6)
// take 4 so that coordinates + per-tile + running accumulators stay in registers
// FP32 (R sweep of profiles/r02j_variant_sweep.txt: R = 6/7/8/10/12 -> 1.444/1.448/1.450/1.479/1.443e12 pairs/s for Stokes FxU,
// 221 registers at R = 10, spills at 12): 10 where the accumulators and the payload leave room, else as in f64
template <class Ker> constexpr int rbig_for() {
#ifdef SCTL_B200_RBIG_F32
  if (IS_F32 && Ker::NACC <= 3 && Ker::PAY <= 4) return SCTL_B200_RBIG_F32;
#endif
  return (Ker::NACC > 4 && RBIG > 4) ? 4 : RBIG;
}

template <class Ker, int R, int NEWTON>
cudaError_t eval_t(const EvalParams& p, long long n_ttiles, cudaStream_t st) {
  auto kern = eval_kernel<Ker, Real, R, NEWTON, minb_for<R>()>;
  constexpr size_t smem = eval_smem_bytes<Ker, Real>();
  static thread_local int attr_dev = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_dev = dev;
  }
  dim3 grid((unsigned)n_ttiles, (unsigned)p.n_chunks, 1);
  kern<<<grid, EVAL_THREADS, smem, st>>>(p);
  count_launch();
  return cudaGetLastError();
}

template <class Ker, int R, int NEWTON>
int occ_t() {
  static thread_local int cached_dev = -1, cached = 1;  // queried once per host thread and device
  int dev = 0;
  cudaGetDevice(&dev);
  if (cached_dev == dev) return cached;
  auto kern = eval_kernel<Ker, Real, R, NEWTON, minb_for<R>()>;
  constexpr size_t smem = eval_smem_bytes<Ker, Real>();
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, EVAL_THREADS, smem) != cudaSuccess || n < 1) n = 1;
  cached_dev = dev;
  cached = n;
  return n;
}

template <class Ker, int NEWTON, int TPW>
cudaError_t small_launch(const void* r_trg, const void* r_src, const void* n_src, const void* v_src, const void* src_wts, void* v_trg, long long Nt,
                         long long Ns, double scale, int accumulate, cudaStream_t st) {
  const long long blocks = (Nt + TPW - 1) / TPW;
  if (blocks > 0x7fffffffLL) return cudaErrorInvalidValue;
  eval_small_kernel<Ker, Real, NEWTON, TPW><<<(unsigned)blocks, SMALL_THREADS, 0, st>>>((const Real*)r_trg, (const Real*)r_src, (const Real*)n_src,
                                                                                     (const Real*)v_src, (const Real*)src_wts, (Real*)v_trg, Nt,
                                                                                     Ns, scale, accumulate);
  count_launch();
  return cudaGetLastError();
}
// targets per CTA by target count (measured, profiles/r01w_small_tpw_sweep.txt): 8 below ~3.5k targets, 16 below ~14k, else 32
template <class Ker, int NEWTON>
cudaError_t small_t(const void* r_trg, const void* r_src, const void* n_src, const void* v_src, const void* src_wts, void* v_trg, long long Nt,
                    long long Ns, double scale, int accumulate, int sms, cudaStream_t st) {
  if (const char* e = getenv("SCTL_B200_SMALL_TPW")) {  // tuning override
    const int tpw = atoi(e);
    if (tpw == 32) return small_launch<Ker, NEWTON, 32>(r_trg, r_src, n_src, v_src, src_wts, v_trg, Nt, Ns, scale, accumulate, st);
    if (tpw == 16) return small_launch<Ker, NEWTON, 16>(r_trg, r_src, n_src, v_src, src_wts, v_trg, Nt, Ns, scale, accumulate, st);
    if (tpw == 8) return small_launch<Ker, NEWTON, 8>(r_trg, r_src, n_src, v_src, src_wts, v_trg, Nt, Ns, scale, accumulate, st);
  }
  if (Nt >= 96ll * sms) return small_launch<Ker, NEWTON, 32>(r_trg, r_src, n_src, v_src, src_wts, v_trg, Nt, Ns, scale, accumulate, st);
  if (Nt >= 24ll * sms) return small_launch<Ker, NEWTON, 16>(r_trg, r_src, n_src, v_src, src_wts, v_trg, Nt, Ns, scale, accumulate, st);
  return small_launch<Ker, NEWTON, 8>(r_trg, r_src, n_src, v_src, src_wts, v_trg, Nt, Ns, scale, accumulate, st);
}

template <class Ker>
cudaError_t reduce_t(const double* partial, void* v_trg, long long Nt, long long nt_stride, int n_chunks, double scale, int accumulate,
                     cudaStream_t st) {
  const unsigned blocks = (unsigned)((Nt + 127) / 128);
  reduce_kernel<Real, Ker::K1><<<blocks, 128, 0, st>>>(partial, (Real*)v_trg, Nt, nt_stride, n_chunks, scale, accumulate);
  count_launch();
  return cudaGetLastError();
}

template <class Ker, int NEWTON>
cudaError_t kmat_t(const void* r_trg, const void* r_src, const void* n_src, void* M, long long Nt, long long Ns, double scale, int sb, cudaStream_t st) {
  constexpr int CT = kmat_cta_targets<Ker, Real, false>();
  const long long gx = (Nt + CT - 1) / CT;
  const long long gy = (Ns + sb - 1) / sb;
  if (gy > 65535) return cudaErrorInvalidValue;  // callers slab the source range
  dim3 grid((unsigned)gx, (unsigned)gy, 1);
  kmat_kernel<Ker, Real, NEWTON><<<grid, KMAT_THREADS, 0, st>>>((const Real*)r_trg, (const Real*)r_src, (const Real*)n_src, (Real*)M, Nt, Ns,
                                                             (Real)scale, sb);
  count_launch();
  return cudaGetLastError();
}

template <class Ker, int NEWTON, bool DOTN>
cudaError_t kmat_batch_t(const void* r_trg, const void* r_src, const void* n_src, const void* src_wts, const void* n_trg, void* M,
                         const KmatBatchDesc& D, long long n_ctas, double scale, cudaStream_t st) {
  // 1-D grid; walk it in slabs of 2^30 CTAs (far beyond any real batch, but the limit is 2^31-1)
  for (long long c0 = 0; c0 < n_ctas; c0 += (1ll << 30)) {
    const long long n = n_ctas - c0 < (1ll << 30) ? n_ctas - c0 : (1ll << 30);
    kmat_batch_kernel<Ker, Real, NEWTON, DOTN><<<(unsigned)n, KMAT_THREADS, 0, st>>>((const Real*)r_trg, (const Real*)r_src, (const Real*)n_src,
                                                                                  (const Real*)src_wts, (const Real*)n_trg, (Real*)M, D, c0, (Real)scale);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}
template <class Ker, bool HAS> struct KmatDotN {
  static void fill(EvalEntry& e) {
    e.kmat_cta_targets[1] = kmat_cta_targets<Ker, Real, true>();
    e.kmat_batch[1][0] = kmat_batch_t<Ker, 0, true>;
    e.kmat_batch[1][1] = kmat_batch_t<Ker, 1, true>;
    e.kmat_batch[1][2] = kmat_batch_t<Ker, IS_F32 ? 1 : 2, true>;
  }
};
template <class Ker> struct KmatDotN<Ker, false> {
  static void fill(EvalEntry& e) {
    e.kmat_cta_targets[1] = 0;
    e.kmat_batch[1][0] = e.kmat_batch[1][1] = e.kmat_batch[1][2] = nullptr;
  }
};

template <class Ker, int ID>
EvalEntry make_entry() {
  EvalEntry e{};
  if constexpr (SCTL_B200_ONLY_KERNEL >= 0 && SCTL_B200_ONLY_KERNEL != ID) {
    return e;
  } else {
  e.rec_len = rec_len<Ker, Real>();
  e.k0 = Ker::K0;
  e.k1 = Ker::K1;
  e.nd = Ker::ND;
  e.R[0] = RSMALL;
  e.R[1] = rbig_for<Ker>();
  e.x2_shift[0] = eval_x2<Real, 0>() ? Ker::X2_SHIFT : 0;
  e.x2_shift[1] = eval_x2<Real, 1>() ? Ker::X2_SHIFT : 0;
  e.x2_shift[2] = eval_x2<Real, (IS_F32 ? 1 : 2)>() ? Ker::X2_SHIFT : 0;
  e.pack = pack_t<Ker>;
  e.reduce = reduce_t<Ker>;
  // f32 has refinement levels 0 and 1 only; level 2 aliases level 1
  e.eval[0][0] = eval_t<Ker, RSMALL, 0>;
  e.eval[0][1] = eval_t<Ker, RSMALL, 1>;
  e.eval[0][2] = eval_t<Ker, RSMALL, IS_F32 ? 1 : 2>;
  e.eval[1][0] = eval_t<Ker, rbig_for<Ker>(), 0>;
  e.eval[1][1] = eval_t<Ker, rbig_for<Ker>(), 1>;
  e.eval[1][2] = eval_t<Ker, rbig_for<Ker>(), IS_F32 ? 1 : 2>;
  e.occupancy[0][0] = occ_t<Ker, RSMALL, 0>;
  e.occupancy[0][1] = occ_t<Ker, RSMALL, 1>;
  e.occupancy[0][2] = occ_t<Ker, RSMALL, IS_F32 ? 1 : 2>;
  e.occupancy[1][0] = occ_t<Ker, rbig_for<Ker>(), 0>;
  e.occupancy[1][1] = occ_t<Ker, rbig_for<Ker>(), 1>;
  e.occupancy[1][2] = occ_t<Ker, rbig_for<Ker>(), IS_F32 ? 1 : 2>;
  e.small[0] = small_t<Ker, 0>;
  e.small[1] = small_t<Ker, 1>;
  e.small[2] = small_t<Ker, IS_F32 ? 1 : 2>;
  e.kmat[0] = kmat_t<Ker, 0>;
  e.kmat[1] = kmat_t<Ker, 1>;
  e.kmat[2] = kmat_t<Ker, IS_F32 ? 1 : 2>;
  e.kmat_cta_targets[0] = kmat_cta_targets<Ker, Real, false>();
  e.kmat_batch[0][0] = kmat_batch_t<Ker, 0, false>;
  e.kmat_batch[0][1] = kmat_batch_t<Ker, 1, false>;
  e.kmat_batch[0][2] = kmat_batch_t<Ker, IS_F32 ? 1 : 2, false>;
  KmatDotN<Ker, Ker::K1 % 3 == 0>::fill(e);
  return e;
  }
}

}  // namespace

cudaError_t SCTL_B200_CONTRACT_FN(const void* Ufull, const void* n_trg, void* U, long long Nt, int K1out, cudaStream_t st) {
  const long long n = Nt * K1out;
  if (n == 0) return cudaSuccess;
  contract_normal_kernel<Real><<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const Real*)Ufull, (const Real*)n_trg, (Real*)U, Nt, K1out);
  count_launch();
  return cudaGetLastError();
}

const EvalEntry* SCTL_B200_TABLE_FN() {
  static const EvalEntry table[NUM_KERNELS] = {
      make_entry<KLaplaceFxU, 0>(), make_entry<KLaplaceDxU, 1>(), make_entry<KLaplaceFxdU, 2>(), make_entry<KStokesFxU, 3>(),
      make_entry<KStokesDxU, 4>(),  make_entry<KStokesFxT, 5>(),  make_entry<KStokesFSxU, 6>(),  make_entry<KStokesFxUP, 7>(),
  };
  return table;
}

}  // namespace sctl_b200
