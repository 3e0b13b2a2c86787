// One-launch direct summation for SMALL problems (a few thousand points per side): the regime of the per-element calls
// the reference makes from BoundaryIntegralOp and of FMM leaf boxes, where the tiled path (pack -> eval -> reduce, 256
// sources walked serially by every thread) is bound by launch count and by the FP64 dependency chain, not by the pipe.
// Same sum as GenericKernel<uKernel>::Eval (reference: include/sctl/generic-kernel.txx:153-186), different mapping:
//   * ONE kernel: the AoS sources are read and packed into shared-memory records by the CTA itself (no pack launch, no
//     record buffer in HBM), the scale factor and the accumulation into v_trg happen in the same kernel (no reduce launch);
//   * a CTA owns TPW = 32, 16 or 8 targets and its 256 threads split every staged batch of 256 sources 256/TPW ways
//     (lane % TPW = target, the rest of the thread index = source lane), so a thread walks Ns*TPW/256 sources instead of
//     Ns and even a few hundred targets give every SM a CTA; the partial sums of a target are added in a fixed order
//     through shared memory (deterministic) - the split-source reduction of the design;
//   * r2 == 0 is masked in the loop (exactly-zero contribution, kernel_functions.hpp:28) - this path is latency-bound,
//     the unmasked fast path of eval_kernel would buy nothing here.
#pragma once
#include "ukernels.cuh"

namespace sctl_b200 {

constexpr int SMALL_WARPS = 8;                   // source lanes per target
constexpr int SMALL_THREADS = 32 * SMALL_WARPS;  // = sources staged per batch (one per thread)

template <class Ker, class Real, int NEWTON, int TPW>
__global__ void __launch_bounds__(SMALL_THREADS) eval_small_kernel(const Real* __restrict__ r_trg, const Real* __restrict__ r_src,
                                                                  const Real* __restrict__ n_src, const Real* __restrict__ v_src,
                                                                  const Real* __restrict__ src_wts, Real* __restrict__ v_trg, long long Nt,
                                                                  long long Ns, double scale, int accumulate) {
  constexpr int REC = rec_len<Ker, Real>();
  constexpr int NACC = Ker::NACC;
  constexpr int K1 = Ker::K1;
  using V = typename VecOf<Real>::type;
  constexpr int NV = REC / VecOf<Real>::N;
  __shared__ __align__(16) Real recs[SMALL_THREADS * REC];
  constexpr int SL = SMALL_THREADS / TPW;  // source lanes per target
  __shared__ double red[NACC][SMALL_THREADS];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tl = lane % TPW, sl = warp * (32 / TPW) + lane / TPW;
  const long long t = (long long)blockIdx.x * TPW + tl;
  const long long tc = t < Nt ? t : Nt - 1;  // clamp: out-of-range lanes recompute the last target and are not stored
  const Real xt = r_trg[tc * 3 + 0], yt = r_trg[tc * 3 + 1], zt = r_trg[tc * 3 + 2];
  double acc[NACC];
#pragma unroll
  for (int q = 0; q < NACC; q++) acc[q] = 0.0;

  for (long long s0 = 0; s0 < Ns; s0 += SMALL_THREADS) {
    const int nb = (int)(Ns - s0 < SMALL_THREADS ? Ns - s0 : SMALL_THREADS);
    if (s0) __syncthreads();  // the previous batch has been consumed
    if (tid < nb) {
      const long long s = s0 + tid;
      Real out[REC], n[3] = {0, 0, 0}, f[Ker::K0];
#pragma unroll
      for (int i = 0; i < REC; i++) out[i] = (Real)0;
      out[0] = r_src[s * 3 + 0];
      out[1] = r_src[s * 3 + 1];
      out[2] = r_src[s * 3 + 2];
#pragma unroll
      for (int i = 0; i < Ker::ND; i++) n[i] = n_src[s * Ker::ND + i];
#pragma unroll
      for (int i = 0; i < Ker::K0; i++) f[i] = v_src[s * Ker::K0 + i];
      if (src_wts) {
        const Real w = src_wts[s];
#pragma unroll
        for (int i = 0; i < Ker::K0; i++) f[i] *= w;
      }
      Ker::pack(out + 3, n, f);
      V* dst = reinterpret_cast<V*>(recs + tid * REC);
#pragma unroll
      for (int i = 0; i < NV; i++) dst[i] = reinterpret_cast<const V*>(out)[i];
    }
    __syncthreads();

    Real tacc[NACC];
#pragma unroll
    for (int q = 0; q < NACC; q++) tacc[q] = (Real)0;
    const V* tile = reinterpret_cast<const V*>(recs);
#pragma unroll 4
    for (int s = sl; s < nb; s += SL) {
      Real src[REC];
#pragma unroll
      for (int i = 0; i < NV; i++) reinterpret_cast<V*>(src)[i] = tile[s * NV + i];  // 32/TPW distinct records per warp, broadcast
      const Real dx = xt - src[0];
      const Real dy = yt - src[1];
      const Real dz = zt - src[2];
      Real r2 = dx * dx;
      r2 = fma(dy, dy, r2);
      r2 = fma(dz, dz, r2);
      const Real rinv = rsqrt_masked<NEWTON>(r2);
      Ker::apply(tacc, dx, dy, dz, rinv, src + 3);
    }
#pragma unroll
    for (int q = 0; q < NACC; q++) acc[q] += (double)tacc[q];
  }

#pragma unroll
  for (int q = 0; q < NACC; q++) red[q][sl * TPW + tl] = acc[q];
  __syncthreads();
  if (tid < TPW && t < Nt) {  // tid < TPW: tl == tid, sl == 0
    double a[NACC], out[K1];
#pragma unroll
    for (int q = 0; q < NACC; q++) {
      double sum = red[q][tid];
#pragma unroll 8
      for (int w = 1; w < SL; w++) sum += red[q][w * TPW + tid];
      a[q] = sum;
    }
    Ker::finalize(out, a);
    Real* v = v_trg + t * K1;
#pragma unroll
    for (int q = 0; q < K1; q++) {
      const double prev = accumulate ? (double)v[q] : 0.0;
      v[q] = (Real)(prev + out[q] * scale);
    }
  }
}

}  // namespace sctl_b200
