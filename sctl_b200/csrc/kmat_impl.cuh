// Dense kernel-matrix assembly.  Replaces GenericKernel<uKernel>::KernelMatrix
// (reference: include/sctl/generic-kernel.txx:191-307):
//     M[(s*K0+k0) * (Nt*K1) + (t*K1+k1)] = scale * K[k0][k1](x_t - x_s, n_s)        row-major, overwritten.
//
// HBM-WRITE-bound: sizeof(Real)*K0*K1 bytes per pair against ~20-50 flops. Mapping:
//   * a lane owns one target (coordinates in registers), a warp owns 32 consecutive targets, i.e. a
//     contiguous span of 32*K1 elements in every row of M; a CTA (KMAT_WARPS warps) owns a block of
//     KMAT_SB consecutive sources (blockIdx.y) and streams over them;
//   * per source, each lane evaluates the K0 x K1 block once (no redundant arithmetic), the warp
//     transposes it through a PRIVATE shared-memory slab (only __syncwarp, never __syncthreads) and
//     writes each of the K0 rows as fully coalesced stores — consecutive lanes write consecutive
//     elements, so every store instruction covers one contiguous 32*sizeof(Real)-byte run and HBM sees
//     whole 32-byte sectors;
//   * the CTA's source coordinates / normals / row weights are staged ONCE in shared memory (coalesced) and read back as
//     warp-uniform LDS broadcasts: with per-iteration global loads the loop sat in long-scoreboard stalls (6.9 per issue,
//     profiles/r01z_kmat.metrics.txt) waiting for the next source's coordinates.
#pragma once
#include "ukernels.cuh"

namespace sctl_b200 {

constexpr int KMAT_WARPS = 4;
constexpr int KMAT_THREADS = KMAT_WARPS * 32;
constexpr int KMAT_SB = 32;  // sources per CTA for large matrices; the launcher passes 16 / 8 / 4 when the grid would otherwise hold
                             // only a few waves of CTAs (a 256 MB matrix is a 40-90 us kernel: one ragged wave is 15 % of it)

// Kernels whose rows hold ONE value per target (Laplace FxU / DxU, and every target-normal contraction of a 3-vector) would
// write 4- or 8-byte elements per lane: there a lane owns V = 16 B / sizeof(Real) consecutive targets and stores them as one
// vector (512 B per warp-wide store instead of 128 / 256 B), which also gives V independent pairs per lane.
template <class Ker, class Real, bool DOTN>
__host__ __device__ constexpr int kmat_vec() {
  return ((DOTN ? Ker::K1 / 3 : Ker::K1) == 1) ? (int)(16 / sizeof(Real)) : 1;
}
template <class Ker, class Real, bool DOTN>
__host__ __device__ constexpr int kmat_cta_targets() {  // targets per CTA (blockIdx.x granularity)
  return KMAT_THREADS * kmat_vec<Ker, Real, DOTN>();
}

// One CTA's share of one dense block: targets [tile_x*128, +128) x sources [slab_y*32, +32) of an Nt x Ns problem.
//   DOTN = false: M[(s*K0+k0)*ld + t*K1+k1]   = w_s * scale * K[k0][k1]                  (KernelMatrix, generic-kernel.txx:284;
//                                                 w_s = src_wts[s] or 1: the row scaling of boundary_integral.txx:990-997)
//   DOTN = true : M[(s*K0+k0)*ld + t*K1'+k1'] = w_s * scale * sum_l K[k0][k1'*3+l] * n_trg[t][l],   K1' = K1/3
//                                                 (the target-normal contraction of boundary_integral.txx:968-984)
// ld = Nt*K1 (or Nt*K1'). slab: KMAT_WARPS * K0 * 32*K1OUT Reals of shared memory.
template <class Ker, class Real, int NEWTON, bool DOTN>
__device__ __forceinline__ void kmat_block(const Real* __restrict__ r_trg, const Real* __restrict__ r_src, const Real* __restrict__ n_src,
                                           const Real* __restrict__ src_wts, const Real* __restrict__ n_trg, Real* __restrict__ M, long long Nt,
                                           long long Ns, long long tile_x, long long slab_y, Real scale, Real* slab_all, Real* srcs, int sb) {
  constexpr int K0 = Ker::K0, K1 = Ker::K1, ND = Ker::ND;
  constexpr int K1OUT = DOTN ? K1 / 3 : K1;
  constexpr int SPAN = 32 * K1OUT;  // elements per row owned by one warp
  constexpr int SW = 3 + ND + 1;    // staged values per source: x, y, z, normal, weight * scale

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;

  // stage this CTA's sources (every thread takes part: the early exits for out-of-range warps come after the barrier)
  const long long s_begin = slab_y * sb;
  const long long s_end = min(Ns, s_begin + sb);
  const int ns_cta = (int)(s_end - s_begin);
  for (int i = threadIdx.x; i < ns_cta * 3; i += KMAT_THREADS) {
    const int j = i / 3;
    srcs[j * SW + (i - j * 3)] = r_src[s_begin * 3 + i];
  }
  for (int i = threadIdx.x; i < ns_cta * ND; i += KMAT_THREADS) {
    const int j = i / (ND ? ND : 1);
    srcs[j * SW + 3 + (i - j * (ND ? ND : 1))] = n_src[s_begin * ND + i];
  }
  for (int j = threadIdx.x; j < ns_cta; j += KMAT_THREADS) srcs[j * SW + 3 + ND] = src_wts ? src_wts[s_begin + j] * scale : scale;
  __syncthreads();

  if constexpr (K1OUT == 1) {
    constexpr int V = kmat_vec<Ker, Real, DOTN>();
    using Vec = typename VecOf<Real>::type;
    const long long t0 = (tile_x * KMAT_WARPS + warp) * (32 * V);
    if (t0 >= Nt) return;
    const long long tl = t0 + (long long)lane * V;  // first of this lane's V consecutive targets
    const int nvalid = (int)max(0ll, min((long long)V, Nt - tl));
    Real xt[V], yt[V], zt[V], nt[V][3];
#pragma unroll
    for (int j = 0; j < V; j++) {
      const long long tc = tl + j < Nt ? tl + j : Nt - 1;
      xt[j] = r_trg[tc * 3 + 0];
      yt[j] = r_trg[tc * 3 + 1];
      zt[j] = r_trg[tc * 3 + 2];
      nt[j][0] = nt[j][1] = nt[j][2] = 0;
      if (DOTN) {
        nt[j][0] = n_trg[tc * 3 + 0];
        nt[j][1] = n_trg[tc * 3 + 1];
        nt[j][2] = n_trg[tc * 3 + 2];
      }
    }
    for (long long s = s_begin; s < s_end; s++) {
      const Real* sp = srcs + (int)(s - s_begin) * SW;
      const Real xs = sp[0], ys = sp[1], zs = sp[2];
      Real n[3] = {0, 0, 0};
#pragma unroll
      for (int i = 0; i < ND; i++) n[i] = sp[3 + i];
      const Real w = sp[3 + ND];
      Real v[K0][V];
#pragma unroll
      for (int j = 0; j < V; j++) {
        const Real dx = xt[j] - xs, dy = yt[j] - ys, dz = zt[j] - zs;
        Real r2 = dx * dx;
        r2 = fma(dy, dy, r2);
        r2 = fma(dz, dz, r2);
        const Real rinv = rsqrt_masked<NEWTON>(r2);
        Real u[K0][K1];
        Ker::matrix(u, dx, dy, dz, rinv, n);
#pragma unroll
        for (int k0 = 0; k0 < K0; k0++) {
          if (DOTN) {
            Real a = u[k0][0] * nt[j][0];
            a = fma(u[k0][1 % K1], nt[j][1], a);
            a = fma(u[k0][2 % K1], nt[j][2], a);
            v[k0][j] = a * w;
          } else {
            v[k0][j] = u[k0][0] * w;
          }
        }
      }
#pragma unroll
      for (int k0 = 0; k0 < K0; k0++) {
        Real* dst = M + (size_t)(s * K0 + k0) * (size_t)Nt + (size_t)tl;  // ld = Nt (one value per target)
        if (nvalid == V && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
          *reinterpret_cast<Vec*>(dst) = *reinterpret_cast<const Vec*>(v[k0]);
        } else {
#pragma unroll
          for (int j = 0; j < V; j++)
            if (j < nvalid) dst[j] = v[k0][j];
        }
      }
    }
    return;
  }

  Real(*slab)[SPAN] = reinterpret_cast<Real(*)[SPAN]>(slab_all + (size_t)warp * K0 * SPAN);
  const long long t0 = (tile_x * KMAT_WARPS + warp) * 32;  // first target of this warp
  if (t0 >= Nt) return;                                    // whole warp out of range (no CTA-wide sync below)
  const long long t = t0 + lane;
  const long long tc = t < Nt ? t : Nt - 1;
  const Real xt = r_trg[tc * 3 + 0], yt = r_trg[tc * 3 + 1], zt = r_trg[tc * 3 + 2];
  Real nt[3] = {0, 0, 0};
  if (DOTN) {
    nt[0] = n_trg[tc * 3 + 0];
    nt[1] = n_trg[tc * 3 + 1];
    nt[2] = n_trg[tc * 3 + 2];
  }
  const long long span_valid = min((long long)SPAN, (Nt - t0) * K1OUT);  // valid elements of this warp's span
  const long long ld = Nt * K1OUT;

  for (long long s = s_begin; s < s_end; s++) {
    const Real* sp = srcs + (int)(s - s_begin) * SW;
    const Real xs = sp[0], ys = sp[1], zs = sp[2];
    Real n[3] = {0, 0, 0};
#pragma unroll
    for (int i = 0; i < ND; i++) n[i] = sp[3 + i];
    const Real w = sp[3 + ND];
    const Real dx = xt - xs, dy = yt - ys, dz = zt - zs;
    Real r2 = dx * dx;
    r2 = fma(dy, dy, r2);
    r2 = fma(dz, dz, r2);
    const Real rinv = rsqrt_masked<NEWTON>(r2);
    Real u[K0][K1];
    Ker::matrix(u, dx, dy, dz, rinv, n);
    Real v[K0][K1OUT];
#pragma unroll
    for (int k0 = 0; k0 < K0; k0++) {
#pragma unroll
      for (int k1 = 0; k1 < K1OUT; k1++) {
        if (DOTN) {
          Real a = u[k0][k1 * 3 + 0] * nt[0];
          a = fma(u[k0][k1 * 3 + 1], nt[1], a);
          a = fma(u[k0][k1 * 3 + 2], nt[2], a);
          v[k0][k1] = a * w;
        } else {
          v[k0][k1] = u[k0][k1] * w;
        }
      }
    }
    Real* row = M + (size_t)(s * K0) * (size_t)ld + (size_t)t0 * K1OUT;
    {
      __syncwarp();  // previous iteration's reads of the slab are complete
#pragma unroll
      for (int k0 = 0; k0 < K0; k0++) {
#pragma unroll
        for (int k1 = 0; k1 < K1OUT; k1++) slab[k0][lane * K1OUT + k1] = v[k0][k1];
      }
      __syncwarp();
#pragma unroll
      for (int k0 = 0; k0 < K0; k0++) {
#pragma unroll
        for (int j = 0; j < K1OUT; j++) {
          const int e = j * 32 + lane;
          if (e < span_valid) row[(size_t)k0 * (size_t)ld + e] = slab[k0][e];
        }
      }
    }
  }
}

template <class Ker, class Real, int NEWTON>
__global__ void __launch_bounds__(KMAT_THREADS) kmat_kernel(const Real* __restrict__ r_trg, const Real* __restrict__ r_src,
                                                            const Real* __restrict__ n_src, Real* __restrict__ M, long long Nt,
                                                            long long Ns, Real scale, int sb) {
  __shared__ Real slab[KMAT_WARPS * Ker::K0 * 32 * Ker::K1];
  __shared__ Real srcs[KMAT_SB * (3 + Ker::ND + 1)];
  kmat_block<Ker, Real, NEWTON, false>(r_trg, r_src, n_src, nullptr, nullptr, M, Nt, Ns, blockIdx.x, blockIdx.y, scale, slab, srcs, sb);
}

// ---------------------------------------------------------------------------------------------
// Batched small-block mode: the reference's BoundaryIntegralOp::SetupNear calls KernelMatrix once per
// element on (that element's far-field nodes) x (its near targets) — thousands of tiny blocks per setup
// (boundary_integral.txx:946-1008). Here ALL blocks are one launch: blocks are described CSR-style by point
// displacements into the concatenated coordinate arrays (exactly the reference's near_elem_dsp / elem_nds_dsp_far),
// each CTA finds its block by binary search in the prefix sum of CTAs per block.
// ---------------------------------------------------------------------------------------------
struct KmatBatchDesc {
  const long long* trg_dsp;  // [nbatch+1] points
  const long long* src_dsp;  // [nbatch+1] points
  const long long* M_dsp;    // [nbatch]   element offset of block b in M
  const long long* cta_dsp;  // [nbatch+1] prefix sum of CTAs per block
  long long nbatch;
  int sb;                    // sources per CTA the prefix sum was built with
};

template <class Ker, class Real, int NEWTON, bool DOTN>
__global__ void __launch_bounds__(KMAT_THREADS) kmat_batch_kernel(const Real* __restrict__ r_trg, const Real* __restrict__ r_src,
                                                                  const Real* __restrict__ n_src, const Real* __restrict__ src_wts,
                                                                  const Real* __restrict__ n_trg, Real* __restrict__ M, KmatBatchDesc D,
                                                                  long long cta0, Real scale) {
  __shared__ Real slab[KMAT_WARPS * Ker::K0 * 32 * (DOTN ? Ker::K1 / 3 : Ker::K1)];
  __shared__ Real srcs[KMAT_SB * (3 + Ker::ND + 1)];
  const long long cta = cta0 + blockIdx.x;
  long long lo = 0, hi = D.nbatch;  // block b with cta_dsp[b] <= cta < cta_dsp[b+1]
  while (hi - lo > 1) {
    const long long mid = (lo + hi) >> 1;
    if (D.cta_dsp[mid] <= cta) lo = mid; else hi = mid;
  }
  const long long b = lo;
  const long long Nt = D.trg_dsp[b + 1] - D.trg_dsp[b];
  const long long Ns = D.src_dsp[b + 1] - D.src_dsp[b];
  constexpr int CT = kmat_cta_targets<Ker, Real, DOTN>();
  const long long ntx = (Nt + CT - 1) / CT;
  const long long c = cta - D.cta_dsp[b];
  const long long so = D.src_dsp[b], to = D.trg_dsp[b];
  kmat_block<Ker, Real, NEWTON, DOTN>(r_trg + to * 3, r_src + so * 3, Ker::ND ? n_src + so * Ker::ND : nullptr, src_wts ? src_wts + so : nullptr,
                                      DOTN ? n_trg + to * 3 : nullptr, M + D.M_dsp[b], Nt, Ns, c % ntx, c / ntx, scale, slab, srcs, D.sb);
}

}  // namespace sctl_b200
