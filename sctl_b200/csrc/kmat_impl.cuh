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
//   * source coordinates / normals are warp-uniform read-only loads (L1 broadcast).
#pragma once
#include "ukernels.cuh"

namespace sctl_b200 {

constexpr int KMAT_WARPS = 4;
constexpr int KMAT_THREADS = KMAT_WARPS * 32;
constexpr int KMAT_SB = 32;  // sources per CTA

template <class Ker, class Real, int NEWTON>
__global__ void __launch_bounds__(KMAT_THREADS) kmat_kernel(const Real* __restrict__ r_trg, const Real* __restrict__ r_src,
                                                            const Real* __restrict__ n_src, Real* __restrict__ M, long long Nt,
                                                            long long Ns, Real scale) {
  constexpr int K0 = Ker::K0, K1 = Ker::K1, ND = Ker::ND;
  constexpr int SPAN = 32 * K1;  // elements per row owned by one warp
  __shared__ Real slab[KMAT_WARPS][K0][SPAN];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const long long t0 = ((long long)blockIdx.x * KMAT_WARPS + warp) * 32;  // first target of this warp
  if (t0 >= Nt) return;                                                   // whole warp out of range (no CTA-wide sync below)
  const long long t = t0 + lane;
  const long long tc = t < Nt ? t : Nt - 1;
  const Real xt = r_trg[tc * 3 + 0], yt = r_trg[tc * 3 + 1], zt = r_trg[tc * 3 + 2];
  const long long span_valid = min((long long)SPAN, (Nt - t0) * K1);  // valid elements of this warp's span
  const long long ld = Nt * K1;

  const long long s_begin = (long long)blockIdx.y * KMAT_SB;
  const long long s_end = min(Ns, s_begin + KMAT_SB);
  for (long long s = s_begin; s < s_end; s++) {
    const Real xs = __ldg(r_src + s * 3 + 0), ys = __ldg(r_src + s * 3 + 1), zs = __ldg(r_src + s * 3 + 2);
    Real n[3] = {0, 0, 0};
#pragma unroll
    for (int i = 0; i < ND; i++) n[i] = __ldg(n_src + s * ND + i);
    const Real dx = xt - xs, dy = yt - ys, dz = zt - zs;
    Real r2 = dx * dx;
    r2 = fma(dy, dy, r2);
    r2 = fma(dz, dz, r2);
    const Real rinv = rsqrt_masked<NEWTON>(r2);
    Real u[K0][K1];
    Ker::matrix(u, dx, dy, dz, rinv, n);

    __syncwarp();  // previous iteration's reads of the slab are complete
#pragma unroll
    for (int k0 = 0; k0 < K0; k0++) {
#pragma unroll
      for (int k1 = 0; k1 < K1; k1++) slab[warp][k0][lane * K1 + k1] = u[k0][k1] * scale;
    }
    __syncwarp();
    Real* row = M + (size_t)(s * K0) * (size_t)ld + (size_t)t0 * K1;
#pragma unroll
    for (int k0 = 0; k0 < K0; k0++) {
#pragma unroll
      for (int j = 0; j < K1; j++) {
        const int e = j * 32 + lane;
        if (e < span_valid) row[(size_t)k0 * (size_t)ld + e] = slab[warp][k0][e];
      }
    }
  }
}

}  // namespace sctl_b200
