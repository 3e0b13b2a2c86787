// Shared device helpers: mbarrier / 1-D bulk-TMA wrappers and the masked reciprocal square roots.
// sm_100a only (cp.async.bulk + mbarrier expect_tx need sm_90+, the build passes -gencode arch=compute_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sctl_b200 {

// ---------------------------------------------------------------------------------------------
// mbarrier + bulk async copy (TMA 1-D, SASS: UBLKCP) — PTX ISA "cp.async.bulk", "mbarrier".
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "LAB_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra LAB_DONE_%=;\n"
      "bra LAB_WAIT_%=;\n"
      "LAB_DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared bulk copy, completion signalled on `bar` (bytes must be a multiple of 16, both
// addresses 16-byte aligned).
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ---------------------------------------------------------------------------------------------
// rinv = (r2 > 0) ? 1/sqrt(r2) : 0      — the arithmetic of approx_rsqrt<digits>(r2, r2 > 0)
// (reference: kernel_functions.hpp:28, vec.txx:361-364, intrin-wrapper.hpp:516-535,3025-3050).
//
// The reference refines a 14-bit hardware seed with 0/1/2 Newton steps chosen from `digits`.
// Here the seed is MUFU.RSQ64H (PTX rsqrt.approx.ftz.f64: ~22 good bits, computed from the high
// word only) and NEWTON selects the refinement on the FP64 pipe:
//   NEWTON 0: seed only                       (rel. err ~2^-22  -> digits <= 6)
//   NEWTON 1: one quadratic Newton step       (rel. err ~1.5*2^-44 -> digits <= 12), 3 FP64 ops
//   NEWTON 2: one cubic (Halley-type) step    (rel. err ~1 ulp),  5 FP64 ops
//       e = 1 - x*y0^2 ;  y1 = y0 * (1 + e/2 + 3e^2/8)     [1/sqrt(1-e) to second order]
// The mask is the reference's: r2 > 0 (kernel_functions.hpp:28) - r2 == 0 gives exactly 0, every positive r2 a finite
// rinv, SUBNORMAL r2 included (the reference's seed instruction handles denormals, intrin-wrapper.hpp:3038-3049;
// MUFU.RSQ flushes them, so they are rescaled by an exact power of four first: rsqrt(r2) = 2^k rsqrt(4^k r2)).
// The test is an integer compare on the high word (ALU pipe), the FP64 pipe sees no compare/select; the rare
// zero / subnormal case is a branch.
// ---------------------------------------------------------------------------------------------
template <int NEWTON>
__device__ __forceinline__ double rsqrt_pos_normal(double r2) {  // r2 positive and normal
  double seed;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(seed) : "d"(r2));
  const int hi = __double2hiint(seed);
  const double y0 = __hiloint2double(hi, 0);
  if (NEWTON == 0) return y0;
  const double t = r2 * y0;
  const double e = fma(-t, y0, 1.0);
  if (NEWTON == 1) {
    // y0/2 by an exponent decrement on the integer pipe (y0 is a positive normal number here)
    const double h = __hiloint2double(hi - 0x00100000, 0);
    return fma(h, e, y0);
  }
  const double p = fma(e, 0.375, 0.5);
  const double q = e * p;
  return fma(y0, q, y0);
}
template <int NEWTON>
__device__ __forceinline__ double rsqrt_masked(double r2) {
  if (__double2hiint(r2) < 0x00100000) {  // zero or subnormal (r2 is a sum of squares: never negative)
    if (!(r2 > 0.0)) return 0.0;          // the r2 > 0 mask: a coincident pair contributes exactly 0
    return rsqrt_pos_normal<NEWTON>(r2 * 0x1p108) * 0x1p54;
  }
  return rsqrt_pos_normal<NEWTON>(r2);
}
// Branch-free form for the tile-repair loop of eval_kernel: zero AND subnormal r2 give 0 here (the seed is masked with integer
// selects, as the reference masks its seed, intrin-wrapper.hpp:3045), and `subnormal` is raised when a POSITIVE subnormal r2
// was seen, in which case the caller repeats the tile with rsqrt_masked above. Costs ISETP + SEL + a few predicate ops per
// pair instead of a divergent branch: the repair of a tile that holds coincident points (every tile on the diagonal of a
// self-interaction) then runs close to the speed of the unmasked loop instead of 4x slower.
// Branch-free AND complete: the r2 > 0 mask with subnormal r2 handled by selects - the seed is taken of r2 * 2^108 when r2 is
// below the normal range (exact scaling; MUFU flushes subnormals) and multiplied by 2^54 afterwards, and it is zeroed only for
// r2 == 0. One loop, no flag, no second pass; a few more integer / FP64 instructions per pair than rsqrt_masked_flag, which
// only the repair path of eval_kernel pays.
template <int NEWTON>
__device__ __forceinline__ double rsqrt_masked_sel(double r2) {
  const int rh = __double2hiint(r2);
  const bool ok = rh >= 0x00100000;                    // positive normal number
  const bool nonzero = (rh | __double2loint(r2)) != 0;  // r2 is a sum of squares: never negative
  const double xs = ok ? r2 : r2 * 0x1p108;
  double seed;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(seed) : "d"(xs));
  double y0 = __hiloint2double(nonzero ? __double2hiint(seed) : 0, 0);
  y0 = ok ? y0 : y0 * 0x1p54;
  if (NEWTON == 0) return y0;
  const double t = r2 * y0;
  const double e = fma(-t, y0, 1.0);
  if (NEWTON == 1) return fma(0.5 * y0, e, y0);
  const double p = fma(e, 0.375, 0.5);
  const double q = e * p;
  return fma(y0, q, y0);
}
template <int NEWTON>
__device__ __forceinline__ double rsqrt_masked_flag(double r2, bool& subnormal) {
  double seed;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(seed) : "d"(r2));
  const int rh = __double2hiint(r2);
  const bool ok = rh >= 0x00100000;  // positive normal number
  subnormal |= (!ok) & ((rh | __double2loint(r2)) != 0);
  const int hi = ok ? __double2hiint(seed) : 0;
  const double y0 = __hiloint2double(hi, 0);
  if (NEWTON == 0) return y0;
  const double t = r2 * y0;
  const double e = fma(-t, y0, 1.0);
  if (NEWTON == 1) {
    const double h = __hiloint2double(ok ? hi - 0x00100000 : 0, 0);
    return fma(h, e, y0);
  }
  const double p = fma(e, 0.375, 0.5);
  const double q = e * p;
  return fma(y0, q, y0);
}

// ---------------------------------------------------------------------------------------------
// UNMASKED variants for the fast path of eval_kernel: no r2 == 0 test at all (r2 == 0 or denormal gives
// inf/NaN, which the kernel detects once per source tile and repairs by re-running that tile with
// rsqrt_masked). On the FP64 pipe every co-issued integer instruction costs ~1.2 issue cycles
// (profiles/r01b_fp64_micro.txt), so dropping ISETP + SEL + the zeroing of the seed's low word is worth ~7 %.
// The seed's LOW word is left undefined: MUFU.RSQ64H only defines the high word and the cubic step below
// tolerates a 2^-20 relative perturbation of the seed (remainder 5/16 e^3 with |e| < 2^-18: < 2e-17).
// ---------------------------------------------------------------------------------------------
// `lo_any` supplies the (irrelevant) low word of the seed: pass a value whose 64-bit register dies at this point
// (the partial sum dx^2 + dy^2), so that the seed can be built in that register pair without a move.
// NEWTON = 2: cubic step (full f64 precision). NEWTON = 1: one Newton step written so that the halving costs no
// integer instruction (e' = 1/2 - (r2 y0)(y0/2) as one DFMA on -y0/2; 4 FP64 instructions instead of 5, ~2^-38).
// NEWTON = 0: the seed itself (~2^-19 with the undefined low word), 0 extra instructions.
template <int NEWTON>
__device__ __forceinline__ double rsqrt_unmasked(double r2, double lo_any) {
  double seed;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(seed) : "d"(r2));
  const double y0 = __hiloint2double(__double2hiint(seed), __double2loint(lo_any));
  if (NEWTON == 0) return y0;
  const double t = r2 * y0;
  if (NEWTON == 1) {
    const double yh = -0.5 * y0;
    const double e = fma(t, yh, 0.5);
    return fma(y0, e, y0);
  }
  const double e = fma(-t, y0, 1.0);
  const double p = fma(e, 0.375, 0.5);
  const double q = e * p;
  return fma(y0, q, y0);
}
// Y = 2/sqrt(r2) after ONE Newton step with the halving deferred - the reference's own form of the iteration
// (intrin-wrapper.hpp:519-528): y0 (3 - r2 y0^2), 3 FP instructions; the factor 2 is folded into the kernel's scale factor
// (ukernels.cuh, X2 mode). Same quadratic convergence as the NEWTON = 1 forms above (~2^-38 in f64, ~1.5 ulp in f32).
__device__ __forceinline__ double rsqrt_unmasked_x2(double r2, double lo_any) {
  double seed;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(seed) : "d"(r2));
  const double y0 = __hiloint2double(__double2hiint(seed), __double2loint(lo_any));
  const double t = r2 * y0;
  const double e3 = fma(-t, y0, 3.0);
  return y0 * e3;
}
__device__ __forceinline__ float rsqrt_unmasked_x2(float r2, float) {
  float y0;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y0) : "f"(r2));
  const float t = r2 * y0;
  const float e3 = fmaf(-t, y0, 3.0f);
  return y0 * e3;
}
template <int NEWTON>
__device__ __forceinline__ float rsqrt_unmasked(float r2, float) {
  float y0;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y0) : "f"(r2));
  if (NEWTON == 0) return y0;
  const float t = r2 * y0;
  const float e = fmaf(-t, y0, 1.0f);
  const float h = 0.5f * y0;
  return fmaf(h, e, y0);
}

// FP32: MUFU.RSQ (rsqrt.approx.ftz.f32, ~2^-22.4) + one Newton step for full precision, as the
// reference's 1 iteration at digits = 6 (intrin-wrapper.hpp:3026).
template <int NEWTON>
__device__ __forceinline__ float rsqrt_pos_normal(float r2) {
  float y0;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y0) : "f"(r2));
  if (NEWTON == 0) return y0;
  const float t = r2 * y0;
  const float e = fmaf(-t, y0, 1.0f);
  const float h = 0.5f * y0;
  return fmaf(h, e, y0);
}
template <int NEWTON>
__device__ __forceinline__ float rsqrt_masked_sel(float r2) {
  const int rb = __float_as_int(r2);
  const bool ok = rb >= 0x00800000;
  const float xs = ok ? r2 : r2 * 0x1p48f;
  float seed;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(seed) : "f"(xs));
  float y0 = rb != 0 ? seed : 0.0f;
  y0 = ok ? y0 : y0 * 0x1p24f;
  if (NEWTON == 0) return y0;
  const float t = r2 * y0;
  const float e = fmaf(-t, y0, 1.0f);
  return fmaf(0.5f * y0, e, y0);
}
template <int NEWTON>
__device__ __forceinline__ float rsqrt_masked_flag(float r2, bool& subnormal) {
  float seed;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(seed) : "f"(r2));
  const int rb = __float_as_int(r2);
  const bool ok = rb >= 0x00800000;
  subnormal |= (!ok) & (rb != 0);
  const float y0 = ok ? seed : 0.0f;
  if (NEWTON == 0) return y0;
  const float t = r2 * y0;
  const float e = fmaf(-t, y0, 1.0f);
  const float h = 0.5f * y0;
  return fmaf(h, e, y0);
}
template <int NEWTON>
__device__ __forceinline__ float rsqrt_masked(float r2) {
  if (__float_as_int(r2) < 0x00800000) {  // zero or subnormal (|dX| < 1.1e-19): same rule as the f64 overload
    if (!(r2 > 0.0f)) return 0.0f;
    return rsqrt_pos_normal<NEWTON>(r2 * 0x1p48f) * 0x1p24f;
  }
  return rsqrt_pos_normal<NEWTON>(r2);
}

template <class Real> struct VecOf;
template <> struct VecOf<double> {
  using type = double2;
  static constexpr int N = 2;
};
template <> struct VecOf<float> {
  using type = float4;
  static constexpr int N = 4;
};

__host__ __device__ constexpr int round_up_int(int a, int b) { return (a + b - 1) / b * b; }

}  // namespace sctl_b200
