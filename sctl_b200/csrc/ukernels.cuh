// Device micro-kernels: the CUDA counterparts of kernel_impl::*::uKerMatrix
// (reference: include/sctl/kernel_functions.hpp:15-198) in two forms:
//
//   matrix(u, dx,dy,dz, rinv, n)  : the K0 x K1 block u[k0][k1] exactly as uKerMatrix defines it
//                                   (used by KernelMatrix, generic-kernel.txx:191-307);
//   apply(acc, dx,dy,dz, rinv, pay): acc[k1] += sum_k0 u[k0][k1] * f[k0] with the contraction over k0
//                                   (generic-kernel.txx:85-89) carried out algebraically, so that a pair
//                                   costs the minimum number of FP64/FP32 pipe instructions.
//
// `pay` is the per-source payload stored behind (x,y,z) in the packed source record; pack() builds
// it from the source normal n[ND] and density f[K0] (e.g. Laplace DxU stores f*n so the inner loop
// needs no separate density multiply). finalize() expands the accumulators to the K1 outputs
// (Stokes FxT keeps only the 6 distinct entries of its symmetric 3x3 result).
//
// FP-pipe instructions per pair, including dX (3), r^2 (3) and the full-precision rsqrt refinement (5 in
// f64, 4 in f32): Laplace FxU 12, DxU 17, FxdU 17; Stokes FxU 22, DxU 25, FxT 27, FSxU 22, FxUP 23.
//
// X2 mode (one-Newton-step refinement level only): apply() is fed Y = 2*rinv, the Newton step with its halving deferred
// exactly as the reference defers it (intrin-wrapper.hpp:519-528: y*(x*y*y - 3) scaled once at the end), which costs 3 FP
// instructions instead of 4. A kernel that is homogeneous of degree X2_SHIFT in rinv then accumulates 2^X2_SHIFT times its
// value and the exact power of two is folded into the scale factor. The Stokeslet-type kernels mix degrees 1 and 3: their
// additive density term is pre-multiplied by 4 once per SOURCE (prep_x2 -> aux, NAUX values shared by all targets of the
// thread), after which they accumulate 8x their value.
#pragma once
#include "common.cuh"

namespace sctl_b200 {

struct KLaplaceFxU {  // kernel_functions.hpp:15-31
  static constexpr int K0 = 1, K1 = 1, ND = 0, PAY = 1, NACC = 1, FLOPS = 6;
  static constexpr int X2_SHIFT = 1, NAUX = 0;
  static constexpr double SCALE_NUM = 1.0, SCALE_DEN_PI = 4.0;
  template <class Real> __device__ __forceinline__ static void pack(Real* pay, const Real*, const Real* f) { pay[0] = f[0]; }
  template <class Real> __device__ __forceinline__ static void prep_x2(Real*, const Real*) {}
  template <bool X2 = false, class Real> __device__ __forceinline__ static void apply(Real (&acc)[NACC], Real, Real, Real, Real rinv, const Real* pay, const Real* = nullptr) {
    acc[0] = fma(pay[0], rinv, acc[0]);
  }
  template <class Real> __device__ __forceinline__ static void finalize(Real (&out)[K1], const Real (&acc)[NACC]) { out[0] = acc[0]; }
  template <class Real> __device__ __forceinline__ static void matrix(Real (&u)[K0][K1], Real, Real, Real, Real rinv, const Real*) {
    u[0][0] = rinv;
  }
};

struct KLaplaceDxU {  // kernel_functions.hpp:33-51
  static constexpr int K0 = 1, K1 = 1, ND = 3, PAY = 3, NACC = 1, FLOPS = 14;
  static constexpr int X2_SHIFT = 3, NAUX = 0;
  template <class Real> __device__ __forceinline__ static void prep_x2(Real*, const Real*) {}
  static constexpr double SCALE_NUM = 1.0, SCALE_DEN_PI = 4.0;
  template <class Real> __device__ __forceinline__ static void pack(Real* pay, const Real* n, const Real* f) {
    pay[0] = f[0] * n[0];
    pay[1] = f[0] * n[1];
    pay[2] = f[0] * n[2];
  }
  template <bool X2 = false, class Real> __device__ __forceinline__ static void apply(Real (&acc)[NACC], Real dx, Real dy, Real dz, Real rinv, const Real* pay, const Real* = nullptr) {
    Real d = dx * pay[0];
    d = fma(dy, pay[1], d);
    d = fma(dz, pay[2], d);
    const Real rinv2 = rinv * rinv;
    const Real rinv3 = rinv2 * rinv;
    acc[0] = fma(d, rinv3, acc[0]);
  }
  template <class Real> __device__ __forceinline__ static void finalize(Real (&out)[K1], const Real (&acc)[NACC]) { out[0] = acc[0]; }
  template <class Real> __device__ __forceinline__ static void matrix(Real (&u)[K0][K1], Real dx, Real dy, Real dz, Real rinv, const Real* n) {
    Real d = dx * n[0];
    d = fma(dy, n[1], d);
    d = fma(dz, n[2], d);
    u[0][0] = d * (rinv * rinv * rinv);
  }
};

struct KLaplaceFxdU {  // kernel_functions.hpp:53-72
  static constexpr int K0 = 1, K1 = 3, ND = 0, PAY = 1, NACC = 3, FLOPS = 11;
  static constexpr int X2_SHIFT = 3, NAUX = 0;
  static constexpr double SCALE_NUM = -1.0, SCALE_DEN_PI = 4.0;
  template <class Real> __device__ __forceinline__ static void pack(Real* pay, const Real*, const Real* f) { pay[0] = f[0]; }
  template <class Real> __device__ __forceinline__ static void prep_x2(Real*, const Real*) {}
  template <bool X2 = false, class Real> __device__ __forceinline__ static void apply(Real (&acc)[NACC], Real dx, Real dy, Real dz, Real rinv, const Real* pay, const Real* = nullptr) {
    const Real rinv2 = rinv * rinv;
    const Real s = (pay[0] * rinv) * rinv2;
    acc[0] = fma(dx, s, acc[0]);
    acc[1] = fma(dy, s, acc[1]);
    acc[2] = fma(dz, s, acc[2]);
  }
  template <class Real> __device__ __forceinline__ static void finalize(Real (&out)[K1], const Real (&acc)[NACC]) {
    out[0] = acc[0];
    out[1] = acc[1];
    out[2] = acc[2];
  }
  template <class Real> __device__ __forceinline__ static void matrix(Real (&u)[K0][K1], Real dx, Real dy, Real dz, Real rinv, const Real*) {
    const Real rinv3 = rinv * rinv * rinv;
    u[0][0] = dx * rinv3;
    u[0][1] = dy * rinv3;
    u[0][2] = dz * rinv3;
  }
};

// shared by the three Stokeslet-type kernels: u[i][j] = delta_ij rinv + r_i r_j rinv^3
template <class Real>
__device__ __forceinline__ void stokeslet_block(Real (&u)[3][3], Real dx, Real dy, Real dz, Real rinv, Real rinv3) {
  const Real r[3] = {dx, dy, dz};
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) u[i][j] = (i == j ? rinv : (Real)0) + r[i] * r[j] * rinv3;
  }
}

struct KStokesFxU {  // kernel_functions.hpp:74-95
  static constexpr int K0 = 3, K1 = 3, ND = 0, PAY = 3, NACC = 3, FLOPS = 23;
  static constexpr int X2_SHIFT = 3, NAUX = 3;
  template <class Real> __device__ __forceinline__ static void prep_x2(Real* aux, const Real* pay) {
    aux[0] = (Real)4 * pay[0];
    aux[1] = (Real)4 * pay[1];
    aux[2] = (Real)4 * pay[2];
  }
  static constexpr double SCALE_NUM = 1.0, SCALE_DEN_PI = 8.0;
  template <class Real> __device__ __forceinline__ static void pack(Real* pay, const Real*, const Real* f) {
    pay[0] = f[0];
    pay[1] = f[1];
    pay[2] = f[2];
  }
  // u_j += rinv * (f_j + r_j * (r.f) * rinv^2);   X2: Y (4 f_j + r_j (r.f) Y^2) = 8x that
  template <bool X2 = false, class Real> __device__ __forceinline__ static void apply(Real (&acc)[NACC], Real dx, Real dy, Real dz, Real rinv, const Real* pay, const Real* aux = nullptr) {
    Real d = dx * pay[0];
    d = fma(dy, pay[1], d);
    d = fma(dz, pay[2], d);
    const Real rinv2 = rinv * rinv;
    const Real s = d * rinv2;
    const Real* add = X2 ? aux : pay;
    acc[0] = fma(rinv, fma(dx, s, add[0]), acc[0]);
    acc[1] = fma(rinv, fma(dy, s, add[1]), acc[1]);
    acc[2] = fma(rinv, fma(dz, s, add[2]), acc[2]);
  }
  template <class Real> __device__ __forceinline__ static void finalize(Real (&out)[K1], const Real (&acc)[NACC]) {
    out[0] = acc[0];
    out[1] = acc[1];
    out[2] = acc[2];
  }
  template <class Real> __device__ __forceinline__ static void matrix(Real (&u)[K0][K1], Real dx, Real dy, Real dz, Real rinv, const Real*) {
    stokeslet_block(u, dx, dy, dz, rinv, rinv * rinv * rinv);
  }
};

struct KStokesDxU {  // kernel_functions.hpp:97-120
  static constexpr int K0 = 3, K1 = 3, ND = 3, PAY = 6, NACC = 3, FLOPS = 26;
  static constexpr int X2_SHIFT = 5, NAUX = 0;
  template <class Real> __device__ __forceinline__ static void prep_x2(Real*, const Real*) {}
  static constexpr double SCALE_NUM = 3.0, SCALE_DEN_PI = 4.0;
  template <class Real> __device__ __forceinline__ static void pack(Real* pay, const Real* n, const Real* f) {
    pay[0] = n[0];
    pay[1] = n[1];
    pay[2] = n[2];
    pay[3] = f[0];
    pay[4] = f[1];
    pay[5] = f[2];
  }
  // u_j += r_j * (r.f)(r.n) rinv^5
  template <bool X2 = false, class Real> __device__ __forceinline__ static void apply(Real (&acc)[NACC], Real dx, Real dy, Real dz, Real rinv, const Real* pay, const Real* = nullptr) {
    Real dn = dx * pay[0];
    dn = fma(dy, pay[1], dn);
    dn = fma(dz, pay[2], dn);
    Real df = dx * pay[3];
    df = fma(dy, pay[4], df);
    df = fma(dz, pay[5], df);
    const Real rinv2 = rinv * rinv;
    const Real rinv5 = (rinv2 * rinv2) * rinv;
    const Real s = (dn * df) * rinv5;
    acc[0] = fma(dx, s, acc[0]);
    acc[1] = fma(dy, s, acc[1]);
    acc[2] = fma(dz, s, acc[2]);
  }
  template <class Real> __device__ __forceinline__ static void finalize(Real (&out)[K1], const Real (&acc)[NACC]) {
    out[0] = acc[0];
    out[1] = acc[1];
    out[2] = acc[2];
  }
  template <class Real> __device__ __forceinline__ static void matrix(Real (&u)[K0][K1], Real dx, Real dy, Real dz, Real rinv, const Real* n) {
    const Real rinv2 = rinv * rinv;
    const Real rinv5 = rinv2 * rinv2 * rinv;
    Real dn = dx * n[0];
    dn = fma(dy, n[1], dn);
    dn = fma(dz, n[2], dn);
    const Real w = dn * rinv5;
    const Real r[3] = {dx, dy, dz};
#pragma unroll
    for (int i = 0; i < 3; i++) {
#pragma unroll
      for (int j = 0; j < 3; j++) u[i][j] = r[i] * r[j] * w;
    }
  }
};

struct KStokesFxT {  // kernel_functions.hpp:122-146   u[i][j*3+k] = r_i r_j r_k rinv^5
  static constexpr int K0 = 3, K1 = 9, ND = 0, PAY = 3, NACC = 6, FLOPS = 39;
  static constexpr int X2_SHIFT = 5, NAUX = 0;
  template <class Real> __device__ __forceinline__ static void prep_x2(Real*, const Real*) {}
  static constexpr double SCALE_NUM = -3.0, SCALE_DEN_PI = 4.0;
  template <class Real> __device__ __forceinline__ static void pack(Real* pay, const Real*, const Real* f) {
    pay[0] = f[0];
    pay[1] = f[1];
    pay[2] = f[2];
  }
  // T_jk += r_j r_k (r.f) rinv^5 : symmetric in (j,k) -> 6 accumulators (00,01,02,11,12,22)
  template <bool X2 = false, class Real> __device__ __forceinline__ static void apply(Real (&acc)[NACC], Real dx, Real dy, Real dz, Real rinv, const Real* pay, const Real* = nullptr) {
    Real d = dx * pay[0];
    d = fma(dy, pay[1], d);
    d = fma(dz, pay[2], d);
    const Real rinv2 = rinv * rinv;
    const Real rinv5 = (rinv2 * rinv2) * rinv;
    const Real s = d * rinv5;
    const Real sx = dx * s, sy = dy * s, sz = dz * s;
    acc[0] = fma(dx, sx, acc[0]);
    acc[1] = fma(dx, sy, acc[1]);
    acc[2] = fma(dx, sz, acc[2]);
    acc[3] = fma(dy, sy, acc[3]);
    acc[4] = fma(dy, sz, acc[4]);
    acc[5] = fma(dz, sz, acc[5]);
  }
  template <class Real> __device__ __forceinline__ static void finalize(Real (&out)[K1], const Real (&acc)[NACC]) {
    out[0] = acc[0];
    out[1] = acc[1];
    out[2] = acc[2];
    out[3] = acc[1];
    out[4] = acc[3];
    out[5] = acc[4];
    out[6] = acc[2];
    out[7] = acc[4];
    out[8] = acc[5];
  }
  template <class Real> __device__ __forceinline__ static void matrix(Real (&u)[K0][K1], Real dx, Real dy, Real dz, Real rinv, const Real*) {
    const Real rinv2 = rinv * rinv;
    const Real rinv5 = rinv2 * rinv2 * rinv;
    const Real r[3] = {dx, dy, dz};
#pragma unroll
    for (int i = 0; i < 3; i++) {
#pragma unroll
      for (int j = 0; j < 3; j++) {
#pragma unroll
        for (int k = 0; k < 3; k++) u[i][j * 3 + k] = r[i] * r[j] * r[k] * rinv5;
      }
    }
  }
};

struct KStokesFSxU {  // kernel_functions.hpp:148-172   Stokeslet + source/sink row u[3][j] = r_j rinv^3
  static constexpr int K0 = 4, K1 = 3, ND = 0, PAY = 4, NACC = 3, FLOPS = 26;
  static constexpr int X2_SHIFT = 3, NAUX = 3;
  template <class Real> __device__ __forceinline__ static void prep_x2(Real* aux, const Real* pay) {
    aux[0] = (Real)4 * pay[0];
    aux[1] = (Real)4 * pay[1];
    aux[2] = (Real)4 * pay[2];
  }
  static constexpr double SCALE_NUM = 1.0, SCALE_DEN_PI = 8.0;
  template <class Real> __device__ __forceinline__ static void pack(Real* pay, const Real*, const Real* f) {
    pay[0] = f[0];
    pay[1] = f[1];
    pay[2] = f[2];
    pay[3] = f[3];
  }
  // u_j += rinv * (f_j + r_j * (r.f + f_3) * rinv^2)
  template <bool X2 = false, class Real> __device__ __forceinline__ static void apply(Real (&acc)[NACC], Real dx, Real dy, Real dz, Real rinv, const Real* pay, const Real* aux = nullptr) {
    Real d = fma(dx, pay[0], pay[3]);
    d = fma(dy, pay[1], d);
    d = fma(dz, pay[2], d);
    const Real rinv2 = rinv * rinv;
    const Real s = d * rinv2;
    const Real* add = X2 ? aux : pay;
    acc[0] = fma(rinv, fma(dx, s, add[0]), acc[0]);
    acc[1] = fma(rinv, fma(dy, s, add[1]), acc[1]);
    acc[2] = fma(rinv, fma(dz, s, add[2]), acc[2]);
  }
  template <class Real> __device__ __forceinline__ static void finalize(Real (&out)[K1], const Real (&acc)[NACC]) {
    out[0] = acc[0];
    out[1] = acc[1];
    out[2] = acc[2];
  }
  template <class Real> __device__ __forceinline__ static void matrix(Real (&u)[K0][K1], Real dx, Real dy, Real dz, Real rinv, const Real*) {
    const Real rinv3 = rinv * rinv * rinv;
    Real b[3][3];
    stokeslet_block(b, dx, dy, dz, rinv, rinv3);
#pragma unroll
    for (int i = 0; i < 3; i++) {
#pragma unroll
      for (int j = 0; j < 3; j++) u[i][j] = b[i][j];
    }
    u[3][0] = dx * rinv3;
    u[3][1] = dy * rinv3;
    u[3][2] = dz * rinv3;
  }
};

struct KStokesFxUP {  // kernel_functions.hpp:174-198   Stokeslet + pressure column u[i][3] = r_i rinv^3
  static constexpr int K0 = 3, K1 = 4, ND = 0, PAY = 3, NACC = 4, FLOPS = 26;
  static constexpr int X2_SHIFT = 3, NAUX = 3;
  template <class Real> __device__ __forceinline__ static void prep_x2(Real* aux, const Real* pay) {
    aux[0] = (Real)4 * pay[0];
    aux[1] = (Real)4 * pay[1];
    aux[2] = (Real)4 * pay[2];
  }
  static constexpr double SCALE_NUM = 1.0, SCALE_DEN_PI = 8.0;
  template <class Real> __device__ __forceinline__ static void pack(Real* pay, const Real*, const Real* f) {
    pay[0] = f[0];
    pay[1] = f[1];
    pay[2] = f[2];
  }
  template <bool X2 = false, class Real> __device__ __forceinline__ static void apply(Real (&acc)[NACC], Real dx, Real dy, Real dz, Real rinv, const Real* pay, const Real* aux = nullptr) {
    Real d = dx * pay[0];
    d = fma(dy, pay[1], d);
    d = fma(dz, pay[2], d);
    const Real rinv2 = rinv * rinv;
    const Real s = d * rinv2;
    const Real* add = X2 ? aux : pay;
    acc[0] = fma(rinv, fma(dx, s, add[0]), acc[0]);
    acc[1] = fma(rinv, fma(dy, s, add[1]), acc[1]);
    acc[2] = fma(rinv, fma(dz, s, add[2]), acc[2]);
    acc[3] = fma(s, rinv, acc[3]);  // p += (r.f) rinv^3
  }
  template <class Real> __device__ __forceinline__ static void finalize(Real (&out)[K1], const Real (&acc)[NACC]) {
    out[0] = acc[0];
    out[1] = acc[1];
    out[2] = acc[2];
    out[3] = acc[3];
  }
  template <class Real> __device__ __forceinline__ static void matrix(Real (&u)[K0][K1], Real dx, Real dy, Real dz, Real rinv, const Real*) {
    const Real rinv3 = rinv * rinv * rinv;
    Real b[3][3];
    stokeslet_block(b, dx, dy, dz, rinv, rinv3);
#pragma unroll
    for (int i = 0; i < 3; i++) {
#pragma unroll
      for (int j = 0; j < 3; j++) u[i][j] = b[i][j];
    }
    u[0][3] = dx * rinv3;
    u[1][3] = dy * rinv3;
    u[2][3] = dz * rinv3;
  }
};

// length (in Reals) of one packed source record: (x,y,z,payload) rounded up so that every record is
// a whole number of 16-byte shared-memory vector loads (double2 / float4).
template <class Ker, class Real>
__host__ __device__ constexpr int rec_len() {
  return round_up_int(3 + Ker::PAY, VecOf<Real>::N);
}

}  // namespace sctl_b200
