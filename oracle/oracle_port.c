/* TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT PATH.
 *
 * Plain-C CPU restatement ("port") of the reference's direct particle-kernel summation:
 *   GenericKernel<uKernel>::Eval          /root/reference/include/sctl/generic-kernel.txx:76-189
 *   GenericKernel<uKernel>::KernelMatrix  /root/reference/include/sctl/generic-kernel.txx:191-307
 *   the eight micro-kernels               /root/reference/include/sctl/kernel_functions.hpp:15-198
 *   ParticleFMM::EvalDirect (np = 1)      /root/reference/include/sctl/fmm-wrapper.txx:490-562
 *
 * PARITY PINNING: the reference ships no golden vector or known-answer test for this path
 * (SURVEY.md section 4 / 8c), so this port is pinned against OUTPUTS OF THE REFERENCE ITSELF:
 * tests/golden/ *.npz were produced by scripts/make_golden.py calling oracle/_ref/libsctl_ref_*.so
 * (the unmodified reference headers compiled by oracle/Makefile), and
 * tests/test_oracle_golden.py checks this file against them on every CPU test run.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library. The product (sctl_b200/csrc) never does.
 *
 * Arithmetic notes (what is restated, line by line):
 *  - dX = x_trg - x_src                                  generic-kernel.txx:83
 *  - r2 = r0*r0 + r1*r1 + r2*r2 (no FMA contraction here) kernel_functions.hpp:27,45,65,86,109
 *  - rinv = (r2 > 0) ? 1/sqrt(r2) : 0                     kernel_functions.hpp:28 + the scalar
 *    fallback of rsqrt_approx_intrin (intrin-wrapper.hpp:539-555); the SIMD builds use a 14-bit
 *    seed + Newton steps (intrin-wrapper.hpp:516-535,3025-3050) which agree to rounding.
 *  - vt[k1] = fma(U[k0][k1], vs[k0], vt[k1]) for k0, k1    generic-kernel.txx:85-89
 *  - single in-order accumulator per target over s = 0..Ns-1   generic-kernel.txx:159-164
 *  - v_trg[t*K1+k] += vt[k] * scale, scale applied once    generic-kernel.txx:182-186
 *  - M[s*K0+k0][t*K1+k1] = U[k0][k1] * scale (overwritten) generic-kernel.txx:246,284
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>

#define MAXK0 4
#define MAXK1 9

typedef struct {
  const char* name;
  int normal_dim, src_dim, trg_dim, flops;
  double scale_num, scale_den_pi; /* scale = scale_num / (scale_den_pi * pi) */
} kinfo_t;

/* kernel_functions.hpp:16-25,34-43,54-63,75-84,98-107,122-131,148-157,174-183 */
static const kinfo_t KINFO[8] = {
    {"Laplace3D-FxU", 0, 1, 1, 6, 1.0, 4.0},   {"Laplace3D-DxU", 3, 1, 1, 14, 1.0, 4.0},
    {"Laplace3D-FxdU", 0, 1, 3, 11, -1.0, 4.0}, {"Stokes3D-FxU", 0, 3, 3, 23, 1.0, 8.0},
    {"Stokes3D-DxU", 3, 3, 3, 26, 3.0, 4.0},    {"Stokes3D-FxT", 0, 3, 9, 39, -3.0, 4.0},
    {"Stokes3D-FSxU", 0, 4, 3, 26, 1.0, 8.0},   {"Stokes3D-FxUP", 0, 3, 4, 26, 1.0, 8.0},
};

#define PI_LD 3.14159265358979323846264338327950288L

#define DEFINE_UKER(REAL, SUFFIX, SQRT)                                                          \
  static void uker_##SUFFIX(int kid, REAL u[MAXK0][MAXK1], const REAL r[3], const REAL* n) {    \
    const REAL r2 = r[0] * r[0] + r[1] * r[1] + r[2] * r[2];                                     \
    const REAL rinv = (r2 > (REAL)0) ? (REAL)1 / SQRT(r2) : (REAL)0;                              \
    switch (kid) {                                                                               \
      case 0: /* Laplace3D_FxU  kernel_functions.hpp:26-30 */                                    \
        u[0][0] = rinv;                                                                          \
        break;                                                                                   \
      case 1: { /* Laplace3D_DxU  :44-50 */                                                      \
        const REAL rdotn = r[0] * n[0] + r[1] * n[1] + r[2] * n[2];                               \
        const REAL rinv3 = rinv * rinv * rinv;                                                   \
        u[0][0] = rdotn * rinv3;                                                                 \
      } break;                                                                                   \
      case 2: { /* Laplace3D_FxdU :64-71 */                                                      \
        const REAL rinv3 = rinv * rinv * rinv;                                                   \
        u[0][0] = r[0] * rinv3;                                                                  \
        u[0][1] = r[1] * rinv3;                                                                  \
        u[0][2] = r[2] * rinv3;                                                                  \
      } break;                                                                                   \
      case 3: { /* Stokes3D_FxU   :85-94 */                                                      \
        const REAL rinv3 = rinv * rinv * rinv;                                                   \
        for (int i = 0; i < 3; i++)                                                              \
          for (int j = 0; j < 3; j++) u[i][j] = (i == j ? rinv : (REAL)0) + r[i] * r[j] * rinv3;  \
      } break;                                                                                   \
      case 4: { /* Stokes3D_DxU   :108-119 */                                                    \
        const REAL rinv2 = rinv * rinv;                                                          \
        const REAL rinv5 = rinv2 * rinv2 * rinv;                                                 \
        const REAL rdotn_rinv5 = (r[0] * n[0] + r[1] * n[1] + r[2] * n[2]) * rinv5;               \
        for (int i = 0; i < 3; i++)                                                              \
          for (int j = 0; j < 3; j++) u[i][j] = r[i] * r[j] * rdotn_rinv5;                        \
      } break;                                                                                   \
      case 5: { /* Stokes3D_FxT   :132-144 */                                                    \
        const REAL rinv2 = rinv * rinv;                                                          \
        const REAL rinv5 = rinv2 * rinv2 * rinv;                                                 \
        for (int i = 0; i < 3; i++)                                                              \
          for (int j = 0; j < 3; j++)                                                            \
            for (int k = 0; k < 3; k++) u[i][j * 3 + k] = r[i] * r[j] * r[k] * rinv5;             \
      } break;                                                                                   \
      case 6: { /* Stokes3D_FSxU  :158-170 */                                                    \
        const REAL rinv3 = rinv * rinv * rinv;                                                   \
        for (int i = 0; i < 3; i++)                                                              \
          for (int j = 0; j < 3; j++) u[i][j] = (i == j ? rinv : (REAL)0) + r[i] * r[j] * rinv3;  \
        for (int j = 0; j < 3; j++) u[3][j] = r[j] * rinv3;                                      \
      } break;                                                                                   \
      case 7: { /* Stokes3D_FxUP  :184-196 */                                                    \
        const REAL rinv3 = rinv * rinv * rinv;                                                   \
        for (int i = 0; i < 3; i++)                                                              \
          for (int j = 0; j < 3; j++) u[i][j] = (i == j ? rinv : (REAL)0) + r[i] * r[j] * rinv3;  \
        for (int i = 0; i < 3; i++) u[i][3] = r[i] * rinv3;                                      \
      } break;                                                                                   \
    }                                                                                            \
  }

DEFINE_UKER(double, f64, sqrt)
DEFINE_UKER(float, f32, sqrtf)

#define DEFINE_EVAL(REAL, SUFFIX, FMA)                                                              \
  static void eval_##SUFFIX(int kid, int64_t Nt, int64_t Ns, const REAL* Xt, const REAL* Xs,       \
                            const REAL* Xn, const REAL* F, REAL* U) {                               \
    const kinfo_t* ki = &KINFO[kid];                                                                \
    const int K0 = ki->src_dim, K1 = ki->trg_dim, Nd = ki->normal_dim;                              \
    const REAL scale = (REAL)((long double)ki->scale_num / ((long double)ki->scale_den_pi * PI_LD)); \
    _Pragma("omp parallel for schedule(static)") for (int64_t t = 0; t < Nt; t++) {                 \
      REAL vt[MAXK1];                                                                               \
      for (int k = 0; k < K1; k++) vt[k] = 0;                                                       \
      for (int64_t s = 0; s < Ns; s++) {                                                            \
        REAL dX[3], u[MAXK0][MAXK1];                                                                \
        for (int k = 0; k < 3; k++) dX[k] = Xt[t * 3 + k] - Xs[s * 3 + k];                          \
        uker_##SUFFIX(kid, u, dX, Nd ? Xn + s * Nd : (const REAL*)0);                               \
        for (int k0 = 0; k0 < K0; k0++)                                                             \
          for (int k1 = 0; k1 < K1; k1++) vt[k1] = FMA(u[k0][k1], F[s * K0 + k0], vt[k1]);          \
      }                                                                                             \
      for (int k = 0; k < K1; k++) U[t * K1 + k] += vt[k] * scale;                                  \
    }                                                                                               \
  }                                                                                                 \
  static void kmat_##SUFFIX(int kid, int64_t Nt, int64_t Ns, const REAL* Xt, const REAL* Xs,       \
                            const REAL* Xn, REAL* M) {                                              \
    const kinfo_t* ki = &KINFO[kid];                                                                \
    const int K0 = ki->src_dim, K1 = ki->trg_dim, Nd = ki->normal_dim;                              \
    const REAL scale = (REAL)((long double)ki->scale_num / ((long double)ki->scale_den_pi * PI_LD)); \
    _Pragma("omp parallel for schedule(static)") for (int64_t s = 0; s < Ns; s++) {                 \
      for (int64_t t = 0; t < Nt; t++) {                                                            \
        REAL dX[3], u[MAXK0][MAXK1];                                                                \
        for (int k = 0; k < 3; k++) dX[k] = Xt[t * 3 + k] - Xs[s * 3 + k];                          \
        uker_##SUFFIX(kid, u, dX, Nd ? Xn + s * Nd : (const REAL*)0);                               \
        for (int k0 = 0; k0 < K0; k0++)                                                             \
          for (int k1 = 0; k1 < K1; k1++)                                                           \
            M[(size_t)(s * K0 + k0) * (size_t)(Nt * K1) + (size_t)(t * K1 + k1)] = u[k0][k1] * scale; \
      }                                                                                             \
    }                                                                                               \
  }

DEFINE_EVAL(double, f64, fma)
DEFINE_EVAL(float, f32, fmaf)

/* dtype: 0 = float, 1 = double. Accumulates into v_trg like the reference (generic-kernel.txx:98-101). */
int oracle_eval(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                const void* v_src, void* v_trg) {
  if (kernel_id < 0 || kernel_id > 7) return 1;
  if (dtype == 1) eval_f64(kernel_id, Nt, Ns, (const double*)r_trg, (const double*)r_src, (const double*)n_src, (const double*)v_src, (double*)v_trg);
  else if (dtype == 0) eval_f32(kernel_id, Nt, Ns, (const float*)r_trg, (const float*)r_src, (const float*)n_src, (const float*)v_src, (float*)v_trg);
  else return 2;
  return 0;
}

int oracle_kernel_matrix(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src, void* M) {
  if (kernel_id < 0 || kernel_id > 7) return 1;
  if (dtype == 1) kmat_f64(kernel_id, Nt, Ns, (const double*)r_trg, (const double*)r_src, (const double*)n_src, (double*)M);
  else if (dtype == 0) kmat_f32(kernel_id, Nt, Ns, (const float*)r_trg, (const float*)r_src, (const float*)n_src, (float*)M);
  else return 2;
  return 0;
}

/* ParticleFMM::EvalDirect for one (source type, target type) pair on a single rank:
 * U is overwritten (fmm-wrapper.txx:501-502), then every registered source type is accumulated (:514-559). */
int oracle_eval_direct(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                       const void* v_src, void* v_trg) {
  if (kernel_id < 0 || kernel_id > 7) return 1;
  const size_t esz = dtype == 1 ? sizeof(double) : sizeof(float);
  memset(v_trg, 0, esz * (size_t)Nt * (size_t)KINFO[kernel_id].trg_dim);
  return oracle_eval(kernel_id, dtype, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg);
}

int oracle_kernel_dims(int kernel_id, int* coord, int* normal, int* src, int* trg, double* scale, int* flops, char* name, int name_len) {
  if (kernel_id < 0 || kernel_id > 7) return 1;
  const kinfo_t* ki = &KINFO[kernel_id];
  if (coord) *coord = 3;
  if (normal) *normal = ki->normal_dim;
  if (src) *src = ki->src_dim;
  if (trg) *trg = ki->trg_dim;
  if (scale) *scale = (double)((long double)ki->scale_num / ((long double)ki->scale_den_pi * PI_LD));
  if (flops) *flops = ki->flops;
  if (name && name_len > 0) {
    strncpy(name, ki->name, (size_t)name_len - 1);
    name[name_len - 1] = 0;
  }
  return 0;
}
