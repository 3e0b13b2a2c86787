// Direct summation kernels: pack -> eval (-> reduce).  Replaces the hot loop of
// GenericKernel<uKernel>::Eval (reference: include/sctl/generic-kernel.txx:153-186).
//
// Mapping (B200-first, FP64/FP32 CUDA-core bound; no tensor cores — this is not a dense contraction):
//   * the reference vectorises over TARGETS (SIMD lanes) and broadcasts each source
//     (generic-kernel.txx:155-166). Here a thread owns R targets in registers (coordinates +
//     accumulators), a CTA owns THREADS*R consecutive targets, and every source record is a
//     warp-uniform shared-memory broadcast (LDS.128, conflict-free) reused by all R targets.
//   * sources are first packed (pack_kernel) into records {x,y,z,payload} padded to the tile size with
//     zero-density dummies, so the inner loop has no bounds checks and a tile is ONE contiguous block.
//   * tiles are streamed global->shared by 1-D bulk TMA (cp.async.bulk, SASS UBLKCP) into a
//     STAGES-deep ring guarded by full/empty mbarriers; warp 0 doubles as the producer.
//   * grid = (target tiles) x (source chunks). Splitting the source range gives enough CTAs to kill
//     wave quantisation when a GPU only holds a few hundred target tiles (8-GPU sharding) and for
//     skinny problems (few targets, many sources). Chunk partial sums go to a [chunk][k][t] f64 buffer
//     and a second tiny kernel adds them in a FIXED order -> results are run-to-run deterministic.
//   * the scale factor is applied once after the sum and v_trg is accumulated into
//     (generic-kernel.txx:182-186) by whichever kernel finishes the sum.
//   * every tile is summed into per-tile accumulators that are flushed into f64 running sums after the tile
//     (FP32: 16M-source sums stay inside 1e-5, the reference accumulates in plain f32; FP64: blocked
//     summation, slightly more accurate than the reference's single in-order accumulator).
//   * there is NO r2 == 0 test in the hot loop (at any rsqrt refinement level): a coincident pair turns the
//     tile's partial sums non-finite, which is detected once per tile and repaired by re-running that tile
//     with the masked rsqrt (exactly-zero contribution, kernel_functions.hpp:28).
#pragma once
#include "ukernels.cuh"

namespace sctl_b200 {

// Tuning knobs (compile-time; the -D overrides exist for the variant sweeps of scripts/build_variants.sh)
#ifndef SCTL_B200_EVAL_THREADS
#define SCTL_B200_EVAL_THREADS 128
#endif
#ifndef SCTL_B200_EVAL_TILE
#define SCTL_B200_EVAL_TILE 256
#endif
#ifndef SCTL_B200_EVAL_STAGES
#define SCTL_B200_EVAL_STAGES 3
#endif
#ifndef SCTL_B200_UNROLL
#define SCTL_B200_UNROLL 1
#endif
#ifndef SCTL_B200_PREFETCH_F64
#define SCTL_B200_PREFETCH_F64 0  // 1: the same for EVERY FP64 eval_kernel (sweeps)
#endif
#ifndef SCTL_B200_PREFETCH_F32
#define SCTL_B200_PREFETCH_F32 0  // 1: EVERY FP32 eval_kernel reads the next source record one iteration ahead (sweeps; default: per kernel, PrefetchF32)
#endif
constexpr int EVAL_THREADS = SCTL_B200_EVAL_THREADS;  // 4 warps; 2 CTAs/SM at R = 6 (212 registers): ILP, not occupancy, hides the FP64 latency
constexpr int EVAL_TILE = SCTL_B200_EVAL_TILE;        // sources per shared-memory stage
constexpr int EVAL_STAGES = SCTL_B200_EVAL_STAGES;    // ring depth (tile k-1's slot is refilled while tile k is consumed)
constexpr int EVAL_UNROLL = SCTL_B200_UNROLL;         // source-loop unroll factor

// FP32 instantiations that read the next source record one iteration ahead (see PREFETCH in eval_kernel): bit n of `levels` =
// refinement level NEWTON = n. Only the (kernel, level) pairs that were measured to win (profiles/r02p_f32_prefetch.txt).
template <class Ker> struct PrefetchF32 { static constexpr int levels = 0; };
template <> struct PrefetchF32<KLaplaceFxU> { static constexpr int levels = 3; };   // digits 4: +9.0 %, full: +1.0 %
template <> struct PrefetchF32<KLaplaceFxdU> { static constexpr int levels = 3; };  // digits 4: +2.7 %, full: +8.0 %
template <> struct PrefetchF32<KStokesFSxU> { static constexpr int levels = 2; };   // full: +3.2 % (digits <= 5 not timed)
template <class Ker> struct PrefetchF64 { static constexpr int levels = 0; };  // FP64: none (see PREFETCH)

struct EvalParams {
  const void* r_trg;  // AoS [Nt*3]
  const void* rec;    // packed source records [n_tiles*EVAL_TILE*REC]
  void* v_trg;        // AoS [Nt*K1]                 (used when n_chunks == 1)
  double* partial;    // [n_chunks][K1][nt_stride]   (used when n_chunks  > 1)
  long long Nt;
  long long nt_stride;
  int n_tiles;          // total source tiles
  int tiles_per_chunk;  // source tiles per blockIdx.y
  int n_chunks;
  int accumulate;  // v_trg += (1) or v_trg = (0)
  double scale;
};

// ---------------------------------------------------------------------------------------------
// pack: AoS (r_src, n_src, v_src [* src_wts]) -> records {x,y,z,payload,0-pad}; tail padded with zero records
// ---------------------------------------------------------------------------------------------
template <class Ker, class Real>
__global__ void __launch_bounds__(256) pack_kernel(const Real* __restrict__ r_src, const Real* __restrict__ n_src,
                                                   const Real* __restrict__ v_src, const Real* __restrict__ src_wts,
                                                   Real* __restrict__ rec, long long Ns, long long Ns_pad) {
  constexpr int REC = rec_len<Ker, Real>();
  const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= Ns_pad) return;
  Real out[REC];
#pragma unroll
  for (int i = 0; i < REC; i++) out[i] = (Real)0;
  {
    // Tail padding repeats the LAST real source's position with a zero payload: the unmasked hot loop then never sees
    // geometry the caller did not supply (a dummy at the origin would overflow r2 for far-away f32 targets and poison
    // the tile with inf * 0), and a dummy pair contributes exactly 0 like the reference's absent pair.
    const long long sc = s < Ns ? s : Ns - 1;
    out[0] = r_src[sc * 3 + 0];
    out[1] = r_src[sc * 3 + 1];
    out[2] = r_src[sc * 3 + 2];
  }
  if (s < Ns) {
    Real n[3] = {0, 0, 0}, f[Ker::K0];
#pragma unroll
    for (int i = 0; i < Ker::ND; i++) n[i] = n_src[s * Ker::ND + i];
#pragma unroll
    for (int i = 0; i < Ker::K0; i++) f[i] = v_src[s * Ker::K0 + i];
    if (src_wts) {  // F <- F * wts: the far-field quadrature weights of BoundaryIntegralOp::ComputeFarField (boundary_integral.txx:1036-1051)
      const Real w = src_wts[s];
#pragma unroll
      for (int i = 0; i < Ker::K0; i++) f[i] *= w;
    }
    Ker::pack(out + 3, n, f);
  }
  using V = typename VecOf<Real>::type;
  V* dst = reinterpret_cast<V*>(rec + s * REC);
#pragma unroll
  for (int i = 0; i < REC / VecOf<Real>::N; i++) dst[i] = reinterpret_cast<const V*>(out)[i];
}

// Which instantiations run in X2 mode: the one-Newton-step refinement level - digits 5..10 in f64, and the full-precision
// level of f32 (one step from the MUFU.RSQ seed, the reference's own count at digits = 6, intrin-wrapper.hpp:3026).
template <class Real, int NEWTON>
__host__ __device__ constexpr bool eval_x2() {
  return NEWTON == 1;
}

// ---------------------------------------------------------------------------------------------
// Tile repair. The hot loops run unmasked; a pair with r2 == 0 (or subnormal r2) made rinv inf/NaN and poisoned this thread's
// partial sums of the tile. Detect it with one add chain per target and redo the tile with the masked rsqrt in its branch-free
// form (rsqrt_masked_sel): exactly-zero contribution for r2 == 0 (kernel_functions.hpp:28) and a finite rinv for every
// positive r2, subnormal ones included. Every tile on the diagonal of a self-interaction (targets == sources) takes this path,
// so its speed matters at mid sizes (185 SASS instructions per source against 147 in the hot loop; the branchy general form
// was 277).
// ---------------------------------------------------------------------------------------------
#ifndef SCTL_B200_REPAIR_MODE
#define SCTL_B200_REPAIR_MODE 7
#endif
// How the repair is coded matters to the HOT loop, although it never runs there: the repair code shares the kernel's register
// allocation, and the hot loop's FP64-pipe rate depends on the exact register (bank) assignment of its DFMA operands. Measured
// on B200, Stokes FxU f64 at N = 262144 (profiles/r02q_repair_mode_sweep.txt; round-1 library: 90.2 ms):
//   7  ONE branch-free loop with rsqrt_masked_sel (zero -> 0, subnormal rescaled by selects)              90.1 ms  <- default
//   3  one loop with the select form that sends subnormal r2 to 0 (round 1's semantics, not the mask's)    90.1 ms
//   0  select loop with a "positive subnormal seen" flag + a second, general loop, both inline            91.5 ms
//   (second loop one target at a time, inline or through a small __noinline__ function: 91.5 / 91.7 ms; second loop or both
//    loops out of line with their arguments in a local-memory struct: 92.1 / 92.4 ms - variants removed again)
// Same 147 SASS instructions in the hot loop in every case; only the register numbers differ.
template <class Ker, class Real, int R, int NEWTON, int FORM, class V>
__device__ __forceinline__ bool repair_loop(Real (&tacc)[R][Ker::NACC], const Real (&xt)[R], const Real (&yt)[R], const Real (&zt)[R], const V* tile, int ns) {
  constexpr int REC = rec_len<Ker, Real>();
  constexpr int NACC = Ker::NACC;
  constexpr int NV = REC / VecOf<Real>::N;
  bool subnormal = false;
#pragma unroll
  for (int r = 0; r < R; r++) {
#pragma unroll
    for (int q = 0; q < NACC; q++) tacc[r][q] = (Real)0;
  }
#pragma unroll 1
  for (int s = 0; s < ns; s++) {
    Real src[REC];
#pragma unroll
    for (int i = 0; i < NV; i++) reinterpret_cast<V*>(src)[i] = tile[s * NV + i];
#pragma unroll
    for (int r = 0; r < R; r++) {
      const Real dx = xt[r] - src[0];
      const Real dy = yt[r] - src[1];
      const Real dz = zt[r] - src[2];
      Real r2 = dx * dx;
      r2 = fma(dy, dy, r2);
      r2 = fma(dz, dz, r2);
      // FORM 0: branch-free and complete; 1: branch-free, subnormal -> 0 + flag; 2: general form with branches
      const Real rinv = FORM == 0 ? rsqrt_masked_sel<NEWTON>(r2) : (FORM == 1 ? rsqrt_masked_flag<NEWTON>(r2, subnormal) : rsqrt_masked<NEWTON>(r2));
      Ker::apply(tacc[r], dx, dy, dz, rinv, src + 3);
    }
  }
  return subnormal;
}

template <class Ker, class Real, int R, int NEWTON, class V>
__device__ __forceinline__ bool repair_tile(Real (&tacc)[R][Ker::NACC], const Real (&xt)[R], const Real (&yt)[R], const Real (&zt)[R], const V* tile, int ns) {
  constexpr int NACC = Ker::NACC;
  bool bad = false;
#pragma unroll
  for (int r = 0; r < R; r++) {
    Real sum = tacc[r][0];
#pragma unroll
    for (int q = 1; q < NACC; q++) sum += tacc[r][q];
    bad |= !(fabs(sum) <= (Real)(sizeof(Real) == 8 ? 1.7976931348623157e308 : 3.402823466e38));
  }
  if (!bad) return false;
  constexpr int MODE = SCTL_B200_REPAIR_MODE;
  if (MODE == 7) {
    repair_loop<Ker, Real, R, NEWTON, 0>(tacc, xt, yt, zt, tile, ns);
  } else {
    const bool subnormal = repair_loop<Ker, Real, R, NEWTON, 1>(tacc, xt, yt, zt, tile, ns);
    if (MODE == 0 && subnormal) repair_loop<Ker, Real, R, NEWTON, 2>(tacc, xt, yt, zt, tile, ns);
  }
  return true;
}

// ---------------------------------------------------------------------------------------------
// eval: the O(Ns*Nt) kernel
// ---------------------------------------------------------------------------------------------
template <class Ker, class Real, int R, int NEWTON, int MINB>
__global__ void __launch_bounds__(EVAL_THREADS, MINB) eval_kernel(const EvalParams p) {
  constexpr int REC = rec_len<Ker, Real>();
  constexpr int NACC = Ker::NACC;
  constexpr int K1 = Ker::K1;
  constexpr uint32_t TILE_BYTES = EVAL_TILE * REC * sizeof(Real);
  using V = typename VecOf<Real>::type;
  constexpr int NV = REC / VecOf<Real>::N;
  // one-Newton-step level: Y = 2 rinv with the halving folded into the scale factor (ukernels.cuh "X2 mode"); the caller
  // passes p.scale already divided by 2^X2_SHIFT (eval_x2_shift)
  constexpr bool X2 = eval_x2<Real, NEWTON>();
  // FP32 only: the source record is read from shared memory one iteration ahead. The f32 loop is short enough per source that
  // the LDS latency shows at its top (profiles/r02y_stokes_fxu_f32.hotloop.txt), but what the reordering does to the rest of
  // the schedule depends on the kernel (profiles/r02p_f32_prefetch.txt, all f32 kernels built with -DSCTL_B200_PREFETCH_F32=1:
  // Laplace FxdU +8.0 %, Laplace FxU +1.0 % (+9.0 % at digits 4), Stokes FSxU +3.2 %, Laplace DxU -0.5 %, Stokes FxU -1.5 %,
  // FxUP -3.0 %, DxU -5.8 %, Stokes FxU at digits 4 -7.3 %), so it is on only where it was measured to win (PrefetchF32 above).
  // The f64 loop is bound by the FP64 pipe and is left exactly as it is.
  constexpr bool PREFETCH = R > 1 && (sizeof(Real) == 4 ? (SCTL_B200_PREFETCH_F32 || ((PrefetchF32<Ker>::levels >> NEWTON) & 1))
                                                        : (SCTL_B200_PREFETCH_F64 || ((PrefetchF64<Ker>::levels >> NEWTON) & 1)));

  extern __shared__ __align__(128) unsigned char smem_raw[];
  Real* tiles = reinterpret_cast<Real*>(smem_raw);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_raw + (size_t)EVAL_STAGES * TILE_BYTES);
  uint64_t* empty_bar = full_bar + EVAL_STAGES;

  const int tid = threadIdx.x;
  const int tile0 = blockIdx.y * p.tiles_per_chunk;
  const int ntile = min(p.tiles_per_chunk, p.n_tiles - tile0);
  const Real* rec = reinterpret_cast<const Real*>(p.rec) + (size_t)tile0 * EVAL_TILE * REC;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < EVAL_STAGES; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], EVAL_THREADS / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < EVAL_STAGES; s++) {
      if (s < ntile) {
        mbar_arrive_expect_tx(&full_bar[s], TILE_BYTES);
        bulk_g2s(tiles + (size_t)s * EVAL_TILE * REC, rec + (size_t)s * EVAL_TILE * REC, TILE_BYTES, &full_bar[s]);
      }
    }
  }

  // R register-blocked targets per thread, strided by the CTA width so global accesses coalesce.
  const long long t_base = (long long)blockIdx.x * (EVAL_THREADS * R) + tid;
  const Real* r_trg = reinterpret_cast<const Real*>(p.r_trg);
  Real xt[R], yt[R], zt[R];
  Real tacc[R][NACC];    // partial sums of the current source tile
  double acc[R][NACC];   // running sums over tiles, always f64 (see FP32 note in the header comment)
#pragma unroll
  for (int r = 0; r < R; r++) {
    long long t = t_base + (long long)r * EVAL_THREADS;
    t = t < p.Nt ? t : p.Nt - 1;  // clamp: out-of-range lanes recompute the last target and are not stored
    xt[r] = r_trg[t * 3 + 0];
    yt[r] = r_trg[t * 3 + 1];
    zt[r] = r_trg[t * 3 + 2];
#pragma unroll
    for (int k = 0; k < NACC; k++) acc[r][k] = 0.0;
  }

  // Every tile runs UNMASKED (no r2 == 0 test in the hot loop); the tile that held a coincident pair is repaired (repair_tile).

  for (int k = 0; k < ntile; k++) {
    const int stage = k % EVAL_STAGES;
    const uint32_t parity = (uint32_t)(k / EVAL_STAGES) & 1u;
    mbar_wait(&full_bar[stage], parity);
    const V* tile = reinterpret_cast<const V*>(tiles + (size_t)stage * EVAL_TILE * REC);

#pragma unroll
    for (int r = 0; r < R; r++) {
#pragma unroll
      for (int q = 0; q < NACC; q++) tacc[r][q] = (Real)0;
    }

    V nxt[PREFETCH ? NV : 1];  // PREFETCH: the record of source s + 1, loaded while source s is being evaluated
    if constexpr (PREFETCH) {
#pragma unroll
      for (int i = 0; i < NV; i++) nxt[i] = tile[i];
    }
#pragma unroll EVAL_UNROLL
    for (int s = 0; s < EVAL_TILE; s++) {
      Real src[REC];
      if constexpr (PREFETCH) {
        const int sn = s + 1 < EVAL_TILE ? s + 1 : s;
#pragma unroll
        for (int i = 0; i < NV; i++) {
          reinterpret_cast<V*>(src)[i] = nxt[i];
          nxt[i] = tile[sn * NV + i];
        }
      } else {
#pragma unroll
        for (int i = 0; i < NV; i++) reinterpret_cast<V*>(src)[i] = tile[s * NV + i];  // warp-uniform LDS.128 broadcast
      }
      Real aux[Ker::NAUX > 0 ? Ker::NAUX : 1];
      if (X2) Ker::prep_x2(aux, src + 3);  // once per source, shared by the R targets
#pragma unroll
      for (int r = 0; r < R; r++) {
        const Real dx = xt[r] - src[0];
        const Real dy = yt[r] - src[1];
        const Real dz = zt[r] - src[2];
        Real r2p = dx * dx;
        r2p = fma(dy, dy, r2p);
        const Real r2 = fma(dz, dz, r2p);
        if (X2) {
          Ker::template apply<true>(tacc[r], dx, dy, dz, rsqrt_unmasked_x2(r2, r2p), src + 3, aux);
        } else {
          Ker::apply(tacc[r], dx, dy, dz, rsqrt_unmasked<NEWTON>(r2, r2p), src + 3);
        }
      }
    }

    if (repair_tile<Ker, Real, R, NEWTON>(tacc, xt, yt, zt, tile, EVAL_TILE) && X2) {
      // the repaired sums are the plain ones: bring them to the 2^X2_SHIFT convention of this instantiation (exact)
#pragma unroll
      for (int r = 0; r < R; r++) {
#pragma unroll
        for (int q = 0; q < NACC; q++) tacc[r][q] *= (Real)(1 << Ker::X2_SHIFT);
      }
    }

#pragma unroll
    for (int r = 0; r < R; r++) {
#pragma unroll
      for (int q = 0; q < NACC; q++) acc[r][q] += (double)tacc[r][q];
    }

    // Release the stage. Warp 0 then refills the slot that tile k-1 used (released by every warp about one
    // tile ago, so the wait below practically never blocks) with tile k-1+STAGES. All 32 lanes of warp 0
    // poll the barrier together — a lane-0-only wait would leave the warp diverged for the whole next tile —
    // and a single lane issues the copy.
    __syncwarp();
    if ((tid & 31) == 0) mbar_arrive(&empty_bar[stage]);
    if (tid < 32 && k >= 1 && k - 1 + EVAL_STAGES < ntile) {
      const int j = k - 1;
      const int js = j % EVAL_STAGES;
      mbar_wait(&empty_bar[js], (uint32_t)(j / EVAL_STAGES) & 1u);
      if (tid == 0) {
        mbar_arrive_expect_tx(&full_bar[js], TILE_BYTES);
        bulk_g2s(tiles + (size_t)js * EVAL_TILE * REC, rec + (size_t)(j + EVAL_STAGES) * EVAL_TILE * REC, TILE_BYTES, &full_bar[js]);
      }
      __syncwarp();
    }
  }

  // epilogue
#pragma unroll
  for (int r = 0; r < R; r++) {
    const long long t = t_base + (long long)r * EVAL_THREADS;
    if (t >= p.Nt) continue;
    double a[NACC], out[K1];
#pragma unroll
    for (int q = 0; q < NACC; q++) a[q] = acc[r][q];
    Ker::finalize(out, a);
    if (p.n_chunks == 1) {
      Real* v = reinterpret_cast<Real*>(p.v_trg) + t * K1;
#pragma unroll
      for (int q = 0; q < K1; q++) {
        const double prev = p.accumulate ? (double)v[q] : 0.0;
        v[q] = (Real)(prev + out[q] * p.scale);
      }
    } else {
      double* dst = p.partial + (size_t)blockIdx.y * K1 * p.nt_stride + t;
#pragma unroll
      for (int q = 0; q < K1; q++) dst[(size_t)q * p.nt_stride] = out[q];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// reduce: v_trg[t*K1+k] (+)= scale * sum_c partial[c][k][t]   (fixed chunk order -> deterministic)
// ---------------------------------------------------------------------------------------------
template <class Real, int K1>
__global__ void __launch_bounds__(128) reduce_kernel(const double* __restrict__ partial, Real* __restrict__ v_trg, long long Nt,
                                                     long long nt_stride, int n_chunks, double scale, int accumulate) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= Nt) return;
  double sum[K1];
#pragma unroll
  for (int q = 0; q < K1; q++) sum[q] = 0.0;
  // the additions run in chunk order (deterministic); the loads of 8 chunks are issued together so that a thread does not
  // pay one memory round trip per chunk (27 chunks at 20k x 20k were 16 us of a 330 us evaluation)
  int c = 0;
  for (; c + 8 <= n_chunks; c += 8) {
    double v[8][K1];
#pragma unroll
    for (int u = 0; u < 8; u++) {
#pragma unroll
      for (int q = 0; q < K1; q++) v[u][q] = __ldcs(partial + ((size_t)(c + u) * K1 + q) * nt_stride + t);
    }
#pragma unroll
    for (int u = 0; u < 8; u++) {
#pragma unroll
      for (int q = 0; q < K1; q++) sum[q] += v[u][q];
    }
  }
  for (; c < n_chunks; c++) {
#pragma unroll
    for (int q = 0; q < K1; q++) sum[q] += __ldcs(partial + ((size_t)c * K1 + q) * nt_stride + t);
  }
#pragma unroll
  for (int q = 0; q < K1; q++) {
    const double prev = accumulate ? (double)v_trg[t * K1 + q] : 0.0;
    v_trg[t * K1 + q] = (Real)(prev + sum[q] * scale);
  }
}

// ---------------------------------------------------------------------------------------------
// U[t*K1out+k] = sum_l Ufull[(t*K1out+k)*3+l] * n_trg[t*3+l]: the target-normal dot product that
// BoundaryIntegralOp::ComputeFarField applies after fmm.Eval (boundary_integral.txx:1060-1071). O(Nt).
// ---------------------------------------------------------------------------------------------
template <class Real>
__global__ void __launch_bounds__(256) contract_normal_kernel(const Real* __restrict__ Ufull, const Real* __restrict__ n_trg, Real* __restrict__ U,
                                                              long long Nt, int K1out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // index into U: t*K1out + k
  if (i >= Nt * K1out) return;
  const long long t = i / K1out;
  Real a = Ufull[i * 3 + 0] * n_trg[t * 3 + 0];
  a = fma(Ufull[i * 3 + 1], n_trg[t * 3 + 1], a);
  a = fma(Ufull[i * 3 + 2], n_trg[t * 3 + 2], a);
  U[i] = a;
}

template <class Ker, class Real>
constexpr size_t eval_smem_bytes() {
  return (size_t)EVAL_STAGES * EVAL_TILE * rec_len<Ker, Real>() * sizeof(Real) + 2 * EVAL_STAGES * sizeof(uint64_t);
}

}  // namespace sctl_b200
