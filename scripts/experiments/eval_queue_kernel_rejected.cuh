// eval_queue_kernel: the O(Ns*Nt) sum for MID-SIZE problems (a few thousand to ~100k points a side) - the regime every
// in-tree caller of the reference's Eval uses (FMM leaves, per-element far fields, BoundaryIntegralOp at 1e4-1e5 nodes).
//
// eval_kernel (eval_impl.cuh) launches (target tiles) x (source chunks) EQUAL CTAs; when the grid holds only 2-30 waves of them
// the SMs idle behind the last wave (20k x 20k: 729 CTAs on 296 slots = 2.46 waves -> 80-85 % of the large-N rate) and every
// CTA pays its own pipeline start-up. Here the grid is PERSISTENT - one CTA per resident slot - and the CTAs pull work items
// from a global ticket counter:
//   * an item = (target tile, contiguous range of source UNITS of 64 sources). Every target tile's source range is cut into
//     LEVELS: the first half into items of s0 units, the next quarter into items of s0/2, ... down to single units, with s0
//     chosen so that every level holds about two items per CTA. Tickets run level by level, so the kernel ends on 64-source
//     items (the guided self-scheduling shape: tail imbalance of one tiny item), and a CTA that was slowed down - e.g. by tiles
//     it had to repair for coincident points - simply takes fewer items. (Measured first and rejected: a static even split,
//     1.8x SLOWER at 20k x 20k with the benchmark's 1024 coincident targets, which all land in one CTA's share; and a
//     two-size schedule, whose last items were still 6 % of a CTA's work - profiles/r02f_midsize_schedules_rejected.txt.)
//   * ONE TMA ring runs across items: the producer lane fetches the next ticket while the current item is being summed and has
//     the next item's first tile in flight before the current one ends, so an item costs no pipeline start-up.
//   * an item writes its partial sum to slot (item index within the target tile) of the [slot][k1][t] buffer that reduce_kernel
//     adds in fixed order - run-to-run deterministic, whichever CTA computed which item.
// Hot loop, rsqrt and r2 == 0 repair are those of eval_kernel; this is a SEPARATE instantiation so that eval_kernel keeps its
// code (and its register allocation) for large N.
#pragma once
#include "eval_impl.cuh"

namespace sctl_b200 {

constexpr int SPLIT_UNIT = 64;                               // sources per work unit
constexpr int SPLIT_UNITS_PER_TILE = EVAL_TILE / SPLIT_UNIT;  // a ring stage holds up to one tile = 4 units
constexpr int QUEUE_MAX_LEVELS = 14;

struct EvalQueueParams {
  const void* r_trg;
  const void* rec;   // packed source records, padded to whole tiles
  void* v_trg;       // used when n_slots == 1
  double* partial;   // [n_slots][K1][nt_stride]
  int* ticket;       // global ticket counter, zero on entry
  long long Nt;
  long long nt_stride;
  int units_per_ttile;  // S = padded sources / SPLIT_UNIT
  int n_ttiles;
  int n_slots;          // items per target tile = partial-sum slots
  int n_levels;
  int lvl_ticket_end[QUEUE_MAX_LEVELS];  // tickets [lvl_ticket_end[l-1], lvl_ticket_end[l]) belong to level l
  int lvl_unit0[QUEUE_MAX_LEVELS];       // first source unit of level l (same in every target tile)
  int lvl_size[QUEUE_MAX_LEVELS];        // units per item (the last item of the last level may be shorter)
  int lvl_slot0[QUEUE_MAX_LEVELS];       // slot of the level's first item
  int accumulate;
  double scale;
};

struct QueueLoad {  // one TMA load of the CTA's stream; lives in the ring slot of its stage
  int ttile;
  int slot;
  int nunits;  // 0 = end of the queue
  int flags;   // 1 = first load of an item (load targets, zero sums), 2 = last load of an item (write the partial sum)
};

template <class Ker, class Real, int R, int NEWTON, int MINB>
__global__ void __launch_bounds__(EVAL_THREADS, MINB) eval_queue_kernel(const EvalQueueParams p) {
  constexpr int REC = rec_len<Ker, Real>();
  constexpr int NACC = Ker::NACC;
  constexpr int K1 = Ker::K1;
  constexpr uint32_t TILE_BYTES = EVAL_TILE * REC * sizeof(Real);
  constexpr uint32_t UNIT_BYTES = SPLIT_UNIT * REC * sizeof(Real);
  constexpr int TT = EVAL_THREADS * R;  // targets per target tile
  using V = typename VecOf<Real>::type;
  constexpr int NV = REC / VecOf<Real>::N;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  Real* tiles = reinterpret_cast<Real*>(smem_raw);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_raw + (size_t)EVAL_STAGES * TILE_BYTES);
  uint64_t* empty_bar = full_bar + EVAL_STAGES;
  QueueLoad* loads = reinterpret_cast<QueueLoad*>(empty_bar + EVAL_STAGES);  // [EVAL_STAGES]
  // producer state, touched by thread 0 only; kept in shared memory so that it costs the hot loop no registers:
  // [0] target tile, [1] slot, [2] next unit, [3] end unit of the item being streamed, [4] ticket fetched ahead, [5] first-load flag, [6] drained
  int* ps = reinterpret_cast<int*>(loads + EVAL_STAGES);

  const int tid = threadIdx.x;
  const Real* rec = reinterpret_cast<const Real*>(p.rec);

  // issues the next load of this CTA's stream into ring stage `st` (thread 0)
  auto issue = [&](int st) {
    if (ps[6]) return;
    int it_ttile = ps[0], it_slot = ps[1], it_u = ps[2], it_end = ps[3];
    bool it_first = ps[5] != 0;
    if (it_u >= it_end) {  // current item exhausted: take the ticket fetched ahead, fetch the one after it
      const int k = ps[4];
      if (k >= p.lvl_ticket_end[p.n_levels - 1]) {
        loads[st].nunits = 0;
        ps[6] = 1;
        mbar_arrive(&full_bar[st]);  // completes the phase without a copy: the consumers read the end marker
        return;
      }
      ps[4] = atomicAdd(p.ticket, 1);  // latency hidden behind the item that is about to stream
      int l = 0;
      while (k >= p.lvl_ticket_end[l]) l++;
      const int k2 = k - (l ? p.lvl_ticket_end[l - 1] : 0);
      const int j = k2 / p.n_ttiles;
      it_ttile = k2 - j * p.n_ttiles;  // consecutive tickets = consecutive target tiles on the SAME source range (L2 reuse)
      it_slot = p.lvl_slot0[l] + j;
      it_u = p.lvl_unit0[l] + j * p.lvl_size[l];
      it_end = min(it_u + p.lvl_size[l], p.units_per_ttile);
      it_first = true;
    }
    const int nu = min(SPLIT_UNITS_PER_TILE, it_end - it_u);
    loads[st].ttile = it_ttile;
    loads[st].slot = it_slot;
    loads[st].nunits = nu;
    loads[st].flags = (it_first ? 1 : 0) | (it_u + nu == it_end ? 2 : 0);
    const uint32_t bytes = (uint32_t)nu * UNIT_BYTES;
    mbar_arrive_expect_tx(&full_bar[st], bytes);
    bulk_g2s(tiles + (size_t)st * EVAL_TILE * REC, rec + (size_t)it_u * SPLIT_UNIT * REC, bytes, &full_bar[st]);
    ps[0] = it_ttile, ps[1] = it_slot, ps[2] = it_u + nu, ps[3] = it_end, ps[5] = 0;
  };

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < EVAL_STAGES; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], EVAL_THREADS / 32);
    }
    mbar_fence_init();
    ps[0] = ps[1] = ps[2] = ps[3] = ps[5] = ps[6] = 0;
    ps[4] = atomicAdd(p.ticket, 1);
  }
  __syncthreads();
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < EVAL_STAGES; s++) issue(s);
  }

  const Real* r_trg = reinterpret_cast<const Real*>(p.r_trg);
  Real xt[R], yt[R], zt[R];
  Real tacc[R][NACC];
  double acc[R][NACC];
  long long t_base = 0;

  for (int k = 0;; k++) {
    const int stage = k % EVAL_STAGES;
    const uint32_t parity = (uint32_t)(k / EVAL_STAGES) & 1u;
    mbar_wait(&full_bar[stage], parity);
    const QueueLoad L = loads[stage];
    if (L.nunits == 0) break;  // queue drained
    if (L.flags & 1) {         // new item: this CTA's targets change
      t_base = (long long)L.ttile * TT + tid;
#pragma unroll
      for (int r = 0; r < R; r++) {
        long long t = t_base + (long long)r * EVAL_THREADS;
        t = t < p.Nt ? t : p.Nt - 1;  // clamp: out-of-range lanes recompute the last target and are not stored
        xt[r] = r_trg[t * 3 + 0];
        yt[r] = r_trg[t * 3 + 1];
        zt[r] = r_trg[t * 3 + 2];
#pragma unroll
        for (int q = 0; q < NACC; q++) acc[r][q] = 0.0;
      }
    }
    const V* tile = reinterpret_cast<const V*>(tiles + (size_t)stage * EVAL_TILE * REC);
    const int ns = L.nunits * SPLIT_UNIT;

#pragma unroll
    for (int r = 0; r < R; r++) {
#pragma unroll
      for (int q = 0; q < NACC; q++) tacc[r][q] = (Real)0;
    }
#pragma unroll EVAL_UNROLL
    for (int s = 0; s < ns; s++) {
      Real src[REC];
#pragma unroll
      for (int i = 0; i < NV; i++) reinterpret_cast<V*>(src)[i] = tile[s * NV + i];  // warp-uniform LDS.128 broadcast
#pragma unroll
      for (int r = 0; r < R; r++) {
        const Real dx = xt[r] - src[0];
        const Real dy = yt[r] - src[1];
        const Real dz = zt[r] - src[2];
        Real r2p = dx * dx;
        r2p = fma(dy, dy, r2p);
        const Real r2 = fma(dz, dz, r2p);
        const Real rinv = rsqrt_unmasked<NEWTON>(r2, r2p);
        Ker::apply(tacc[r], dx, dy, dz, rinv, src + 3);
      }
    }
    repair_tile<Ker, Real, R, NEWTON>(tacc, xt, yt, zt, tile, ns);
#pragma unroll
    for (int r = 0; r < R; r++) {
#pragma unroll
      for (int q = 0; q < NACC; q++) acc[r][q] += (double)tacc[r][q];
    }

    // release the stage; warp 0 refills the slot of load k-1 with the next load of the stream (all 32 lanes poll, lane 0 issues)
    __syncwarp();
    if ((tid & 31) == 0) mbar_arrive(&empty_bar[stage]);
    if (tid < 32 && k >= 1) {
      const int j = k - 1;
      const int js = j % EVAL_STAGES;
      mbar_wait(&empty_bar[js], (uint32_t)(j / EVAL_STAGES) & 1u);
      if (tid == 0) issue(js);
      __syncwarp();
    }

    if (L.flags & 2) {  // end of item: its partial sum goes to its slot (or straight to v_trg when the target tile has one item)
      const int n_slots = p.n_slots;
#pragma unroll
      for (int r = 0; r < R; r++) {
        const long long t = t_base + (long long)r * EVAL_THREADS;
        if (t >= p.Nt) continue;
        double a[NACC], out[K1];
#pragma unroll
        for (int q = 0; q < NACC; q++) a[q] = acc[r][q];
        Ker::finalize(out, a);
        if (n_slots == 1) {
          Real* v = reinterpret_cast<Real*>(p.v_trg) + t * K1;
#pragma unroll
          for (int q = 0; q < K1; q++) {
            const double prev = p.accumulate ? (double)v[q] : 0.0;
            v[q] = (Real)(prev + out[q] * p.scale);
          }
        } else {
          double* dst = p.partial + (size_t)L.slot * K1 * p.nt_stride + t;
#pragma unroll
          for (int q = 0; q < K1; q++) dst[(size_t)q * p.nt_stride] = out[q];
        }
      }
    }
  }
}

template <class Ker, class Real>
constexpr size_t eval_queue_smem_bytes() {
  return (size_t)EVAL_STAGES * EVAL_TILE * rec_len<Ker, Real>() * sizeof(Real) + 2 * EVAL_STAGES * sizeof(uint64_t) + EVAL_STAGES * sizeof(QueueLoad) + 8 * sizeof(int);
}

}  // namespace sctl_b200
