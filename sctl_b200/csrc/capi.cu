// C ABI of libsctl_b200.so (declared in include/sctl_b200.h): argument checking, per-thread streams,
// stream-ordered device workspaces, the pack -> eval -> reduce launch plan, host<->device staging and the
// in-process multi-GPU target sharding. No CPU fallback anywhere: a missing device is an error.
#include <atomic>
#include <cmath>
#include <deque>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/sctl_b200.h"
#include "internal.h"
#include "launch.h"

namespace sctl_b200 {

// ------------------------------------------------------------------------------------------------
// bookkeeping
// ------------------------------------------------------------------------------------------------
static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static thread_local std::string t_last_error;
static thread_local double t_timing[5] = {0, 0, 0, 0, 0};

int set_error(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  t_last_error = buf;
  return code;
}
int cuda_error(cudaError_t e, const char* what) {
  return set_error(SCTL_B200_ERR_CUDA, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
}
const char* last_error_cstr() { return t_last_error.c_str(); }

// kernel_functions.hpp:15-198 (Name, FLOPS, uKerScaleFactor) and the dims deduced by generic-kernel.hpp:33-52
static const KernelInfo KINFO[NUM_KERNELS] = {
    {"Laplace3D-FxU", 0, 1, 1, 6, 1.0, 4.0},    {"Laplace3D-DxU", 3, 1, 1, 14, 1.0, 4.0}, {"Laplace3D-FxdU", 0, 1, 3, 11, -1.0, 4.0},
    {"Stokes3D-FxU", 0, 3, 3, 23, 1.0, 8.0},    {"Stokes3D-DxU", 3, 3, 3, 26, 3.0, 4.0},  {"Stokes3D-FxT", 0, 3, 9, 39, -3.0, 4.0},
    {"Stokes3D-FSxU", 0, 4, 3, 26, 1.0, 8.0},   {"Stokes3D-FxUP", 0, 3, 4, 26, 1.0, 8.0},
};
const KernelInfo& kernel_info(int kernel_id) { return KINFO[kernel_id]; }
static double kernel_scale(int kid) {
  return (double)((long double)KINFO[kid].scale_num / ((long double)KINFO[kid].scale_den_pi * 3.14159265358979323846264338327950288L));
}

// digits -> rsqrt refinement level (common.cuh). The reference relaxes the Newton count with digits
// (intrin-wrapper.hpp:3026: 14-bit seed for digits <= 4, one iteration = 2^-28 up to 8, two from 9). Here the levels are
// the bare seed (rel. err <= 2^-19, f32 2^-22), one Newton step (<= 2^-37) and full precision; a level is used only while
// its error stays below 0.2 * 10^-digits, so that the rinv^3 / rinv^5 kernels still hold `digits` digits
// (tests/test_gpu_rsqrt.py measures the three bounds on the device and checks this table against them).
static int newton_level(int dtype, int digits) {
  if (dtype == SCTL_B200_F64) {
    if (digits < 0 || digits > 10) return 2;
    return digits > 4 ? 1 : 0;
  }
  return (digits < 0 || digits > 5) ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------
// per-thread, per-device context: one non-blocking stream + timing events
// ------------------------------------------------------------------------------------------------
constexpr size_t STAGE_BYTES = 1u << 20;  // initial size of the per-thread staging arena (grows by doubling, see stage_reserve)
// The calling thread's contexts, one per device it has used. A deque: push_back never relocates existing elements, so a
// DevCtx* handed out by get_ctx stays valid when the thread later touches another device. The holder's destructor releases
// streams, events and staging blocks when the thread exits (long-lived threads may also call sctl_b200_finalize).
struct ThreadCtxHolder {
  std::deque<DevCtx> v;
  ~ThreadCtxHolder();
};
static thread_local ThreadCtxHolder t_holder;
#define t_ctx (t_holder.v)
static std::mutex g_devices_mutex;
static std::vector<int> g_devices;  // devices SCTL_B200_ALL_GPUS shards over (sctl_b200_init); empty = every visible device
static std::mutex g_pool_mutex;
static std::vector<int> g_pool_configured;

int get_ctx(DevCtx** out) {
  int dev = 0;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev < 1) {
    cudaGetLastError();
    return set_error(SCTL_B200_ERR_CUDA, "no usable CUDA device (%s); libsctl_b200 has no CPU fallback",
                     e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  CU_TRY(cudaGetDevice(&dev));
  for (auto& c : t_ctx)
    if (c.dev == dev) {
      *out = &c;
      return 0;
    }
  DevCtx c;
  c.dev = dev;
  CU_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  for (auto& ev : c.ev) CU_TRY(cudaEventCreate(&ev));
  CU_TRY(cudaDeviceGetAttribute(&c.sms, cudaDevAttrMultiProcessorCount, dev));
  {
    // keep stream-ordered allocations cached in the default pool instead of returning them at every sync
    std::lock_guard<std::mutex> lk(g_pool_mutex);
    bool done = false;
    for (int d : g_pool_configured) done |= (d == dev);
    if (!done) {
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        unsigned long long thr = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
      }
      g_pool_configured.push_back(dev);
    }
  }
  t_ctx.push_back(c);
  *out = &t_ctx.back();
  return 0;
}

// Releases the calling thread's streams, events and staging blocks. Long-lived caller threads (the main thread, an OpenMP
// team) keep theirs for the life of the process; the short-lived workers of eval_host_all_gpus call this before they exit.
void release_thread_ctx() {
  if (t_ctx.empty()) return;
  DeviceGuard guard;
  for (auto& c : t_ctx) {
    if (cudaSetDevice(c.dev) != cudaSuccess) continue;
    if (c.stream) cudaStreamDestroy(c.stream);
    for (auto& ev : c.ev)
      if (ev) cudaEventDestroy(ev);
    if (c.h_stage) cudaFreeHost(c.h_stage);
    if (c.d_stage) cudaFree(c.d_stage);
  }
  t_ctx.clear();
  cudaGetLastError();
}
ThreadCtxHolder::~ThreadCtxHolder() { release_thread_ctx(); }

size_t elem_size(int dtype) { return dtype == SCTL_B200_F64 ? 8 : 4; }

int check_common(int kernel_id, int dtype, int64_t Nt, int64_t Ns) {
  if (kernel_id < 0 || kernel_id >= NUM_KERNELS) return set_error(SCTL_B200_ERR_KERNEL, "unknown kernel id %d", kernel_id);
  if (dtype != SCTL_B200_F32 && dtype != SCTL_B200_F64) return set_error(SCTL_B200_ERR_DTYPE, "unknown dtype %d", dtype);
  if (Nt < 0 || Ns < 0) return set_error(SCTL_B200_ERR_ARG, "negative size Nt=%lld Ns=%lld", (long long)Nt, (long long)Ns);
  return 0;
}

int DevAlloc::alloc(size_t bytes, cudaStream_t s) {
  st = s;
  cudaError_t e = cudaMallocAsync(&p, bytes ? bytes : 16, s);
  if (e != cudaSuccess) {
    p = nullptr;
    cudaGetLastError();
    return set_error(SCTL_B200_ERR_NOMEM, "cudaMallocAsync(%zu bytes): %s", bytes, cudaGetErrorString(e));
  }
  return 0;
}
int DevBuf::reserve(size_t bytes) {
  if (bytes <= cap) return 0;
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) {
    cudaGetLastError();
    p = nullptr;
    return set_error(SCTL_B200_ERR_NOMEM, "cudaMalloc(%zu bytes): %s", bytes, cudaGetErrorString(e));
  }
  cap = bytes;
  return 0;
}

// ------------------------------------------------------------------------------------------------
// launch plan for Eval on device-resident AoS arrays
// ------------------------------------------------------------------------------------------------
struct EvalPlan {
  int rvar, newton, n_chunks, tiles_per_chunk;
  long long n_tiles, n_ttiles, Ns_pad, nt_stride;
};

static int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return (s && *s) ? atoi(s) : dflt;
}

static EvalPlan make_plan(const EvalEntry& E, int dtype, int64_t Nt, int64_t Ns, int digits, int sms) {
  EvalPlan P{};
  P.newton = newton_level(dtype, digits);
  // The register-blocked variant (R targets per thread) needs enough CTAs to balance 148 SMs; problems that give it fewer
  // than ~7 CTAs per SM at one tile per chunk (N = M below ~15k), and small target sets (ragged last tile), run the R=1
  // variant: 128 targets per CTA, 6x the CTAs (profiles/r01q_small_n_rvar.txt: 0.129 vs 0.143 ms at 8k x 8k, 0.065 vs
  // 0.107 ms at 4k x 4k; the blocked variant wins from 20k x 20k up).
  P.n_tiles = (Ns + EVAL_TILE - 1) / EVAL_TILE;
  const long long ttiles_big = (Nt + (long long)EVAL_THREADS * E.R[1] - 1) / ((long long)EVAL_THREADS * E.R[1]);
  P.rvar = (Nt >= (int64_t)env_int("SCTL_B200_RBIG_MIN_NT", 2048) && ttiles_big * P.n_tiles >= (long long)env_int("SCTL_B200_RBIG_MIN_CTAS_PER_SM", 7) * sms) ? 1 : 0;
  const int R = E.R[P.rvar];
  P.Ns_pad = P.n_tiles * EVAL_TILE;
  P.n_ttiles = (Nt + (long long)EVAL_THREADS * R - 1) / ((long long)EVAL_THREADS * R);
  P.nt_stride = (Nt + 31) / 32 * 32;
  // Split the source range into chunks (blockIdx.y): aim at ~64 CTAs per resident slot, at most 1 GiB of partial sums.
  // Many short CTAs beat few long ones: the block scheduler refills slots as they free up, so short CTAs balance the SMs
  // dynamically, while a wave-quantised plan leaves SMs idle behind the last wave (measured with scripts/sweep_chunks.py,
  // profiles/r01o_chunk_sweep.txt: at 20k x 20k 1-3 tiles per chunk run in 0.33-0.35 ms, 4-8 tiles in 0.42-0.56 ms; a
  // cost model built on whole waves picked the latter). Two tiles per chunk keep the per-CTA prologue below 2 %; one tile
  // per chunk only when even that leaves fewer than 4 CTAs per slot.
  const int occ = E.occupancy[P.rvar][P.newton]();
  const long long slots = (long long)occ * sms;
  const long long want = (long long)env_int("SCTL_B200_CTAS_PER_SLOT", 64) * slots;
  long long chunks = (want + P.n_ttiles - 1) / P.n_ttiles;
  long long max_by_tiles = P.n_tiles / 2 > 0 ? P.n_tiles / 2 : 1;
  if (max_by_tiles * P.n_ttiles < 4 * slots) max_by_tiles = P.n_tiles;
  const long long max_by_mem = (long long)((1ull << 30) / ((size_t)E.k1 * 8 * (size_t)P.nt_stride));
  if (chunks > max_by_tiles) chunks = max_by_tiles;
  if (chunks > max_by_mem) chunks = max_by_mem;
  if (chunks > 65535) chunks = 65535;
  if (chunks < 1) chunks = 1;
  const int forced = env_int("SCTL_B200_FORCE_CHUNKS", 0);
  if (forced > 0) chunks = forced < P.n_tiles ? forced : P.n_tiles;
  P.tiles_per_chunk = (int)((P.n_tiles + chunks - 1) / chunks);
  P.n_chunks = (int)((P.n_tiles + P.tiles_per_chunk - 1) / P.tiles_per_chunk);
  return P;
}

// pack: AoS (r_src, n_src, v_src) -> tile-padded source records in `rec` (device). Ns_pad records are written.
static int64_t padded_sources(int64_t Ns) { return (Ns + EVAL_TILE - 1) / EVAL_TILE * EVAL_TILE; }
static int pack_device(int kernel_id, int dtype, int64_t Ns, const void* r_src, const void* n_src, const void* v_src, const void* src_wts, void* rec,
                       cudaStream_t st) {
  const EvalEntry& E = (dtype == SCTL_B200_F64 ? eval_table_f64() : eval_table_f32())[kernel_id];
  if (Ns == 0) return 0;
  CU_TRY(E.pack(r_src, n_src, v_src, src_wts, rec, Ns, padded_sources(Ns), st));
  return 0;
}

// The O(Ns*Nt) part on pre-packed records. All pointers are device pointers on the current device.
// Enqueues on `st`, does not synchronise.
static int eval_packed_device(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* rec, void* v_trg, int digits,
                              bool accumulate, cudaStream_t st, int sms, cudaEvent_t* ev) {
  const EvalEntry& E = (dtype == SCTL_B200_F64 ? eval_table_f64() : eval_table_f32())[kernel_id];
  const EvalPlan P = make_plan(E, dtype, Nt, Ns, digits, sms);
  DevAlloc partial;
  if (P.n_chunks > 1) {
    int rc = partial.alloc((size_t)P.n_chunks * E.k1 * (size_t)P.nt_stride * sizeof(double), st);
    if (rc) return rc;
  }
  EvalParams p{};
  p.r_trg = r_trg;
  p.rec = rec;
  p.v_trg = v_trg;
  p.partial = (double*)partial.p;
  p.Nt = Nt;
  p.nt_stride = P.nt_stride;
  p.n_tiles = (int)P.n_tiles;
  p.tiles_per_chunk = P.tiles_per_chunk;
  p.n_chunks = P.n_chunks;
  p.accumulate = accumulate ? 1 : 0;
  p.scale = std::ldexp(kernel_scale(kernel_id), -E.x2_shift[P.newton]);  // X2 mode accumulates an exact power of two times the sum
  CU_TRY(E.eval[P.rvar][P.newton](p, P.n_ttiles, st));
  if (ev) CU_TRY(cudaEventRecord(ev[0], st));
  if (P.n_chunks > 1) CU_TRY(E.reduce((const double*)partial.p, v_trg, Nt, P.nt_stride, P.n_chunks, p.scale, p.accumulate, st));
  if (ev) CU_TRY(cudaEventRecord(ev[1], st));
  return 0;
}

// Small problems take the one-launch kernel of small_impl.cuh: every thread there walks Ns/8 sources, so it wins while the
// tiled path cannot fill the SMs (thresholds measured with scripts/latency.py, profiles/r01r_small_path.txt).
static bool use_small_path(int64_t Nt, int64_t Ns) {
  return Ns <= (int64_t)env_int("SCTL_B200_SMALL_MAX_NS", 8192) && Nt * Ns <= (int64_t)env_int("SCTL_B200_SMALL_MAX_PAIRS", 50000000);
}

// pack + eval (+ reduce). ev (optional): 4 events recorded around pack | eval | reduce.
int eval_device(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src, const void* v_src,
                void* v_trg, int digits, bool accumulate, cudaStream_t st, int sms, cudaEvent_t* ev) {
  const EvalEntry& E = (dtype == SCTL_B200_F64 ? eval_table_f64() : eval_table_f32())[kernel_id];
  const size_t esz = elem_size(dtype);
  if (Nt == 0) return 0;
  if (Ns == 0) {
    if (!accumulate) CU_TRY(cudaMemsetAsync(v_trg, 0, (size_t)Nt * E.k1 * esz, st));
    if (ev)
      for (int i = 0; i < 4; i++) CU_TRY(cudaEventRecord(ev[i], st));
    return 0;
  }
  if (Ns > (int64_t)EVAL_TILE * 0x7fffffffLL) return set_error(SCTL_B200_ERR_ARG, "Ns too large");
  if (use_small_path(Nt, Ns)) {  // one launch: no record buffer, no partial sums
    if (ev) CU_TRY(cudaEventRecord(ev[0], st));
    if (ev) CU_TRY(cudaEventRecord(ev[1], st));
    CU_TRY(E.small[newton_level(dtype, digits)](r_trg, r_src, n_src, v_src, nullptr, v_trg, Nt, Ns, kernel_scale(kernel_id), accumulate ? 1 : 0, sms, st));
    if (ev) CU_TRY(cudaEventRecord(ev[2], st));
    if (ev) CU_TRY(cudaEventRecord(ev[3], st));
    return 0;
  }
  DevAlloc rec;
  int rc = rec.alloc((size_t)padded_sources(Ns) * E.rec_len * esz, st);
  if (rc) return rc;
  if (ev) CU_TRY(cudaEventRecord(ev[0], st));
  if ((rc = pack_device(kernel_id, dtype, Ns, r_src, n_src, v_src, nullptr, rec.p, st))) return rc;
  if (ev) CU_TRY(cudaEventRecord(ev[1], st));
  return eval_packed_device(kernel_id, dtype, Nt, Ns, r_trg, rec.p, v_trg, digits, accumulate, st, sms, ev ? ev + 2 : nullptr);
}

// The tiled path (pack -> eval_kernel -> reduce) with its working memory placed in a block the CALLER owns instead of two
// stream-ordered allocations: used by sctl_b200_eval_noalloc (noalloc.cu), which hands in a file-scope __device__ array. No
// allocation of any kind happens here. Layout of `scratch` (a device address on the current device, 256-byte aligned):
//   [0, rec_bytes)            packed source records, all Ns_pad of them written by the pack kernel
//   [round_up(rec_bytes,256)) chunk partial sums [n_chunks][K1][nt_stride] f64, every element that reduce_kernel reads
//                             (t < Nt) written by eval_kernel's epilogue
// so nothing in the block has to start at zero and nothing of an earlier call is read. The plan is the one eval_device makes,
// except that the number of source chunks is cut down to what the rest of the block holds (one chunk needs no partial sums
// at all: that CTA column then writes v_trg directly). Records larger than rec_max_bytes are refused.
int eval_device_in_scratch(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                           const void* v_src, void* v_trg, int digits, bool accumulate, cudaStream_t st, int sms, void* scratch,
                           size_t scratch_bytes, size_t rec_max_bytes) {
  const EvalEntry& E = (dtype == SCTL_B200_F64 ? eval_table_f64() : eval_table_f32())[kernel_id];
  const size_t esz = elem_size(dtype);
  if (Nt == 0) return 0;
  if (Ns == 0) {
    if (!accumulate) CU_TRY(cudaMemsetAsync(v_trg, 0, (size_t)Nt * E.k1 * esz, st));
    return 0;
  }
  const size_t rec_one = (size_t)E.rec_len * esz;
  if ((uint64_t)Ns > rec_max_bytes / rec_one || (size_t)padded_sources(Ns) * rec_one > rec_max_bytes)
    return set_error(SCTL_B200_ERR_ARG,
                     "sctl_b200_eval_noalloc: %lld sources of %s need %llu bytes of packed records (%zu bytes each, padded to a multiple of %d "
                     "sources), the static scratch block holds at most %zu: at most %lld sources for this kernel and dtype - use "
                     "sctl_b200_eval_async for larger source sets",
                     (long long)Ns, KINFO[kernel_id].name, (unsigned long long)padded_sources(Ns) * rec_one, rec_one, EVAL_TILE, rec_max_bytes,
                     (long long)(rec_max_bytes / rec_one / EVAL_TILE * EVAL_TILE));
  const size_t rec_bytes = (size_t)padded_sources(Ns) * rec_one;
  const size_t part_off = (rec_bytes + 255) & ~(size_t)255;
  EvalPlan P = make_plan(E, dtype, Nt, Ns, digits, sms);
  const size_t per_chunk = (size_t)E.k1 * (size_t)P.nt_stride * sizeof(double);
  const size_t fit = part_off < scratch_bytes ? (scratch_bytes - part_off) / per_chunk : 0;
  if (P.n_chunks > 1 && (size_t)P.n_chunks > fit) {
    const long long chunks = fit >= 2 ? (long long)fit : 1;
    P.tiles_per_chunk = (int)((P.n_tiles + chunks - 1) / chunks);
    P.n_chunks = (int)((P.n_tiles + P.tiles_per_chunk - 1) / P.tiles_per_chunk);
  }
  char* rec = (char*)scratch;
  CU_TRY(E.pack(r_src, n_src, v_src, nullptr, rec, Ns, padded_sources(Ns), st));
  EvalParams p{};
  p.r_trg = r_trg;
  p.rec = rec;
  p.v_trg = v_trg;
  p.partial = P.n_chunks > 1 ? (double*)((char*)scratch + part_off) : nullptr;
  p.Nt = Nt;
  p.nt_stride = P.nt_stride;
  p.n_tiles = (int)P.n_tiles;
  p.tiles_per_chunk = P.tiles_per_chunk;
  p.n_chunks = P.n_chunks;
  p.accumulate = accumulate ? 1 : 0;
  p.scale = std::ldexp(kernel_scale(kernel_id), -E.x2_shift[P.newton]);
  CU_TRY(E.eval[P.rvar][P.newton](p, P.n_ttiles, st));
  if (P.n_chunks > 1) CU_TRY(E.reduce(p.partial, v_trg, Nt, P.nt_stride, P.n_chunks, p.scale, p.accumulate, st));
  return 0;
}

// Sources per CTA of the KernelMatrix kernels (measured, profiles/r02l_kmat_sources_per_cta.txt). Blocks with several values
// per target (K1out >= 3) write fastest with 8 sources per CTA at every size (Stokes FxU f64: 111 % of the measured copy
// bandwidth against 106 % at 32; the batched SetupNear shape: 6.2 against 5.5 TB/s); thin rows (one value per target, a 16-byte
// store per lane) want the long 32-source loop unless the grid would not even fill two waves of CTAs.
static int kmat_sources_per_cta(long long ctas_at_32, int sms, bool thin) {
  const long long wave = 16ll * sms;  // 16 resident 128-thread CTAs per SM
  int sb;
  if (thin) sb = ctas_at_32 >= 2 * wave ? 32 : (ctas_at_32 >= wave / 2 ? 16 : 8);
  else sb = ctas_at_32 * 4 >= 2 * wave ? 8 : 4;
  const int forced = env_int("SCTL_B200_KMAT_SB", 0);
  return (forced == 4 || forced == 8 || forced == 16 || forced == 32) ? forced : sb;
}

static int kmat_device(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src, void* M,
                       int digits, cudaStream_t st, int sms) {
  const EvalEntry& E = (dtype == SCTL_B200_F64 ? eval_table_f64() : eval_table_f32())[kernel_id];
  if (Nt == 0 || Ns == 0) return 0;
  const size_t esz = elem_size(dtype);
  const int newton = newton_level(dtype, digits);
  const long long ct = E.kmat_cta_targets[0];
  const int sb = kmat_sources_per_cta(((Nt + ct - 1) / ct) * ((Ns + KMAT_SB - 1) / KMAT_SB), sms, E.k1 == 1);
  // grid.y is limited to 65535 blocks of sb sources: walk the source range in slabs
  const int64_t slab = (int64_t)65535 * sb;
  for (int64_t s0 = 0; s0 < Ns; s0 += slab) {
    const int64_t ns = Ns - s0 < slab ? Ns - s0 : slab;
    const char* rs = (const char*)r_src + (size_t)s0 * 3 * esz;
    const char* nn = n_src ? (const char*)n_src + (size_t)s0 * E.nd * esz : nullptr;
    char* Mp = (char*)M + (size_t)s0 * E.k0 * (size_t)Nt * E.k1 * esz;
    CU_TRY(E.kmat[newton](r_trg, rs, nn, Mp, Nt, ns, kernel_scale(kernel_id), sb, st));
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------
// host-pointer paths (what the C++ header shim calls): stage in, run, stage out, synchronise
// ------------------------------------------------------------------------------------------------
static float ev_ms(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0;
  if (cudaEventElapsedTime(&ms, a, b) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return ms;
}

// Host-pointer calls whose arrays total up to STAGE_MAX_BYTES go through a per-thread, grow-only staging ARENA: one pinned host
// block and one device block of the same size, kept for the thread's lifetime. The general path below costs five pageable
// copies (each staged by the driver through its own bounce buffer, one after the other), five stream-ordered allocations and
// six event records - more than the sum itself for the call sizes the reference's in-tree callers make (72 us wall at 64 x 64,
// +40 % at 20k x 20k, profiles/r01r_small_path.txt). Here every array is copied into the pinned block by the CPU and sent with its
// own asynchronous copy straight away, so the CPU copy of array i+1 overlaps the bus transfer of array i; no allocation, two
// events, one synchronisation.
constexpr size_t STAGE_MAX_BYTES = 64u << 20;
static int stage_reserve(DevCtx* c, size_t bytes) {
  if (bytes <= c->stage_cap) return 0;
  size_t cap = STAGE_BYTES;
  while (cap < bytes) cap *= 2;
  if (c->h_stage) cudaFreeHost(c->h_stage);
  if (c->d_stage) cudaFree(c->d_stage);
  c->h_stage = c->d_stage = nullptr;
  c->stage_cap = 0;
  cudaError_t e = cudaMallocHost(&c->h_stage, cap);
  if (e == cudaSuccess) e = cudaMalloc(&c->d_stage, cap);
  if (e != cudaSuccess) {
    if (c->h_stage) cudaFreeHost(c->h_stage);
    c->h_stage = c->d_stage = nullptr;
    cudaGetLastError();
    return set_error(SCTL_B200_ERR_NOMEM, "staging arena of %zu bytes: %s", cap, cudaGetErrorString(e));
  }
  c->stage_cap = cap;
  return 0;
}

static int eval_host_staged(DevCtx* c, int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                            const void* v_src, void* v_trg, int digits, bool accumulate, const size_t (&bytes)[5], const size_t (&off)[6]) {
  cudaStream_t st = c->stream;
  int rc = stage_reserve(c, off[5]);
  if (rc) return rc;
  char* h = (char*)c->h_stage;
  char* d = (char*)c->d_stage;
  const void* in[5] = {r_trg, r_src, n_src, v_src, accumulate ? v_trg : nullptr};
  cudaEvent_t* ev = c->ev;
  // Tiny calls: ZERO-COPY. The pinned block is mapped into the device address space (UVA), so the one-launch kernel reads
  // the arrays over the bus itself and writes v_trg straight back into host memory: one launch + one synchronisation instead
  // of copy + launch + copy (each a separate device operation of 2-3 us).
  if (off[5] <= (size_t)env_int("SCTL_B200_ZEROCOPY_MAX_BYTES", 32768) && Ns > 0 && use_small_path(Nt, Ns)) {
    for (int i = 0; i < 5; i++)
      if (in[i] && bytes[i]) memcpy(h + off[i], in[i], bytes[i]);
    const EvalEntry& E = (dtype == SCTL_B200_F64 ? eval_table_f64() : eval_table_f32())[kernel_id];
    CU_TRY(E.small[newton_level(dtype, digits)](h + off[0], h + off[1], h + off[2], h + off[3], nullptr, h + off[4], Nt, Ns, kernel_scale(kernel_id),
                                                accumulate ? 1 : 0, c->sms, st));
    CU_TRY(cudaStreamSynchronize(st));
    memcpy(v_trg, h + off[4], bytes[4]);
    t_timing[0] = t_timing[1] = t_timing[2] = t_timing[3] = t_timing[4] = 0;  // not timed on the device: no events on this path
    return 0;
  }
  CU_TRY(cudaEventRecord(ev[4], st));
  if (off[5] <= STAGE_BYTES) {  // small: one gathered copy
    for (int i = 0; i < 5; i++)
      if (in[i] && bytes[i]) memcpy(h + off[i], in[i], bytes[i]);
    const size_t in_bytes = accumulate ? off[5] : off[4];
    CU_TRY(cudaMemcpyAsync(d, h, in_bytes, cudaMemcpyHostToDevice, st));
  } else {  // larger: array by array, so that the CPU copy of the next array overlaps the transfer of this one
    for (int i = 0; i < 5; i++) {
      if (!in[i] || !bytes[i]) continue;
      memcpy(h + off[i], in[i], bytes[i]);
      CU_TRY(cudaMemcpyAsync(d + off[i], h + off[i], bytes[i], cudaMemcpyHostToDevice, st));
    }
  }
  rc = eval_device(kernel_id, dtype, Nt, Ns, d + off[0], d + off[1], d + off[2], d + off[3], d + off[4], digits, accumulate, st, c->sms, nullptr);
  if (rc) return rc;
  CU_TRY(cudaMemcpyAsync(h + off[4], d + off[4], bytes[4], cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaEventRecord(ev[5], st));
  CU_TRY(cudaStreamSynchronize(st));
  memcpy(v_trg, h + off[4], bytes[4]);
  t_timing[0] = t_timing[1] = t_timing[3] = t_timing[4] = 0;
  t_timing[2] = ev_ms(ev[4], ev[5]);  // copies included: the stages are not timed separately on this path
  return 0;
}

static int eval_host_one_device(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                                const void* v_src, void* v_trg, int digits, bool accumulate) {
  DevCtx* c = nullptr;
  int rc = get_ctx(&c);
  if (rc) return rc;
  const KernelInfo& ki = KINFO[kernel_id];
  const size_t esz = elem_size(dtype);
  cudaStream_t st = c->stream;
  if (Nt == 0) return 0;
  {
    const size_t bytes[5] = {(size_t)Nt * 3 * esz, (size_t)Ns * 3 * esz, (size_t)Ns * ki.normal_dim * esz, (size_t)Ns * ki.src_dim * esz,
                             (size_t)Nt * ki.trg_dim * esz};
    size_t off[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 5; i++) off[i + 1] = off[i] + (bytes[i] + 255) / 256 * 256;
    if (off[5] <= STAGE_MAX_BYTES && (size_t)env_int("SCTL_B200_STAGE_MAX_BYTES", (int)STAGE_MAX_BYTES) >= off[5])
      return eval_host_staged(c, kernel_id, dtype, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg, digits, accumulate, bytes, off);
  }
  DevAlloc d_rt, d_rs, d_ns, d_vs, d_vt;
  const size_t b_rt = (size_t)Nt * 3 * esz, b_rs = (size_t)Ns * 3 * esz, b_ns = (size_t)Ns * ki.normal_dim * esz,
               b_vs = (size_t)Ns * ki.src_dim * esz, b_vt = (size_t)Nt * ki.trg_dim * esz;
  if ((rc = d_rt.alloc(b_rt, st)) || (rc = d_rs.alloc(b_rs, st)) || (rc = d_ns.alloc(b_ns, st)) || (rc = d_vs.alloc(b_vs, st)) ||
      (rc = d_vt.alloc(b_vt, st)))
    return rc;
  cudaEvent_t* ev = c->ev;
  CU_TRY(cudaEventRecord(ev[4], st));
  CU_TRY(cudaMemcpyAsync(d_rt.p, r_trg, b_rt, cudaMemcpyHostToDevice, st));
  if (b_rs) CU_TRY(cudaMemcpyAsync(d_rs.p, r_src, b_rs, cudaMemcpyHostToDevice, st));
  if (b_ns) CU_TRY(cudaMemcpyAsync(d_ns.p, n_src, b_ns, cudaMemcpyHostToDevice, st));
  if (b_vs) CU_TRY(cudaMemcpyAsync(d_vs.p, v_src, b_vs, cudaMemcpyHostToDevice, st));
  if (accumulate) CU_TRY(cudaMemcpyAsync(d_vt.p, v_trg, b_vt, cudaMemcpyHostToDevice, st));
  rc = eval_device(kernel_id, dtype, Nt, Ns, d_rt.p, d_rs.p, d_ns.p, d_vs.p, d_vt.p, digits, accumulate, st, c->sms, ev);
  if (rc) return rc;
  CU_TRY(cudaMemcpyAsync(v_trg, d_vt.p, b_vt, cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaEventRecord(ev[5], st));
  CU_TRY(cudaStreamSynchronize(st));
  t_timing[0] = ev_ms(ev[4], ev[0]);
  t_timing[1] = ev_ms(ev[0], ev[1]);
  t_timing[2] = ev_ms(ev[1], ev[2]);
  t_timing[3] = ev_ms(ev[2], ev[3]);
  t_timing[4] = ev_ms(ev[3], ev[5]);
  return 0;
}

// SCTL_B200_ALL_GPUS: contiguous target blocks [Nt*g/G, Nt*(g+1)/G) per visible device — the partition
// formula of ParticleFMM::EvalDirect (fmm-wrapper.txx:507) — sources replicated by independent H2D
// copies (they start on the host, so there is no collective), one host thread per device.
static int eval_host_all_gpus(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                              const void* v_src, void* v_trg, int digits, bool accumulate) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
    cudaGetLastError();
    return set_error(SCTL_B200_ERR_CUDA, "no usable CUDA device; libsctl_b200 has no CPU fallback");
  }
  std::vector<int> devs;
  {
    std::lock_guard<std::mutex> lk(g_devices_mutex);
    devs = g_devices;
  }
  if (devs.empty())
    for (int g = 0; g < ndev; g++) devs.push_back(g);
  ndev = (int)devs.size();
  int dev0 = 0;
  CU_TRY(cudaGetDevice(&dev0));
  if (ndev == 1 || Nt < ndev) {
    if (ndev == 1 && devs[0] != dev0) CU_TRY(cudaSetDevice(devs[0]));
    const int rc1 = eval_host_one_device(kernel_id, dtype, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg, digits, accumulate);
    if (ndev == 1 && devs[0] != dev0) cudaSetDevice(dev0);
    return rc1;
  }
  const KernelInfo& ki = KINFO[kernel_id];
  const size_t esz = elem_size(dtype);
  // Sources: every GPU uploads only ITS 1/G block of the caller's arrays and the blocks are replicated over NVLink by the NCCL
  // driver of comm.cu (ncclCommInitAll communicators, one per device, cached for the life of the process) - the caller's pageable
  // memory crosses PCIe once instead of G times. SCTL_B200_ALL_GPUS_MODE=replicate (or a process without NCCL) keeps the
  // collective-free form: G independent uploads of all sources.
  static std::mutex all_gpus_mutex;  // the communicators carry one collective sequence at a time
  std::lock_guard<std::mutex> all_lk(all_gpus_mutex);
  std::vector<Comm*> comms;
  const char* mode = getenv("SCTL_B200_ALL_GPUS_MODE");
  const bool want_nccl = !(mode && !strcmp(mode, "replicate")) && Ns >= ndev;
  if (want_nccl && comm_all_gpus(devs, comms) != 0) comms.clear();  // NCCL unavailable: replicate
  std::vector<int64_t> counts(ndev);
  for (int g = 0; g < ndev; g++) counts[g] = Ns * (g + 1) / ndev - Ns * g / ndev;
  std::vector<int> rcs(ndev, 0);
  std::vector<std::string> msgs(ndev);
  std::vector<std::thread> th;
  for (int g = 0; g < ndev; g++) {
    th.emplace_back([&, g]() {
      if (cudaSetDevice(devs[g]) != cudaSuccess) {
        rcs[g] = SCTL_B200_ERR_CUDA;
        msgs[g] = "cudaSetDevice failed";
        return;
      }
      const int64_t t0 = Nt * g / ndev, t1 = Nt * (g + 1) / ndev;
      const char* xt = (const char*)r_trg + (size_t)t0 * 3 * esz;
      char* ut = (char*)v_trg + (size_t)t0 * ki.trg_dim * esz;
      if (!comms.empty()) {
        const int64_t s0 = Ns * g / ndev;
        DevCtx* c = nullptr;
        rcs[g] = get_ctx(&c);
        if (!rcs[g])
          rcs[g] = eval_sharded_impl(comms[g], kernel_id, dtype, t1 - t0, counts[g], xt, (const char*)r_src + (size_t)s0 * 3 * esz,
                                     ki.normal_dim ? (const char*)n_src + (size_t)s0 * ki.normal_dim * esz : nullptr,
                                     (const char*)v_src + (size_t)s0 * ki.src_dim * esz, ut, digits, accumulate ? 0u : (unsigned)SCTL_B200_OVERWRITE,
                                     counts.data(), c->stream, true);
      } else {
        rcs[g] = eval_host_one_device(kernel_id, dtype, t1 - t0, Ns, xt, r_src, n_src, v_src, ut, digits, accumulate);
      }
      if (rcs[g]) msgs[g] = t_last_error;
      release_thread_ctx();
    });
  }
  for (auto& t : th) t.join();
  cudaSetDevice(dev0);
  for (int g = 0; g < ndev; g++)
    if (rcs[g]) return set_error(rcs[g], "device %d: %s", devs[g], msgs[g].c_str());
  return 0;
}

static int kmat_host(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src, void* M,
                     int digits) {
  DevCtx* c = nullptr;
  int rc = get_ctx(&c);
  if (rc) return rc;
  if (Nt == 0 || Ns == 0) return 0;
  const KernelInfo& ki = KINFO[kernel_id];
  const size_t esz = elem_size(dtype);
  cudaStream_t st = c->stream;
  DevAlloc d_rt, d_rs, d_ns, d_M;
  const size_t b_rt = (size_t)Nt * 3 * esz, b_rs = (size_t)Ns * 3 * esz, b_ns = (size_t)Ns * ki.normal_dim * esz;
  const size_t row_bytes = (size_t)ki.src_dim * (size_t)Nt * ki.trg_dim * esz;  // bytes of M per source
  // device slab of at most ~1 GiB of M at a time (rows of consecutive sources are contiguous in M)
  int64_t slab = (int64_t)((1ull << 30) / (row_bytes ? row_bytes : 1));
  if (slab < 1) slab = 1;
  if (slab > Ns) slab = Ns;
  if ((rc = d_rt.alloc(b_rt, st)) || (rc = d_rs.alloc(b_rs, st)) || (rc = d_ns.alloc(b_ns, st)) || (rc = d_M.alloc((size_t)slab * row_bytes, st)))
    return rc;
  cudaEvent_t* ev = c->ev;
  CU_TRY(cudaEventRecord(ev[4], st));
  CU_TRY(cudaMemcpyAsync(d_rt.p, r_trg, b_rt, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(d_rs.p, r_src, b_rs, cudaMemcpyHostToDevice, st));
  if (b_ns) CU_TRY(cudaMemcpyAsync(d_ns.p, n_src, b_ns, cudaMemcpyHostToDevice, st));
  CU_TRY(cudaEventRecord(ev[0], st));
  for (int64_t s0 = 0; s0 < Ns; s0 += slab) {
    const int64_t ns = Ns - s0 < slab ? Ns - s0 : slab;
    rc = kmat_device(kernel_id, dtype, Nt, ns, d_rt.p, (const char*)d_rs.p + (size_t)s0 * 3 * esz,
                     b_ns ? (const char*)d_ns.p + (size_t)s0 * ki.normal_dim * esz : nullptr, d_M.p, digits, st, c->sms);
    if (rc) return rc;
    CU_TRY(cudaMemcpyAsync((char*)M + (size_t)s0 * row_bytes, d_M.p, (size_t)ns * row_bytes, cudaMemcpyDeviceToHost, st));
  }
  CU_TRY(cudaEventRecord(ev[5], st));
  CU_TRY(cudaStreamSynchronize(st));
  t_timing[0] = ev_ms(ev[4], ev[0]);
  t_timing[1] = 0;
  t_timing[2] = ev_ms(ev[0], ev[5]);
  t_timing[3] = 0;
  t_timing[4] = 0;
  return 0;
}

// ------------------------------------------------------------------------------------------------
// batched small-block KernelMatrix (one launch for all blocks)
// ------------------------------------------------------------------------------------------------
static int kmat_batch_impl(int kernel_id, int dtype, int64_t nbatch, const int64_t* trg_dsp, const int64_t* src_dsp, const int64_t* M_dsp,
                           const void* r_trg, const void* r_src, const void* n_src, const void* src_wts, const void* n_trg, void* M, int64_t M_len,
                           int digits, bool device_ptrs) {
  DevCtx* c = nullptr;
  int rc = get_ctx(&c);
  if (rc) return rc;
  const KernelInfo& ki = KINFO[kernel_id];
  const EvalEntry& E = (dtype == SCTL_B200_F64 ? eval_table_f64() : eval_table_f32())[kernel_id];
  const size_t esz = elem_size(dtype);
  const int dotn = n_trg ? 1 : 0;
  const int newton = newton_level(dtype, digits);
  if (dotn && !E.kmat_batch[1][newton]) return set_error(SCTL_B200_ERR_ARG, "target-normal contraction needs TrgDim %% 3 == 0 (%s has %d)", ki.name, ki.trg_dim);
  const int k1out = dotn ? ki.trg_dim / 3 : ki.trg_dim;
  // validate the CSR description and build the CTA prefix sum (host side, O(nbatch))
  std::vector<long long> desc(4 * (size_t)nbatch + 3);
  long long* h_trg = desc.data();
  long long* h_src = h_trg + nbatch + 1;
  long long* h_M = h_src + nbatch + 1;
  long long* h_cta = h_M + nbatch;
  h_cta[0] = 0;
  int sb = KMAT_SB;
  {  // sources per CTA: fewer when the whole batch would be only a few waves of CTAs
    long long ctas32 = 0;
    const long long ct = E.kmat_cta_targets[dotn];
    for (int64_t b = 0; b < nbatch; b++) {
      const long long nt = trg_dsp[b + 1] - trg_dsp[b], ns = src_dsp[b + 1] - src_dsp[b];
      if (nt > 0 && ns > 0) ctas32 += ((nt + ct - 1) / ct) * ((ns + KMAT_SB - 1) / KMAT_SB);
    }
    sb = kmat_sources_per_cta(ctas32, c->sms, k1out == 1);
  }
  for (int64_t b = 0; b <= nbatch; b++) {
    h_trg[b] = trg_dsp[b];
    h_src[b] = src_dsp[b];
  }
  if (trg_dsp[0] < 0 || src_dsp[0] < 0) return set_error(SCTL_B200_ERR_ARG, "negative displacement");
  for (int64_t b = 0; b < nbatch; b++) {
    const long long nt = trg_dsp[b + 1] - trg_dsp[b], ns = src_dsp[b + 1] - src_dsp[b];
    if (nt < 0 || ns < 0) return set_error(SCTL_B200_ERR_ARG, "block %lld: displacements must be non-decreasing", (long long)b);
    const long long len = ns * ki.src_dim * nt * k1out;
    if (M_dsp[b] < 0 || M_dsp[b] + len > M_len) return set_error(SCTL_B200_ERR_ARG, "block %lld does not fit in M (offset %lld + %lld > %lld)", (long long)b, (long long)M_dsp[b], len, (long long)M_len);
    h_M[b] = M_dsp[b];
    const long long ct = E.kmat_cta_targets[dotn];
    const long long ctas = (nt && ns) ? ((nt + ct - 1) / ct) * ((ns + sb - 1) / sb) : 0;
    h_cta[b + 1] = h_cta[b] + ctas;
  }
  const long long n_ctas = h_cta[nbatch];
  if (n_ctas == 0) return 0;
  const int64_t Nt_all = trg_dsp[nbatch], Ns_all = src_dsp[nbatch];
  cudaStream_t st = c->stream;
  DevAlloc d_desc, d_rt, d_rs, d_ns, d_w, d_nt, d_M;
  if ((rc = d_desc.alloc(desc.size() * sizeof(long long), st))) return rc;
  CU_TRY(cudaMemcpyAsync(d_desc.p, desc.data(), desc.size() * sizeof(long long), cudaMemcpyHostToDevice, st));
  const void *p_rt = r_trg, *p_rs = r_src, *p_ns = n_src, *p_w = src_wts, *p_nt = n_trg;
  void* p_M = M;
  if (!device_ptrs) {
    auto up = [&](DevAlloc& d, const void* h, size_t bytes, const void** out) -> int {
      if (!h || !bytes) {
        *out = nullptr;
        return 0;
      }
      int r = d.alloc(bytes, st);
      if (r) return r;
      cudaError_t e = cudaMemcpyAsync(d.p, h, bytes, cudaMemcpyHostToDevice, st);
      if (e != cudaSuccess) return cuda_error(e, "cudaMemcpyAsync");
      *out = d.p;
      return 0;
    };
    if ((rc = up(d_rt, r_trg, (size_t)Nt_all * 3 * esz, &p_rt)) || (rc = up(d_rs, r_src, (size_t)Ns_all * 3 * esz, &p_rs)) ||
        (rc = up(d_ns, ki.normal_dim ? n_src : nullptr, (size_t)Ns_all * ki.normal_dim * esz, &p_ns)) ||
        (rc = up(d_w, src_wts, (size_t)Ns_all * esz, &p_w)) || (rc = up(d_nt, n_trg, (size_t)Nt_all * 3 * esz, &p_nt)))
      return rc;
    if ((rc = d_M.alloc((size_t)M_len * esz, st))) return rc;
    // entries of M that no block covers keep the caller's values: stage the caller's buffer in first - unless the blocks tile
    // M exactly (the layout KernelMatrixBatch of kernel-matrix-batch-b200.hpp builds: consecutive blocks, no gaps), where the
    // upload of a multi-GB matrix that is about to be overwritten would only cost time
    bool tiled = true;
    long long at = 0;
    for (int64_t b = 0; b < nbatch && tiled; b++) {
      tiled = (M_dsp[b] == at);
      at += (src_dsp[b + 1] - src_dsp[b]) * ki.src_dim * (trg_dsp[b + 1] - trg_dsp[b]) * k1out;
    }
    tiled = tiled && at == M_len;
    if (!tiled) CU_TRY(cudaMemcpyAsync(d_M.p, M, (size_t)M_len * esz, cudaMemcpyHostToDevice, st));
    p_M = d_M.p;
  }
  KmatBatchDesc D{};
  D.trg_dsp = (const long long*)d_desc.p;
  D.src_dsp = D.trg_dsp + nbatch + 1;
  D.M_dsp = D.src_dsp + nbatch + 1;
  D.cta_dsp = D.M_dsp + nbatch;
  D.nbatch = nbatch;
  D.sb = sb;
  CU_TRY(E.kmat_batch[dotn][newton](p_rt, p_rs, p_ns, p_w, p_nt, p_M, D, n_ctas, kernel_scale(kernel_id), st));
  if (!device_ptrs) CU_TRY(cudaMemcpyAsync(M, p_M, (size_t)M_len * esz, cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Device-resident particle context: the GPU-aware counterpart of ParticleFMM's registry + EvalDirect
// (reference: fmm-wrapper.txx:297-479 registry, :490-562 EvalDirect). Coordinates, normals, densities and
// the packed source records stay in HBM between calls; records are re-packed only when the coordinates or
// the density of that source changed (coordinates are set once, the density changes every solver
// iteration: boundary_integral.txx:777-778 vs :1054).
// ------------------------------------------------------------------------------------------------
struct FmmSrc {
  int src_dim = 0, normal_dim = 0;
  int64_t Ns = 0;    // from the coordinates
  int64_t n_den = 0; // elements of the density last set
  DevBuf X, Xn, F, W;
  int64_t n_wts = 0;  // 0 = no quadrature weights, else must equal Ns
  unsigned long long version = 0;  // bumped by SetSrcCoord / SetSrcDensity / SetSrcWeights
  std::vector<int64_t> counts;     // multi-rank: sources held by every rank (Ns is their sum); empty until SetSrcCoord
};
struct FmmTrg {
  int trg_dim = 0;
  int64_t Nt = 0;      // targets this rank EVALUATES (its own, or its block of the re-partitioned set)
  int64_t Nt_own = 0;  // targets this rank handed in = rows eval_direct returns
  int64_t n_nrm = 0;   // 0 = no target-normal contraction, else must equal Nt*3
  DevBuf X, U, Xn, Ufull;
  // multi-rank: targets per rank; moved = re-partitioned evenly (fmm-wrapper.txx:504-510) because the ranks hold uneven numbers
  std::vector<int64_t> counts;
  bool moved = false;
  int64_t lo = 0, hi = 0;  // the evaluated block within the concatenated target set
  DevBuf Ublk, Ugath;
};
struct FmmS2T {
  int kernel_id = -1;
  DevBuf rec;
  unsigned long long packed_version = ~0ull;
};
struct FmmCtx {
  int dtype = SCTL_B200_F64;
  int dev = 0;
  int digits = -1;  // ParticleFMM default digits_ = 10 is an FMM tolerance; the direct sum is evaluated at SetAccuracy's value
  Comm* comm = nullptr;  // multi-rank registry (sctl_b200_fmm_set_comm); not owned
  std::mutex mu;
  std::vector<std::pair<std::string, FmmSrc*>> src;
  std::vector<std::pair<std::string, FmmTrg*>> trg;
  std::vector<std::pair<std::pair<std::string, std::string>, FmmS2T*>> s2t;  // kept sorted like std::map (src, trg)
  ~FmmCtx() {
    for (auto& a : src) delete a.second;
    for (auto& a : trg) delete a.second;
    for (auto& a : s2t) delete a.second;
  }
  FmmSrc* find_src(const char* n) {
    for (auto& a : src)
      if (a.first == n) return a.second;
    return nullptr;
  }
  FmmTrg* find_trg(const char* n) {
    for (auto& a : trg)
      if (a.first == n) return a.second;
    return nullptr;
  }
};

static bool fmm_multi(const FmmCtx* C) { return C->comm && comm_size(C->comm) > 1; }

// Multi-rank Set*: this rank's block (host or device pointer, counts[rank] particles of `width` bytes) -> replica of all
// ranks' blocks in `b`. `slice` (optional): keep only particles [lo, hi) of the replica (re-partitioned targets).
static int fmm_gather(FmmCtx* C, DevBuf& b, const void* src, size_t width, const std::vector<int64_t>& counts, bool device_ptrs, cudaStream_t st,
                      bool slice = false, int64_t lo = 0, int64_t hi = 0) {
  const int rank = comm_rank(C->comm);
  int64_t total = 0;
  for (int64_t v : counts) total += v;
  const size_t mine = (size_t)counts[(size_t)rank] * width;
  DevAlloc tmp, all;
  const void* d_local = src;
  int rc;
  if (!device_ptrs) {
    if ((rc = tmp.alloc(mine, st))) return rc;
    if (mine) CU_TRY(cudaMemcpyAsync(tmp.p, src, mine, cudaMemcpyHostToDevice, st));
    d_local = tmp.p;
  }
  if (!slice) {
    if ((rc = b.reserve((size_t)total * width + 16))) return rc;
    return comm_allgatherv(C->comm, d_local, width, counts, b.p, st);
  }
  if ((rc = all.alloc((size_t)total * width, st))) return rc;
  if ((rc = comm_allgatherv(C->comm, d_local, width, counts, all.p, st))) return rc;
  if ((rc = b.reserve((size_t)(hi - lo) * width + 16))) return rc;
  if (hi > lo) CU_TRY(cudaMemcpyAsync(b.p, (const char*)all.p + (size_t)lo * width, (size_t)(hi - lo) * width, cudaMemcpyDeviceToDevice, st));
  return 0;
}

static int fmm_upload(DevBuf& b, const void* src, size_t bytes, bool device_ptrs, cudaStream_t st) {
  int rc = b.reserve(bytes ? bytes : 16);
  if (rc) return rc;
  if (bytes) CU_TRY(cudaMemcpyAsync(b.p, src, bytes, device_ptrs ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
  return 0;
}

}  // namespace sctl_b200

using namespace sctl_b200;

#define FMM_ENTER(h)                                                                       \
  if (!(h)) return set_error(SCTL_B200_ERR_ARG, "null context");                           \
  FmmCtx* C = (FmmCtx*)(h);                                                                \
  std::lock_guard<std::mutex> lk__(C->mu);                                                 \
  DeviceGuard guard__; /* the caller's current device is restored on return */             \
  CU_TRY(cudaSetDevice(C->dev));                                                           \
  DevCtx* c = nullptr;                                                                     \
  {                                                                                        \
    int rc__ = get_ctx(&c);                                                                \
    if (rc__) return rc__;                                                                 \
  }

extern "C" {

int sctl_b200_fmm_create(void** fmm, int dtype) {
  if (!fmm) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if (dtype != SCTL_B200_F32 && dtype != SCTL_B200_F64) return set_error(SCTL_B200_ERR_DTYPE, "unknown dtype %d", dtype);
  DevCtx* c = nullptr;
  int rc = get_ctx(&c);
  if (rc) return rc;
  FmmCtx* C = new FmmCtx();
  C->dtype = dtype;
  C->dev = c->dev;
  *fmm = C;
  return 0;
}
int sctl_b200_fmm_destroy(void* fmm) {
  if (!fmm) return 0;
  FmmCtx* C = (FmmCtx*)fmm;
  DeviceGuard guard;
  cudaSetDevice(C->dev);
  cudaDeviceSynchronize();
  delete C;
  return 0;
}
int sctl_b200_fmm_set_comm(void* fmm, void* comm) {
  if (!fmm) return set_error(SCTL_B200_ERR_ARG, "null context");
  FmmCtx* C = (FmmCtx*)fmm;
  std::lock_guard<std::mutex> lk(C->mu);
  if (comm && comm_device((Comm*)comm) != C->dev)
    return set_error(SCTL_B200_ERR_ARG, "communicator lives on device %d, the registry on device %d", comm_device((Comm*)comm), C->dev);
  C->comm = (Comm*)comm;
  return 0;
}
int sctl_b200_fmm_set_accuracy(void* fmm, int digits) {
  if (!fmm) return set_error(SCTL_B200_ERR_ARG, "null context");
  FmmCtx* C = (FmmCtx*)fmm;
  std::lock_guard<std::mutex> lk(C->mu);
  C->digits = digits;
  return 0;
}
int sctl_b200_fmm_add_src(void* fmm, const char* name, int src_dim, int normal_dim) {
  if (!fmm || !name) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  FmmCtx* C = (FmmCtx*)fmm;
  std::lock_guard<std::mutex> lk(C->mu);
  if (C->find_src(name)) return set_error(SCTL_B200_ERR_ARG, "Source name already exists.");
  if (src_dim < 1 || src_dim > 4 || (normal_dim != 0 && normal_dim != 3)) return set_error(SCTL_B200_ERR_ARG, "bad source dims");
  FmmSrc* s = new FmmSrc();
  s->src_dim = src_dim;
  s->normal_dim = normal_dim;
  C->src.emplace_back(name, s);
  return 0;
}
int sctl_b200_fmm_add_trg(void* fmm, const char* name, int trg_dim) {
  if (!fmm || !name) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  FmmCtx* C = (FmmCtx*)fmm;
  std::lock_guard<std::mutex> lk(C->mu);
  if (C->find_trg(name)) return set_error(SCTL_B200_ERR_ARG, "Target name already exists.");
  if (trg_dim < 1 || trg_dim > 9) return set_error(SCTL_B200_ERR_ARG, "bad target dim");
  FmmTrg* t = new FmmTrg();
  t->trg_dim = trg_dim;
  C->trg.emplace_back(name, t);
  return 0;
}
int sctl_b200_fmm_set_kernel_s2t(void* fmm, const char* src_name, const char* trg_name, int kernel_id) {
  if (!fmm || !src_name || !trg_name) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if (kernel_id < 0 || kernel_id >= NUM_KERNELS) return set_error(SCTL_B200_ERR_KERNEL, "unknown kernel id %d", kernel_id);
  FmmCtx* C = (FmmCtx*)fmm;
  std::lock_guard<std::mutex> lk(C->mu);
  FmmSrc* s = C->find_src(src_name);
  FmmTrg* t = C->find_trg(trg_name);
  if (!s) return set_error(SCTL_B200_ERR_ARG, "Source name does not exists.");
  if (!t) return set_error(SCTL_B200_ERR_ARG, "Target name does not exists.");
  const KernelInfo& ki = KINFO[kernel_id];
  // CheckKernelDims (fmm-wrapper.txx:567-605): the S2T kernel must agree with the dims the source / target were declared with
  if (ki.src_dim != s->src_dim || ki.normal_dim != s->normal_dim || ki.trg_dim != t->trg_dim)
    return set_error(SCTL_B200_ERR_ARG, "kernel %s (src %d, normal %d, trg %d) does not match source '%s' (%d, %d) / target '%s' (%d)", ki.name,
                     ki.src_dim, ki.normal_dim, ki.trg_dim, src_name, s->src_dim, s->normal_dim, trg_name, t->trg_dim);
  const std::pair<std::string, std::string> key(src_name, trg_name);
  for (auto& a : C->s2t)
    if (a.first == key) {  // replaces an existing S2T kernel (fmm-wrapper.txx:376)
      a.second->kernel_id = kernel_id;
      a.second->packed_version = ~0ull;
      return 0;
    }
  FmmS2T* d = new FmmS2T();
  d->kernel_id = kernel_id;
  auto it = C->s2t.begin();
  while (it != C->s2t.end() && it->first < key) ++it;  // std::map iteration order of the reference (:514)
  C->s2t.insert(it, std::make_pair(key, d));
  return 0;
}
int sctl_b200_fmm_delete_src(void* fmm, const char* name) {
  if (!fmm || !name) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  FmmCtx* C = (FmmCtx*)fmm;
  std::lock_guard<std::mutex> lk(C->mu);
  DeviceGuard guard;
  cudaSetDevice(C->dev);
  cudaDeviceSynchronize();
  bool found = false;
  for (size_t i = 0; i < C->src.size(); i++)
    if (C->src[i].first == name) {
      delete C->src[i].second;
      C->src.erase(C->src.begin() + i);
      found = true;
      break;
    }
  if (!found) return set_error(SCTL_B200_ERR_ARG, "Source name does not exist.");
  for (size_t i = C->s2t.size(); i-- > 0;)
    if (C->s2t[i].first.first == name) {
      delete C->s2t[i].second;
      C->s2t.erase(C->s2t.begin() + i);
    }
  return 0;
}
int sctl_b200_fmm_delete_trg(void* fmm, const char* name) {
  if (!fmm || !name) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  FmmCtx* C = (FmmCtx*)fmm;
  std::lock_guard<std::mutex> lk(C->mu);
  DeviceGuard guard;
  cudaSetDevice(C->dev);
  cudaDeviceSynchronize();
  bool found = false;
  for (size_t i = 0; i < C->trg.size(); i++)
    if (C->trg[i].first == name) {
      delete C->trg[i].second;
      C->trg.erase(C->trg.begin() + i);
      found = true;
      break;
    }
  if (!found) return set_error(SCTL_B200_ERR_ARG, "Target name does not exist.");
  for (size_t i = C->s2t.size(); i-- > 0;)
    if (C->s2t[i].first.second == name) {
      delete C->s2t[i].second;
      C->s2t.erase(C->s2t.begin() + i);
    }
  return 0;
}

int sctl_b200_fmm_set_src_coord(void* fmm, const char* name, int64_t Ns, const void* r_src, const void* n_src, unsigned flags) {
  FMM_ENTER(fmm);
  FmmSrc* s = name ? C->find_src(name) : nullptr;
  if (!s) return set_error(SCTL_B200_ERR_ARG, "Source name does not exist.");
  if (Ns < 0 || (Ns > 0 && (!r_src || (s->normal_dim && !n_src)))) return set_error(SCTL_B200_ERR_ARG, "bad source coordinates");
  const size_t esz = elem_size(C->dtype);
  const bool dp = flags & SCTL_B200_DEVICE_PTRS;
  int rc;
  if (fmm_multi(C)) {  // collective: every rank hands in its block; the registry holds the replica of all of them
    if ((rc = comm_counts(C->comm, Ns, s->counts, c->stream))) return rc;
    if ((rc = fmm_gather(C, s->X, r_src, 3 * esz, s->counts, dp, c->stream))) return rc;
    if (s->normal_dim && (rc = fmm_gather(C, s->Xn, n_src, (size_t)s->normal_dim * esz, s->counts, dp, c->stream))) return rc;
    Ns = 0;
    for (int64_t v : s->counts) Ns += v;
  } else {
    s->counts.clear();
    if ((rc = fmm_upload(s->X, r_src, (size_t)Ns * 3 * esz, dp, c->stream))) return rc;
    if (s->normal_dim && (rc = fmm_upload(s->Xn, n_src, (size_t)Ns * s->normal_dim * esz, dp, c->stream))) return rc;
  }
  CU_TRY(cudaStreamSynchronize(c->stream));  // the caller may free its buffers after the call (the reference copies, fmm-wrapper.txx:447-448)
  s->Ns = Ns;
  s->version++;
  return 0;
}
int sctl_b200_fmm_set_src_density(void* fmm, const char* name, int64_t n, const void* v_src, unsigned flags) {
  FMM_ENTER(fmm);
  FmmSrc* s = name ? C->find_src(name) : nullptr;
  if (!s) return set_error(SCTL_B200_ERR_ARG, "Source name does not exist.");
  if (n < 0 || (n > 0 && !v_src)) return set_error(SCTL_B200_ERR_ARG, "bad source density");
  int rc;
  if (fmm_multi(C)) {
    if (s->counts.empty()) return set_error(SCTL_B200_ERR_ARG, "SetSrcCoord must come first: it fixes how many sources every rank holds");
    const int64_t mine = s->counts[(size_t)comm_rank(C->comm)];
    if (n != mine * s->src_dim) return set_error(SCTL_B200_ERR_ARG, "density has %lld values, expected Ns_local*SrcDim = %lld", (long long)n, (long long)(mine * s->src_dim));
    if ((rc = fmm_gather(C, s->F, v_src, (size_t)s->src_dim * elem_size(C->dtype), s->counts, flags & SCTL_B200_DEVICE_PTRS, c->stream))) return rc;
    n = s->Ns * s->src_dim;
  } else if ((rc = fmm_upload(s->F, v_src, (size_t)n * elem_size(C->dtype), flags & SCTL_B200_DEVICE_PTRS, c->stream))) {
    return rc;
  }
  CU_TRY(cudaStreamSynchronize(c->stream));
  s->n_den = n;
  s->version++;
  return 0;
}
int sctl_b200_fmm_set_trg_coord(void* fmm, const char* name, int64_t Nt, const void* r_trg, unsigned flags) {
  FMM_ENTER(fmm);
  FmmTrg* t = name ? C->find_trg(name) : nullptr;
  if (!t) return set_error(SCTL_B200_ERR_ARG, "Target name does not exist.");
  if (Nt < 0 || (Nt > 0 && !r_trg)) return set_error(SCTL_B200_ERR_ARG, "bad target coordinates");
  int rc;
  t->Nt_own = Nt;
  t->moved = false;
  if (fmm_multi(C)) {
    const int rank = comm_rank(C->comm), size = comm_size(C->comm);
    if ((rc = comm_counts(C->comm, Nt, t->counts, c->stream))) return rc;
    int64_t total = 0, mx = 0, before = 0;
    for (int r = 0; r < size; r++) {
      total += t->counts[(size_t)r];
      mx = t->counts[(size_t)r] > mx ? t->counts[(size_t)r] : mx;
      if (r < rank) before += t->counts[(size_t)r];
    }
    // keep targets with their owner unless the ranks are uneven (> 5 % above the mean): then every rank evaluates block
    // [total*r/np, total*(r+1)/np) of the concatenated set, the reference's PartitionN (fmm-wrapper.txx:504-510)
    t->moved = total > 0 && (double)mx > 1.05 * (double)total / size + 1.0;
    if (t->moved) {
      t->lo = (int64_t)((__int128)total * rank / size);
      t->hi = (int64_t)((__int128)total * (rank + 1) / size);
      if ((rc = fmm_gather(C, t->X, r_trg, 3 * elem_size(C->dtype), t->counts, flags & SCTL_B200_DEVICE_PTRS, c->stream, true, t->lo, t->hi))) return rc;
      Nt = t->hi - t->lo;
    } else {
      t->lo = before;
      t->hi = before + Nt;
    }
  } else {
    t->counts.clear();
  }
  if (!t->moved && (rc = fmm_upload(t->X, r_trg, (size_t)Nt * 3 * elem_size(C->dtype), flags & SCTL_B200_DEVICE_PTRS, c->stream))) return rc;
  CU_TRY(cudaStreamSynchronize(c->stream));
  t->Nt = Nt;
  return 0;
}
// Optional far-field quadrature weights of a source: the density is multiplied by wts[s] when the records are packed
// (BoundaryIntegralOp::ComputeFarField scales F_far by wts_far on the host before SetSrcDensity, boundary_integral.txx:1036-1051).
// n == 0 removes the weights.
int sctl_b200_fmm_set_src_weights(void* fmm, const char* name, int64_t n, const void* wts, unsigned flags) {
  FMM_ENTER(fmm);
  FmmSrc* s = name ? C->find_src(name) : nullptr;
  if (!s) return set_error(SCTL_B200_ERR_ARG, "Source name does not exist.");
  if (n < 0 || (n > 0 && !wts)) return set_error(SCTL_B200_ERR_ARG, "bad source weights");
  int rc;
  if (fmm_multi(C)) {
    // collective in multi-rank mode, also for a rank that holds no sources: the ranks first agree whether weights are being
    // set (some rank passes n > 0) or removed (every rank passes 0)
    std::vector<int64_t> wc;
    if ((rc = comm_counts(C->comm, n, wc, c->stream))) return rc;
    int64_t total = 0;
    for (int64_t v : wc) total += v;
    if (total == 0) {
      s->n_wts = 0;
      s->version++;
      return 0;
    }
    if (s->counts.empty()) return set_error(SCTL_B200_ERR_ARG, "SetSrcCoord must come first: it fixes how many sources every rank holds");
    if (n != s->counts[(size_t)comm_rank(C->comm)]) return set_error(SCTL_B200_ERR_ARG, "%lld weights, expected Ns_local = %lld", (long long)n, (long long)s->counts[(size_t)comm_rank(C->comm)]);
    if ((rc = fmm_gather(C, s->W, wts, elem_size(C->dtype), s->counts, flags & SCTL_B200_DEVICE_PTRS, c->stream))) return rc;
    n = s->Ns;
  } else if ((rc = fmm_upload(s->W, wts, (size_t)n * elem_size(C->dtype), flags & SCTL_B200_DEVICE_PTRS, c->stream))) {
    return rc;
  }
  CU_TRY(cudaStreamSynchronize(c->stream));
  s->n_wts = n;
  s->version++;
  return 0;
}
// Optional target normals: eval_direct then returns U[t*K1'+k] = sum_l Ufull[(t*K1'+k)*3+l] * n_trg[t*3+l] with K1' = TrgDim/3
// (the trg_normal_dot_prod_ branch of ComputeFarField, boundary_integral.txx:1060-1071). n == 0 removes them.
int sctl_b200_fmm_set_trg_normal(void* fmm, const char* name, int64_t n, const void* n_trg, unsigned flags) {
  FMM_ENTER(fmm);
  FmmTrg* t = name ? C->find_trg(name) : nullptr;
  if (!t) return set_error(SCTL_B200_ERR_ARG, "Target name does not exist.");
  if (n < 0 || (n > 0 && !n_trg)) return set_error(SCTL_B200_ERR_ARG, "bad target normals");
  if (n > 0 && t->trg_dim % 3) return set_error(SCTL_B200_ERR_ARG, "target-normal contraction needs TrgDim %% 3 == 0 (target '%s' has %d)", name, t->trg_dim);
  int rc;
  if (fmm_multi(C) && t->moved) {
    // collective (the targets were re-partitioned): ranks that own no target take part too; all-zero = remove the normals
    std::vector<int64_t> nc;
    if ((rc = comm_counts(C->comm, n, nc, c->stream))) return rc;
    int64_t total = 0;
    for (int64_t v : nc) total += v;
    if (total == 0) {
      t->n_nrm = 0;
      return 0;
    }
    if (n != t->Nt_own * 3) return set_error(SCTL_B200_ERR_ARG, "%lld normal values, expected Nt_local*3 = %lld", (long long)n, (long long)(t->Nt_own * 3));
    if ((rc = fmm_gather(C, t->Xn, n_trg, 3 * elem_size(C->dtype), t->counts, flags & SCTL_B200_DEVICE_PTRS, c->stream, true, t->lo, t->hi))) return rc;
    n = t->Nt * 3;
  } else if ((rc = fmm_upload(t->Xn, n_trg, (size_t)n * elem_size(C->dtype), flags & SCTL_B200_DEVICE_PTRS, c->stream))) {
    return rc;
  }
  CU_TRY(cudaStreamSynchronize(c->stream));
  t->n_nrm = n;
  return 0;
}
int sctl_b200_fmm_trg_count(void* fmm, const char* name, int64_t* Nt, int* trg_dim) {
  if (!fmm || !name) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  FmmCtx* C = (FmmCtx*)fmm;
  std::lock_guard<std::mutex> lk(C->mu);
  FmmTrg* t = C->find_trg(name);
  if (!t) return set_error(SCTL_B200_ERR_ARG, "Target name does not exist.");
  if (Nt) *Nt = t->Nt_own;
  if (trg_dim) *trg_dim = t->n_nrm ? t->trg_dim / 3 : t->trg_dim;  // values per target that eval_direct returns
  return 0;
}

// U[Nt*TrgDim] = sum over every source registered with an S2T kernel for this target (overwritten, fmm-wrapper.txx:501-502)
int sctl_b200_fmm_eval_direct(void* fmm, const char* trg_name, void* U, int64_t n, unsigned flags) {
  FMM_ENTER(fmm);
  FmmTrg* t = trg_name ? C->find_trg(trg_name) : nullptr;
  if (!t) return set_error(SCTL_B200_ERR_ARG, "Target name does not exist.");
  const bool dotn = t->n_nrm != 0;
  if (dotn && t->n_nrm != t->Nt * 3) return set_error(SCTL_B200_ERR_ARG, "target '%s': %lld normal values, expected Nt*3 = %lld", trg_name, (long long)t->n_nrm, (long long)(t->Nt * 3));
  const int k1out = dotn ? t->trg_dim / 3 : t->trg_dim;
  if (n != t->Nt_own * k1out) return set_error(SCTL_B200_ERR_ARG, "U must hold Nt*%d = %lld values", k1out, (long long)(t->Nt_own * k1out));
  const bool moved = fmm_multi(C) && t->moved;  // this rank evaluates a block of the re-partitioned targets; the potentials travel back below
  if (n == 0 && !moved) return 0;
  if (n > 0 && !U) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  const size_t esz = elem_size(C->dtype);
  const bool dp = flags & SCTL_B200_DEVICE_PTRS;
  cudaStream_t st = c->stream;
  void* dU = U;  // where the result of the EVALUATED targets is produced on the device
  int rc;
  if (moved) {
    if ((rc = t->Ublk.reserve((size_t)t->Nt * k1out * esz + 16))) return rc;
    dU = t->Ublk.p;
  } else if (!dp) {
    if ((rc = t->U.reserve((size_t)n * esz))) return rc;
    dU = t->U.p;
  }
  void* dSum = dU;  // where the kernels accumulate (full TrgDim values per target)
  const int64_t n_full = t->Nt * t->trg_dim;
  if (dotn) {
    if ((rc = t->Ufull.reserve((size_t)n_full * esz))) return rc;
    dSum = t->Ufull.p;
  }
  bool first = true;
  for (auto& a : C->s2t) {
    if (a.first.second != trg_name) continue;
    FmmSrc* s = C->find_src(a.first.first.c_str());
    FmmS2T* d = a.second;
    if (!s) return set_error(SCTL_B200_ERR_ARG, "Source name does not exist.");
    if (s->n_den != s->Ns * s->src_dim)  // the reference asserts F.Dim() == Ns*SrcDim (fmm-wrapper.txx:533)
      return set_error(SCTL_B200_ERR_ARG, "source '%s': density has %lld values, expected Ns*SrcDim = %lld", a.first.first.c_str(),
                       (long long)s->n_den, (long long)(s->Ns * s->src_dim));
    if (s->n_wts && s->n_wts != s->Ns)
      return set_error(SCTL_B200_ERR_ARG, "source '%s': %lld weights, expected Ns = %lld", a.first.first.c_str(), (long long)s->n_wts, (long long)s->Ns);
    if (s->Ns == 0 || t->Nt == 0) continue;
    const EvalEntry& E = (C->dtype == SCTL_B200_F64 ? eval_table_f64() : eval_table_f32())[d->kernel_id];
    if (use_small_path(t->Nt, s->Ns)) {  // one launch straight from the resident AoS arrays; weights fused, no records needed
      CU_TRY(E.small[newton_level(C->dtype, C->digits)](t->X.p, s->X.p, s->normal_dim ? s->Xn.p : nullptr, s->F.p, s->n_wts ? s->W.p : nullptr, dSum, t->Nt,
                                                        s->Ns, kernel_scale(d->kernel_id), first ? 0 : 1, c->sms, st));
      first = false;
      continue;
    }
    if (d->packed_version != s->version) {
      if ((rc = d->rec.reserve((size_t)padded_sources(s->Ns) * E.rec_len * esz))) return rc;
      if ((rc = pack_device(d->kernel_id, C->dtype, s->Ns, s->X.p, s->normal_dim ? s->Xn.p : nullptr, s->F.p, s->n_wts ? s->W.p : nullptr, d->rec.p, st)))
        return rc;
      d->packed_version = s->version;
    }
    if ((rc = eval_packed_device(d->kernel_id, C->dtype, t->Nt, s->Ns, t->X.p, d->rec.p, dSum, C->digits, !first, st, c->sms, nullptr))) return rc;
    first = false;
  }
  if (first && n_full) CU_TRY(cudaMemsetAsync(dSum, 0, (size_t)n_full * esz, st));
  if (dotn) CU_TRY((C->dtype == SCTL_B200_F64 ? contract_normal_f64 : contract_normal_f32)(dSum, t->Xn.p, dU, t->Nt, k1out, st));
  if (moved) {
    // potentials back to the ranks that own the targets (fmm-wrapper.txx:560-561): all-gather the evenly sized blocks, keep own rows
    const int size = comm_size(C->comm);
    int64_t total = 0;
    for (int64_t v : t->counts) total += v;
    std::vector<int64_t> blk((size_t)size);
    for (int r = 0; r < size; r++) blk[(size_t)r] = (int64_t)((__int128)total * (r + 1) / size) - (int64_t)((__int128)total * r / size);
    if ((rc = t->Ugath.reserve((size_t)total * k1out * esz + 16))) return rc;
    if ((rc = comm_allgatherv(C->comm, dU, (size_t)k1out * esz, blk, t->Ugath.p, st))) return rc;
    int64_t own0 = 0;
    for (int r = 0; r < comm_rank(C->comm); r++) own0 += t->counts[(size_t)r];
    if (n) CU_TRY(cudaMemcpyAsync(U, (const char*)t->Ugath.p + (size_t)own0 * k1out * esz, (size_t)n * esz, dp ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
  } else if (!dp) {
    CU_TRY(cudaMemcpyAsync(U, dU, (size_t)n * esz, cudaMemcpyDeviceToHost, st));
  }
  CU_TRY(cudaStreamSynchronize(st));
  return 0;
}

}  // extern "C"


extern "C" {

int sctl_b200_kernel_dims(int kernel_id, int* coord_dim, int* normal_dim, int* src_dim, int* trg_dim, double* scale, int* flops) {
  if (kernel_id < 0 || kernel_id >= NUM_KERNELS) return set_error(SCTL_B200_ERR_KERNEL, "unknown kernel id %d", kernel_id);
  const KernelInfo& ki = KINFO[kernel_id];
  if (coord_dim) *coord_dim = 3;
  if (normal_dim) *normal_dim = ki.normal_dim;
  if (src_dim) *src_dim = ki.src_dim;
  if (trg_dim) *trg_dim = ki.trg_dim;
  if (scale) *scale = kernel_scale(kernel_id);
  if (flops) *flops = ki.flops;
  return 0;
}

const char* sctl_b200_kernel_name(int kernel_id) {
  if (kernel_id < 0 || kernel_id >= NUM_KERNELS) return nullptr;
  return KINFO[kernel_id].name;
}

int sctl_b200_eval(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                   const void* v_src, void* v_trg, int digits, unsigned flags) {
  int rc = check_common(kernel_id, dtype, Nt, Ns);
  if (rc) return rc;
  const KernelInfo& ki = KINFO[kernel_id];
  if ((Nt > 0 && (!r_trg || !v_trg)) || (Ns > 0 && Nt > 0 && (!r_src || !v_src || (ki.normal_dim && !n_src))))
    return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if ((flags & SCTL_B200_DEVICE_PTRS) && (flags & SCTL_B200_ALL_GPUS))
    return set_error(SCTL_B200_ERR_ARG, "SCTL_B200_ALL_GPUS needs host pointers");
  const bool accumulate = !(flags & SCTL_B200_OVERWRITE);
  if (flags & SCTL_B200_DEVICE_PTRS) {
    DevCtx* c = nullptr;
    if ((rc = get_ctx(&c))) return rc;
    rc = eval_device(kernel_id, dtype, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg, digits, accumulate, c->stream, c->sms, nullptr);
    if (rc) return rc;
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
  }
  if (flags & SCTL_B200_ALL_GPUS) return eval_host_all_gpus(kernel_id, dtype, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg, digits, accumulate);
  return eval_host_one_device(kernel_id, dtype, Nt, Ns, r_trg, r_src, n_src, v_src, v_trg, digits, accumulate);
}

int sctl_b200_eval_async(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* d_r_trg, const void* d_r_src, const void* d_n_src,
                         const void* d_v_src, void* d_v_trg, int digits, unsigned flags, void* cuda_stream) {
  int rc = check_common(kernel_id, dtype, Nt, Ns);
  if (rc) return rc;
  const KernelInfo& ki = KINFO[kernel_id];
  if ((Nt > 0 && (!d_r_trg || !d_v_trg)) || (Ns > 0 && Nt > 0 && (!d_r_src || !d_v_src || (ki.normal_dim && !d_n_src))))
    return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if (flags & SCTL_B200_ALL_GPUS) return set_error(SCTL_B200_ERR_ARG, "SCTL_B200_ALL_GPUS needs host pointers");
  DevCtx* c = nullptr;
  if ((rc = get_ctx(&c))) return rc;
  return eval_device(kernel_id, dtype, Nt, Ns, d_r_trg, d_r_src, d_n_src, d_v_src, d_v_trg, digits, !(flags & SCTL_B200_OVERWRITE),
                     (cudaStream_t)cuda_stream, c->sms, nullptr);
}

int sctl_b200_kernel_matrix(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                            void* M, int digits, unsigned flags) {
  int rc = check_common(kernel_id, dtype, Nt, Ns);
  if (rc) return rc;
  const KernelInfo& ki = KINFO[kernel_id];
  if (Nt > 0 && Ns > 0 && (!r_trg || !r_src || !M || (ki.normal_dim && !n_src))) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if (flags & SCTL_B200_ALL_GPUS) return set_error(SCTL_B200_ERR_ARG, "SCTL_B200_ALL_GPUS is not supported for kernel_matrix");
  if (flags & SCTL_B200_DEVICE_PTRS) {
    DevCtx* c = nullptr;
    if ((rc = get_ctx(&c))) return rc;
    if ((rc = kmat_device(kernel_id, dtype, Nt, Ns, r_trg, r_src, n_src, M, digits, c->stream, c->sms))) return rc;
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
  }
  return kmat_host(kernel_id, dtype, Nt, Ns, r_trg, r_src, n_src, M, digits);
}

int sctl_b200_kernel_matrix_async(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* d_r_trg, const void* d_r_src,
                                  const void* d_n_src, void* d_M, int digits, unsigned flags, void* cuda_stream) {
  int rc = check_common(kernel_id, dtype, Nt, Ns);
  if (rc) return rc;
  const KernelInfo& ki = KINFO[kernel_id];
  if (Nt > 0 && Ns > 0 && (!d_r_trg || !d_r_src || !d_M || (ki.normal_dim && !d_n_src))) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if (flags & SCTL_B200_ALL_GPUS) return set_error(SCTL_B200_ERR_ARG, "SCTL_B200_ALL_GPUS is not supported for kernel_matrix");
  DevCtx* c = nullptr;
  if ((rc = get_ctx(&c))) return rc;
  return kmat_device(kernel_id, dtype, Nt, Ns, d_r_trg, d_r_src, d_n_src, d_M, digits, (cudaStream_t)cuda_stream, c->sms);
}

int sctl_b200_kernel_matrix_batch(int kernel_id, int dtype, int64_t nbatch, const int64_t* trg_dsp, const int64_t* src_dsp, const int64_t* M_dsp,
                                  const void* r_trg, const void* r_src, const void* n_src, const void* src_wts, const void* n_trg, void* M,
                                  int64_t M_len, int digits, unsigned flags) {
  int rc = check_common(kernel_id, dtype, 0, 0);
  if (rc) return rc;
  if (nbatch < 0 || M_len < 0) return set_error(SCTL_B200_ERR_ARG, "negative size");
  if (nbatch == 0) return 0;
  if (!trg_dsp || !src_dsp || !M_dsp) return set_error(SCTL_B200_ERR_ARG, "null block description");
  const KernelInfo& ki = KINFO[kernel_id];
  if (trg_dsp[nbatch] > 0 && src_dsp[nbatch] > 0 && (!r_trg || !r_src || !M || (ki.normal_dim && !n_src)))
    return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if (flags & SCTL_B200_ALL_GPUS) return set_error(SCTL_B200_ERR_ARG, "SCTL_B200_ALL_GPUS is not supported for kernel_matrix_batch");
  return kmat_batch_impl(kernel_id, dtype, nbatch, trg_dsp, src_dsp, M_dsp, r_trg, r_src, n_src, src_wts, n_trg, M, M_len, digits,
                         (flags & SCTL_B200_DEVICE_PTRS) != 0);
}

int sctl_b200_buf_create(void** d_ptr, int dtype, int64_t n) {
  if (!d_ptr || n < 0) return set_error(SCTL_B200_ERR_ARG, "bad argument");
  if (dtype != SCTL_B200_F32 && dtype != SCTL_B200_F64) return set_error(SCTL_B200_ERR_DTYPE, "unknown dtype %d", dtype);
  DevCtx* c = nullptr;
  int rc = get_ctx(&c);
  if (rc) return rc;
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, n ? (size_t)n * elem_size(dtype) : 16);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return set_error(SCTL_B200_ERR_NOMEM, "cudaMalloc: %s", cudaGetErrorString(e));
  }
  *d_ptr = p;
  return 0;
}
int sctl_b200_buf_upload(void* d_ptr, const void* host, int dtype, int64_t n) {
  if ((n > 0 && (!d_ptr || !host)) || n < 0) return set_error(SCTL_B200_ERR_ARG, "bad argument");
  if (dtype != SCTL_B200_F32 && dtype != SCTL_B200_F64) return set_error(SCTL_B200_ERR_DTYPE, "unknown dtype %d", dtype);
  if (n) CU_TRY(cudaMemcpy(d_ptr, host, (size_t)n * elem_size(dtype), cudaMemcpyHostToDevice));
  return 0;
}
int sctl_b200_buf_download(const void* d_ptr, void* host, int dtype, int64_t n) {
  if ((n > 0 && (!d_ptr || !host)) || n < 0) return set_error(SCTL_B200_ERR_ARG, "bad argument");
  if (dtype != SCTL_B200_F32 && dtype != SCTL_B200_F64) return set_error(SCTL_B200_ERR_DTYPE, "unknown dtype %d", dtype);
  if (n) CU_TRY(cudaMemcpy(host, d_ptr, (size_t)n * elem_size(dtype), cudaMemcpyDeviceToHost));
  return 0;
}
int sctl_b200_buf_zero(void* d_ptr, int dtype, int64_t n) {
  if ((n > 0 && !d_ptr) || n < 0) return set_error(SCTL_B200_ERR_ARG, "bad argument");
  if (dtype != SCTL_B200_F32 && dtype != SCTL_B200_F64) return set_error(SCTL_B200_ERR_DTYPE, "unknown dtype %d", dtype);
  if (n) CU_TRY(cudaMemset(d_ptr, 0, (size_t)n * elem_size(dtype)));
  return 0;
}
int sctl_b200_buf_destroy(void* d_ptr) {
  if (d_ptr) CU_TRY(cudaFree(d_ptr));
  return 0;
}

int sctl_b200_device_count(int* n) {
  if (!n) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  int c = 0;
  cudaError_t e = cudaGetDeviceCount(&c);
  if (e != cudaSuccess) {
    cudaGetLastError();
    *n = 0;
    return cuda_error(e, "cudaGetDeviceCount");
  }
  *n = c;
  return 0;
}
int sctl_b200_init(int ndev, const int* dev_ids) {
  int visible = 0;
  cudaError_t e = cudaGetDeviceCount(&visible);
  if (e != cudaSuccess || visible < 1) {
    cudaGetLastError();
    return set_error(SCTL_B200_ERR_CUDA, "no usable CUDA device (%s); libsctl_b200 has no CPU fallback",
                     e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  if (ndev < 0) return set_error(SCTL_B200_ERR_ARG, "negative device count");
  std::vector<int> devs;
  if (ndev == 0 || !dev_ids) {
    for (int g = 0; g < visible; g++) devs.push_back(g);
  } else {
    for (int i = 0; i < ndev; i++) {
      if (dev_ids[i] < 0 || dev_ids[i] >= visible) return set_error(SCTL_B200_ERR_ARG, "device id %d out of range (0..%d)", dev_ids[i], visible - 1);
      for (int d : devs)
        if (d == dev_ids[i]) return set_error(SCTL_B200_ERR_ARG, "device id %d listed twice", dev_ids[i]);
      devs.push_back(dev_ids[i]);
    }
  }
  int dev0 = 0;
  CU_TRY(cudaGetDevice(&dev0));
  for (int d : devs) {  // create the contexts now
    CU_TRY(cudaSetDevice(d));
    CU_TRY(cudaFree(nullptr));
  }
  CU_TRY(cudaSetDevice(dev0));
  std::lock_guard<std::mutex> lk(g_devices_mutex);
  g_devices = (ndev == 0 || !dev_ids) ? std::vector<int>() : devs;
  return 0;
}
int sctl_b200_finalize(void) {
  comm_all_gpus_release();
  std::vector<int> devs;
  {
    std::lock_guard<std::mutex> lk(g_devices_mutex);
    devs.swap(g_devices);
  }
  int visible = 0;
  if (cudaGetDeviceCount(&visible) != cudaSuccess || visible < 1) {
    cudaGetLastError();
    return 0;  // nothing was ever initialised
  }
  int dev0 = 0;
  CU_TRY(cudaGetDevice(&dev0));
  if (devs.empty())
    for (auto& c : t_ctx) devs.push_back(c.dev);
  for (int d : devs) {
    if (cudaSetDevice(d) == cudaSuccess) cudaDeviceSynchronize();
  }
  release_thread_ctx();
  cudaGetLastError();
  CU_TRY(cudaSetDevice(dev0));
  return 0;
}
int sctl_b200_set_device(int dev) {
  CU_TRY(cudaSetDevice(dev));
  return 0;
}
int sctl_b200_get_device(int* dev) {
  if (!dev) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  CU_TRY(cudaGetDevice(dev));
  return 0;
}
int sctl_b200_sm_count(int* n) {
  if (!n) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  DevCtx* c = nullptr;
  int rc = get_ctx(&c);
  if (rc) return rc;
  *n = c->sms;
  return 0;
}

int64_t sctl_b200_launch_count(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }

int sctl_b200_measure_peak(int dtype, double seconds, double* tflops) {
  if (!tflops) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if (dtype != SCTL_B200_F32 && dtype != SCTL_B200_F64) return set_error(SCTL_B200_ERR_DTYPE, "unknown dtype %d", dtype);
  DevCtx* c = nullptr;
  int rc = get_ctx(&c);
  if (rc) return rc;
  CU_TRY(dtype == SCTL_B200_F64 ? run_peak_f64(seconds, tflops, c->stream) : run_peak_f32(seconds, tflops, c->stream));
  return 0;
}

int sctl_b200_last_timing(double ms[5]) {
  if (!ms) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  for (int i = 0; i < 5; i++) ms[i] = t_timing[i];
  return 0;
}

int sctl_b200_newton_level(int dtype, int digits) {
  if (dtype != SCTL_B200_F32 && dtype != SCTL_B200_F64) return -1;
  return newton_level(dtype, digits);
}

const char* sctl_b200_last_error(void) { return t_last_error.c_str(); }
const char* sctl_b200_version(void) { return "sctl_b200 0.1 (sm_100a)"; }

}  // extern "C"
