// NCCL target-sharding driver behind the C ABI (sctl_b200_comm_* / sctl_b200_eval_sharded*, include/sctl_b200.h).
//
// This is the per-rank body of the reference's multi-rank ParticleFMM::EvalDirect
// (/root/reference/include/sctl/fmm-wrapper.txx:490-562) for ranks that each own a GPU:
//
//   reference (MPI, host memory)                               here (NCCL over NVLink / NVSwitch, device memory)
//   ---------------------------------------------------------  -----------------------------------------------------------
//   partition(): rank r owns block [N r/np, N (r+1)/np) of     the caller hands every rank ITS block of targets and sources
//   the targets and of every source array (:504-529)           (sctl_b200_partition gives the same formula)
//   np ring rounds of blocking Isend/Irecv of (Xs, Xn, F),     ONE ncclAllGather per source array (a group of in-place
//   an Eval accumulating into U after each (:537-558)          ncclBroadcasts when the blocks are ragged) into a contiguous
//                                                              replica, then ONE pack + eval (+ reduce) over all sources
//   U re-partitioned back to the owners (:560-561)             targets never move: results stay with their owner
//
// NCCL is loaded with dlopen at the first communicator, not linked: a Python process that has imported torch already holds
// torch's own libnccl.so.2 and dlopen by soname returns THAT copy, while a C++ program gets the system library; a link-time
// dependency would instead pull the system NCCL into a process where torch later expects its newer one. Processes that never
// create a communicator never load NCCL. SCTL_B200_NCCL_LIB names another library file.
#include <dlfcn.h>
#include <nccl.h>  // types and prototypes only; every call goes through the table below

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "internal.h"

namespace sctl_b200 {

// ------------------------------------------------------------------------------------------------------------------
// the NCCL entry points this driver uses
// ------------------------------------------------------------------------------------------------------------------
struct NcclApi {
  void* handle = nullptr;
  std::string path;
  decltype(&ncclGetVersion) GetVersion = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommInitAll) CommInitAll = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclCommAbort) CommAbort = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclBroadcast) Broadcast = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
};
static std::mutex g_nccl_mutex;
static NcclApi g_nccl;
static std::string g_nccl_why;  // why loading failed

static const NcclApi* nccl_api() {
  std::lock_guard<std::mutex> lk(g_nccl_mutex);
  if (g_nccl.handle) return &g_nccl;
  if (!g_nccl_why.empty()) return nullptr;
  const char* forced = getenv("SCTL_B200_NCCL_LIB");
  const char* names[] = {forced && *forced ? forced : "libnccl.so.2", "libnccl.so", nullptr};
  void* h = nullptr;
  for (int i = 0; names[i] && !h; i++) {
    h = dlopen(names[i], RTLD_NOW | RTLD_LOCAL);
    if (h) g_nccl.path = names[i];
  }
  if (!h) {
    const char* why = dlerror();
    g_nccl_why = std::string("cannot load NCCL (libnccl.so.2): ") + (why ? why : "dlopen failed");
    return nullptr;
  }
#define SCTL_B200_NCCL_SYM(field, sym)                                   \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, sym)); \
  if (!g_nccl.field) {                                                   \
    g_nccl_why = std::string("NCCL library lacks ") + sym;               \
    dlclose(h);                                                          \
    return nullptr;                                                      \
  }
  SCTL_B200_NCCL_SYM(GetVersion, "ncclGetVersion")
  SCTL_B200_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
  SCTL_B200_NCCL_SYM(CommInitRank, "ncclCommInitRank")
  SCTL_B200_NCCL_SYM(CommInitAll, "ncclCommInitAll")
  SCTL_B200_NCCL_SYM(CommDestroy, "ncclCommDestroy")
  SCTL_B200_NCCL_SYM(CommAbort, "ncclCommAbort")
  SCTL_B200_NCCL_SYM(GetErrorString, "ncclGetErrorString")
  SCTL_B200_NCCL_SYM(AllGather, "ncclAllGather")
  SCTL_B200_NCCL_SYM(Broadcast, "ncclBroadcast")
  SCTL_B200_NCCL_SYM(GroupStart, "ncclGroupStart")
  SCTL_B200_NCCL_SYM(GroupEnd, "ncclGroupEnd")
#undef SCTL_B200_NCCL_SYM
  g_nccl.handle = h;
  return &g_nccl;
}

static int nccl_error(const NcclApi* api, ncclResult_t r, const char* what) {
  return set_error(SCTL_B200_ERR_NCCL, "%s: %s", what, api ? api->GetErrorString(r) : "NCCL not loaded");
}
#define NCCL_TRY(api, call)                                          \
  do {                                                               \
    ncclResult_t r__ = (call);                                       \
    if (r__ != ncclSuccess) return nccl_error((api), r__, #call);    \
  } while (0)

// ------------------------------------------------------------------------------------------------------------------
// communicator handle
// ------------------------------------------------------------------------------------------------------------------
struct Comm {
  ncclComm_t nccl = nullptr;
  int rank = 0, size = 1, dev = 0;
  DevBuf gx, gn, gf;  // replicas of the source arrays (grow-only, reused by every call)
  DevBuf cnt;         // [size + 1] int64 for the counts exchange
};
int comm_rank(const Comm* cm) { return cm->rank; }
int comm_size(const Comm* cm) { return cm->size; }
int comm_device(const Comm* cm) { return cm->dev; }

static int comm_check_device(const Comm* cm) {
  int dev = -1;
  CU_TRY(cudaGetDevice(&dev));
  if (dev != cm->dev)
    return set_error(SCTL_B200_ERR_ARG, "communicator was created on device %d but the current device is %d (call sctl_b200_set_device first)", cm->dev, dev);
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// collectives used by eval_sharded and the multi-rank particle registry
// ------------------------------------------------------------------------------------------------------------------
int comm_counts(Comm* cm, int64_t n_local, std::vector<int64_t>& counts, cudaStream_t st) {
  counts.assign((size_t)cm->size, 0);
  if (cm->size == 1) {
    counts[0] = n_local;
    return 0;
  }
  const NcclApi* api = nccl_api();
  if (!api) return set_error(SCTL_B200_ERR_NCCL, "%s", g_nccl_why.c_str());
  int rc = cm->cnt.reserve(sizeof(int64_t) * (size_t)(cm->size + 1));
  if (rc) return rc;
  int64_t* d = (int64_t*)cm->cnt.p;
  CU_TRY(cudaMemcpyAsync(d + cm->size, &n_local, sizeof(int64_t), cudaMemcpyHostToDevice, st));
  NCCL_TRY(api, api->AllGather(d + cm->size, d, sizeof(int64_t), ncclChar, cm->nccl, st));
  CU_TRY(cudaMemcpyAsync(counts.data(), d, sizeof(int64_t) * (size_t)cm->size, cudaMemcpyDeviceToHost, st));
  CU_TRY(cudaStreamSynchronize(st));
  for (int r = 0; r < cm->size; r++)
    if (counts[(size_t)r] < 0) return set_error(SCTL_B200_ERR_ARG, "rank %d reported a negative particle count", r);
  return 0;
}

int comm_allgatherv(Comm* cm, const void* d_local, size_t width_bytes, const std::vector<int64_t>& counts, void* d_out, cudaStream_t st) {
  if ((int)counts.size() != cm->size) return set_error(SCTL_B200_ERR_ARG, "counts must hold one entry per rank");
  std::vector<size_t> off((size_t)cm->size + 1, 0);
  bool equal = true;
  for (int r = 0; r < cm->size; r++) {
    off[(size_t)r + 1] = off[(size_t)r] + (size_t)counts[(size_t)r] * width_bytes;
    equal &= counts[(size_t)r] == counts[0];
  }
  const size_t mine = (size_t)counts[(size_t)cm->rank] * width_bytes;
  if (off[(size_t)cm->size] == 0 || width_bytes == 0) return 0;
  if (cm->size == 1) {
    if (d_out != d_local) CU_TRY(cudaMemcpyAsync(d_out, d_local, mine, cudaMemcpyDeviceToDevice, st));
    return 0;
  }
  const NcclApi* api = nccl_api();
  if (!api) return set_error(SCTL_B200_ERR_NCCL, "%s", g_nccl_why.c_str());
  if (equal) {
    NCCL_TRY(api, api->AllGather(d_local, d_out, mine, ncclChar, cm->nccl, st));
    return 0;
  }
  // ragged blocks: every rank broadcasts its block into place; one group = one fused NCCL launch
  char* out = (char*)d_out;
  NCCL_TRY(api, api->GroupStart());
  for (int r = 0; r < cm->size; r++) {
    const size_t n = off[(size_t)r + 1] - off[(size_t)r];
    if (!n) continue;
    const void* send = (r == cm->rank) ? d_local : (const void*)(out + off[(size_t)r]);
    ncclResult_t e = api->Broadcast(send, out + off[(size_t)r], n, ncclChar, r, cm->nccl, st);
    if (e != ncclSuccess) {
      api->GroupEnd();
      return nccl_error(api, e, "ncclBroadcast");
    }
  }
  NCCL_TRY(api, api->GroupEnd());
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// eval_sharded: this rank's targets x ALL ranks' sources
// ------------------------------------------------------------------------------------------------------------------
// flags: SCTL_B200_DEVICE_PTRS (pointers are device pointers of cm's device) | SCTL_B200_OVERWRITE. src_counts: sources per
// rank (nullptr = exchange them with one tiny all-gather + synchronisation). st: stream to enqueue on; own_stream = true
// means "st is the library stream: synchronise before returning" (host pointers always synchronise).
int eval_sharded_impl(Comm* cm, int kernel_id, int dtype, int64_t Nt, int64_t Ns_local, const void* r_trg, const void* r_src, const void* n_src,
                      const void* v_src, void* v_trg, int digits, unsigned flags, const int64_t* src_counts, cudaStream_t st, bool own_stream) {
  int rc = check_common(kernel_id, dtype, Nt, Ns_local);
  if (rc) return rc;
  if ((rc = comm_check_device(cm))) return rc;
  const KernelInfo& ki = kernel_info(kernel_id);
  if ((Nt > 0 && (!r_trg || !v_trg)) || (Ns_local > 0 && (!r_src || !v_src || (ki.normal_dim && !n_src))))
    return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if (flags & SCTL_B200_ALL_GPUS) return set_error(SCTL_B200_ERR_ARG, "SCTL_B200_ALL_GPUS does not combine with a communicator");
  DevCtx* c = nullptr;
  if ((rc = get_ctx(&c))) return rc;
  const bool dev_ptrs = (flags & SCTL_B200_DEVICE_PTRS) != 0;
  const bool accumulate = !(flags & SCTL_B200_OVERWRITE);
  const size_t esz = elem_size(dtype);

  std::vector<int64_t> counts;
  if (src_counts) {
    counts.assign(src_counts, src_counts + cm->size);
    if (counts[(size_t)cm->rank] != Ns_local)
      return set_error(SCTL_B200_ERR_ARG, "src_counts[%d] = %lld but this rank passed Ns_local = %lld", cm->rank, (long long)counts[(size_t)cm->rank], (long long)Ns_local);
    for (int64_t v : counts)
      if (v < 0) return set_error(SCTL_B200_ERR_ARG, "negative entry in src_counts");
  } else if ((rc = comm_counts(cm, Ns_local, counts, st))) {
    return rc;
  }
  int64_t Ns = 0;
  for (int64_t v : counts) Ns += v;

  // host pointers: this rank's blocks cross the bus ONCE (1/size of the sources per GPU; the rest arrives over NVLink)
  DevAlloc d_rt, d_rs, d_ns, d_vs, d_vt;
  const size_t b_rt = (size_t)Nt * 3 * esz, b_rs = (size_t)Ns_local * 3 * esz, b_ns = (size_t)Ns_local * ki.normal_dim * esz,
               b_vs = (size_t)Ns_local * ki.src_dim * esz, b_vt = (size_t)Nt * ki.trg_dim * esz;
  if (!dev_ptrs) {
    if ((rc = d_rt.alloc(b_rt, st)) || (rc = d_rs.alloc(b_rs, st)) || (rc = d_ns.alloc(b_ns, st)) || (rc = d_vs.alloc(b_vs, st)) || (rc = d_vt.alloc(b_vt, st)))
      return rc;
    if (b_rt) CU_TRY(cudaMemcpyAsync(d_rt.p, r_trg, b_rt, cudaMemcpyHostToDevice, st));
    if (b_rs) CU_TRY(cudaMemcpyAsync(d_rs.p, r_src, b_rs, cudaMemcpyHostToDevice, st));
    if (b_ns) CU_TRY(cudaMemcpyAsync(d_ns.p, n_src, b_ns, cudaMemcpyHostToDevice, st));
    if (b_vs) CU_TRY(cudaMemcpyAsync(d_vs.p, v_src, b_vs, cudaMemcpyHostToDevice, st));
    if (accumulate && b_vt) CU_TRY(cudaMemcpyAsync(d_vt.p, v_trg, b_vt, cudaMemcpyHostToDevice, st));
  }
  const void* p_rt = dev_ptrs ? r_trg : d_rt.p;
  const void* p_rs = dev_ptrs ? r_src : d_rs.p;
  const void* p_ns = dev_ptrs ? n_src : d_ns.p;
  const void* p_vs = dev_ptrs ? v_src : d_vs.p;
  void* p_vt = dev_ptrs ? v_trg : d_vt.p;

  const void *gx = p_rs, *gn = p_ns, *gf = p_vs;
  if (cm->size > 1) {
    // replicate: one collective per source array into grow-only buffers owned by the communicator
    if ((rc = cm->gx.reserve((size_t)Ns * 3 * esz + 16)) || (rc = cm->gf.reserve((size_t)Ns * ki.src_dim * esz + 16)) ||
        (ki.normal_dim && (rc = cm->gn.reserve((size_t)Ns * ki.normal_dim * esz + 16))))
      return rc;
    if ((rc = comm_allgatherv(cm, p_rs, 3 * esz, counts, cm->gx.p, st))) return rc;
    if (ki.normal_dim && (rc = comm_allgatherv(cm, p_ns, (size_t)ki.normal_dim * esz, counts, cm->gn.p, st))) return rc;
    if ((rc = comm_allgatherv(cm, p_vs, (size_t)ki.src_dim * esz, counts, cm->gf.p, st))) return rc;
    gx = cm->gx.p;
    gn = ki.normal_dim ? cm->gn.p : nullptr;
    gf = cm->gf.p;
  }
  // ONE Eval over all sources: targets are disjoint between ranks, so there is no reduction step and U stays sharded
  if ((rc = eval_device(kernel_id, dtype, Nt, Ns, p_rt, gx, ki.normal_dim ? gn : nullptr, gf, p_vt, digits, accumulate, st, c->sms, nullptr))) return rc;
  if (!dev_ptrs) {
    if (b_vt) CU_TRY(cudaMemcpyAsync(v_trg, d_vt.p, b_vt, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
  } else if (own_stream) {
    CU_TRY(cudaStreamSynchronize(st));
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// the all-GPU communicators of one process (SCTL_B200_ALL_GPUS): ncclCommInitAll over the selected devices
// ------------------------------------------------------------------------------------------------------------------
static std::mutex g_all_mutex;
static std::vector<int> g_all_devs;
static std::vector<Comm*> g_all_comms;

static void comm_free(Comm* cm) {
  if (!cm) return;
  DeviceGuard guard;
  cudaSetDevice(cm->dev);
  const NcclApi* api = nccl_api();
  if (api && cm->nccl) api->CommDestroy(cm->nccl);
  delete cm;  // DevBuf destructors free on cm->dev (current)
}

void comm_all_gpus_release() {
  std::lock_guard<std::mutex> lk(g_all_mutex);
  for (Comm* cm : g_all_comms) comm_free(cm);
  g_all_comms.clear();
  g_all_devs.clear();
}

int comm_all_gpus(const std::vector<int>& devs, std::vector<Comm*>& out) {
  std::lock_guard<std::mutex> lk(g_all_mutex);
  if (g_all_devs == devs && !g_all_comms.empty()) {
    out = g_all_comms;
    return 0;
  }
  for (Comm* cm : g_all_comms) comm_free(cm);
  g_all_comms.clear();
  g_all_devs.clear();
  const NcclApi* api = nccl_api();
  if (!api) return set_error(SCTL_B200_ERR_NCCL, "%s", g_nccl_why.c_str());
  std::vector<ncclComm_t> nc(devs.size(), nullptr);
  DeviceGuard guard;
  NCCL_TRY(api, api->CommInitAll(nc.data(), (int)devs.size(), devs.data()));
  for (size_t g = 0; g < devs.size(); g++) {
    Comm* cm = new Comm();
    cm->nccl = nc[g];
    cm->rank = (int)g;
    cm->size = (int)devs.size();
    cm->dev = devs[g];
    g_all_comms.push_back(cm);
  }
  g_all_devs = devs;
  out = g_all_comms;
  return 0;
}

}  // namespace sctl_b200

using namespace sctl_b200;

extern "C" {

int sctl_b200_partition(int64_t n, int rank, int nranks, int64_t* begin, int64_t* end) {
  if (n < 0 || nranks < 1 || rank < 0 || rank >= nranks) return set_error(SCTL_B200_ERR_ARG, "bad partition arguments");
  // fmm-wrapper.txx:507: cnt*(rank+1)/np - cnt*rank/np elements for rank `rank`
  if (begin) *begin = (int64_t)((__int128)n * rank / nranks);
  if (end) *end = (int64_t)((__int128)n * (rank + 1) / nranks);
  return 0;
}

int sctl_b200_comm_unique_id(void* id, int64_t id_bytes) {
  if (!id || id_bytes < (int64_t)sizeof(ncclUniqueId)) return set_error(SCTL_B200_ERR_ARG, "id buffer must hold %zu bytes", sizeof(ncclUniqueId));
  const NcclApi* api = nccl_api();
  if (!api) return set_error(SCTL_B200_ERR_NCCL, "%s", g_nccl_why.c_str());
  ncclUniqueId uid;
  NCCL_TRY(api, api->GetUniqueId(&uid));
  memset(id, 0, (size_t)id_bytes);
  memcpy(id, &uid, sizeof(uid));
  return 0;
}

int sctl_b200_comm_create(void** comm, int nranks, int rank, const void* nccl_unique_id) {
  if (!comm || nranks < 1 || rank < 0 || rank >= nranks) return set_error(SCTL_B200_ERR_ARG, "bad communicator arguments");
  if (nranks > 1 && !nccl_unique_id) return set_error(SCTL_B200_ERR_ARG, "nccl_unique_id is required for nranks > 1");
  DevCtx* c = nullptr;
  int rc = get_ctx(&c);  // fails loudly without a device
  if (rc) return rc;
  Comm* cm = new Comm();
  cm->rank = rank;
  cm->size = nranks;
  cm->dev = c->dev;
  if (nranks > 1) {  // a single rank never needs NCCL: every collective degenerates to a copy
    const NcclApi* api = nccl_api();
    if (!api) {
      delete cm;
      return set_error(SCTL_B200_ERR_NCCL, "%s", g_nccl_why.c_str());
    }
    ncclUniqueId uid;
    memcpy(&uid, nccl_unique_id, sizeof(uid));
    ncclResult_t e = api->CommInitRank(&cm->nccl, nranks, uid, rank);
    if (e != ncclSuccess) {
      delete cm;
      return nccl_error(api, e, "ncclCommInitRank");
    }
  }
  *comm = cm;
  return 0;
}

int sctl_b200_comm_destroy(void* comm) {
  if (!comm) return 0;
  Comm* cm = (Comm*)comm;
  {
    DeviceGuard guard;
    if (cudaSetDevice(cm->dev) == cudaSuccess) cudaDeviceSynchronize();
  }
  comm_free(cm);
  return 0;
}

int sctl_b200_comm_info(void* comm, int* rank, int* nranks, int* device) {
  if (!comm) return set_error(SCTL_B200_ERR_ARG, "null communicator");
  const Comm* cm = (const Comm*)comm;
  if (rank) *rank = cm->rank;
  if (nranks) *nranks = cm->size;
  if (device) *device = cm->dev;
  return 0;
}

int sctl_b200_nccl_version(int* version) {
  if (!version) return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  const NcclApi* api = nccl_api();
  if (!api) return set_error(SCTL_B200_ERR_NCCL, "%s", g_nccl_why.c_str());
  NCCL_TRY(api, api->GetVersion(version));
  return 0;
}

int sctl_b200_eval_sharded(void* comm, int kernel_id, int dtype, int64_t Nt_local, int64_t Ns_local, const void* r_trg, const void* r_src,
                           const void* n_src, const void* v_src, void* v_trg, int digits, unsigned flags, const int64_t* src_counts) {
  if (!comm) return set_error(SCTL_B200_ERR_ARG, "null communicator");
  DevCtx* c = nullptr;
  int rc = get_ctx(&c);
  if (rc) return rc;
  return eval_sharded_impl((Comm*)comm, kernel_id, dtype, Nt_local, Ns_local, r_trg, r_src, n_src, v_src, v_trg, digits, flags, src_counts, c->stream, true);
}

int sctl_b200_eval_sharded_async(void* comm, int kernel_id, int dtype, int64_t Nt_local, int64_t Ns_local, const void* d_r_trg, const void* d_r_src,
                                 const void* d_n_src, const void* d_v_src, void* d_v_trg, int digits, unsigned flags, const int64_t* src_counts,
                                 void* cuda_stream) {
  if (!comm) return set_error(SCTL_B200_ERR_ARG, "null communicator");
  return eval_sharded_impl((Comm*)comm, kernel_id, dtype, Nt_local, Ns_local, d_r_trg, d_r_src, d_n_src, d_v_src, d_v_trg, digits,
                           flags | SCTL_B200_DEVICE_PTRS, src_counts, (cudaStream_t)cuda_stream, false);
}

}  // extern "C"
