// Internals of libsctl_b200.so shared between capi.cu (single-GPU entry points, particle registry) and comm.cu (the NCCL
// target-sharding driver). Nothing here is part of the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "../../include/sctl_b200.h"

namespace sctl_b200 {

int set_error(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
int cuda_error(cudaError_t e, const char* what);
const char* last_error_cstr();
#define CU_TRY(call)                                                    \
  do {                                                                  \
    cudaError_t e__ = (call);                                           \
    if (e__ != cudaSuccess) return ::sctl_b200::cuda_error(e__, #call); \
  } while (0)

struct KernelInfo {
  const char* name;
  int normal_dim, src_dim, trg_dim, flops;
  double scale_num, scale_den_pi;
};
const KernelInfo& kernel_info(int kernel_id);  // kernel_id must be valid
size_t elem_size(int dtype);
int check_common(int kernel_id, int dtype, int64_t Nt, int64_t Ns);

// per-thread, per-device context: one non-blocking stream + timing events + staging blocks for small host-pointer calls
struct DevCtx {
  int dev = -1;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  int sms = 0;
  void* h_stage = nullptr;  // pinned host block  (grow-only arena, see stage_reserve)
  void* d_stage = nullptr;  // device block of the same size
  size_t stage_cap = 0;
};
// The context of the calling thread for the CURRENT device. The pointer stays valid until the thread releases its contexts
// (sctl_b200_finalize / thread exit): contexts live in a std::deque, which never relocates its elements.
int get_ctx(DevCtx** out);
void release_thread_ctx();

// RAII: remember the caller's current device, restore it on scope exit
struct DeviceGuard {
  int prev = -1;
  DeviceGuard() {
    if (cudaGetDevice(&prev) != cudaSuccess) {
      cudaGetLastError();
      prev = -1;
    }
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

struct DevAlloc {  // stream-ordered allocation, freed on the same stream
  void* p = nullptr;
  cudaStream_t st = nullptr;
  ~DevAlloc() {
    if (p) cudaFreeAsync(p, st);
  }
  int alloc(size_t bytes, cudaStream_t s);
};

struct DevBuf {  // grow-only device buffer (cudaMalloc), freed with its owner
  void* p = nullptr;
  size_t cap = 0;
  ~DevBuf() {
    if (p) cudaFree(p);
  }
  int reserve(size_t bytes);
};

// pack + eval (+ reduce) on device-resident AoS arrays of the current device; enqueues on `st`, does not synchronise.
// ev (optional): 4 events recorded around pack | eval | reduce.
int eval_device(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src, const void* v_src,
                void* v_trg, int digits, bool accumulate, cudaStream_t st, int sms, cudaEvent_t* ev);

// The same tiled sequence with records and partial sums inside `scratch` (device address, 256-byte aligned, scratch_bytes long)
// instead of stream-ordered allocations; allocates nothing. Refuses source sets whose packed records exceed rec_max_bytes.
int eval_device_in_scratch(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                           const void* v_src, void* v_trg, int digits, bool accumulate, cudaStream_t st, int sms, void* scratch,
                           size_t scratch_bytes, size_t rec_max_bytes);

// ------------------------------------------------------------------------------------------------------------------
// comm.cu: NCCL communicator handle (sctl_b200_comm_*) and the collectives the registry uses
// ------------------------------------------------------------------------------------------------------------------
struct Comm;
int comm_rank(const Comm* cm);
int comm_size(const Comm* cm);
int comm_device(const Comm* cm);
// number of particles every rank holds: one ncclAllGather of an int64 + a stream synchronisation
int comm_counts(Comm* cm, int64_t n_local, std::vector<int64_t>& counts, cudaStream_t st);
// contiguous replica of every rank's block (counts[r] * width_bytes bytes from rank r, in rank order) in d_out; d_local is this
// rank's block (device pointer). ONE ncclAllGather when the blocks are equal, a group of in-place ncclBroadcasts when ragged.
int comm_allgatherv(Comm* cm, const void* d_local, size_t width_bytes, const std::vector<int64_t>& counts, void* d_out, cudaStream_t st);
// the all-GPU communicators of this process (ncclCommInitAll over `devs`), created on first use; nullptr entries if NCCL is unavailable
int comm_all_gpus(const std::vector<int>& devs, std::vector<Comm*>& out);
void comm_all_gpus_release();
int eval_sharded_impl(Comm* cm, int kernel_id, int dtype, int64_t Nt, int64_t Ns_local, const void* r_trg, const void* r_src, const void* n_src,
                      const void* v_src, void* v_trg, int digits, unsigned flags, const int64_t* src_counts, cudaStream_t st, bool own_stream);

}  // namespace sctl_b200
