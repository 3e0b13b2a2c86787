// sctl_b200_eval_noalloc: the device-pointer potential evaluation for callers that must not allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc). It has no counterpart in the reference.
// The tiled path needs working memory (packed source records + per-chunk partial sums, see
// eval_device_in_scratch in capi.cu); here that memory is ONE file-scope __device__ array of fixed size, which every GPU that
// loads the library holds until the process exits.
//
// g_scratch below is, in HOST code, only a handle for the symbol - not a device address. The address on each device comes from
// cudaGetSymbolAddress, asked once per device with that device current, and is what the kernels receive.
//
// Sharing: every call on a device uses the same array. Calls are ordered by the streams they are enqueued on; two calls whose
// kernels can overlap on the GPU (different streams without an event between them, or two host threads) race on the array and
// both results are undefined. Nothing in the array has to start at zero (every byte a call reads it has written itself before),
// so no clearing is done and nothing of an earlier call leaks into a later one.
#include <cuda_runtime.h>

#include <mutex>

#include "internal.h"

namespace sctl_b200 {
namespace {

static __device__ __align__(256) unsigned char g_scratch[SCTL_B200_NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
struct PerDevice {
  void* scratch = nullptr;  // device address of g_scratch on this device
  int sms = 0;
};
PerDevice g_dev[MAX_DEVICES];
std::mutex g_mu;

// set-up, once per device, with that device current (it is: the caller's current device is the one the pointers live on)
int device_setup(int dev, PerDevice* out) {
  std::lock_guard<std::mutex> lk(g_mu);
  PerDevice& d = g_dev[dev];
  if (!d.scratch) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_scratch);
    if (e != cudaSuccess) return cuda_error(e, "cudaGetSymbolAddress(g_scratch)");
    if (!ptr || ((uintptr_t)ptr & 255u)) return set_error(SCTL_B200_ERR_CUDA, "static scratch block: bad device address %p", ptr);
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return cuda_error(e, "cudaDeviceGetAttribute(MultiProcessorCount)");
    d.sms = sms;
    d.scratch = ptr;
  }
  *out = d;
  return 0;
}

}  // namespace
}  // namespace sctl_b200

using namespace sctl_b200;

extern "C" int sctl_b200_eval_noalloc(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* d_r_trg, const void* d_r_src,
                                      const void* d_n_src, const void* d_v_src, void* d_v_trg, int digits, unsigned flags, void* cuda_stream) {
  int rc = check_common(kernel_id, dtype, Nt, Ns);
  if (rc) return rc;
  const KernelInfo& ki = kernel_info(kernel_id);
  if ((Nt > 0 && (!d_r_trg || !d_v_trg)) || (Ns > 0 && Nt > 0 && (!d_r_src || !d_v_src || (ki.normal_dim && !d_n_src))))
    return set_error(SCTL_B200_ERR_ARG, "null pointer argument");
  if (flags & ~(unsigned)(SCTL_B200_DEVICE_PTRS | SCTL_B200_OVERWRITE))
    return set_error(SCTL_B200_ERR_ARG, "sctl_b200_eval_noalloc takes device pointers on one GPU: only SCTL_B200_OVERWRITE may be set");
  int dev = -1;
  CU_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES) return set_error(SCTL_B200_ERR_CUDA, "device ordinal %d out of range", dev);
  PerDevice d;
  if ((rc = device_setup(dev, &d))) return rc;
  return eval_device_in_scratch(kernel_id, dtype, Nt, Ns, d_r_trg, d_r_src, d_n_src, d_v_src, d_v_trg, digits, !(flags & SCTL_B200_OVERWRITE),
                                (cudaStream_t)cuda_stream, d.sms, d.scratch, (size_t)SCTL_B200_NOALLOC_SCRATCH_BYTES,
                                (size_t)SCTL_B200_NOALLOC_RECORD_BYTES);
}
