// Test hooks of libsctl_b200.so (declared at the end of include/sctl_b200.h): the device-side reciprocal square roots of
// common.cuh applied elementwise, so that tests/test_gpu_rsqrt.py can check the arithmetic of the hot loop directly against
// the reference's approx_rsqrt<digits> (vec.txx:361-364) and its own unit test (vec-test.hpp:384-418) instead of only
// through sums. Not used by any compute entry point.
#include <cuda_runtime.h>

#include "../../include/sctl_b200.h"
#include "common.cuh"

namespace sctl_b200 {
namespace {

// mode 0: rsqrt_unmasked, the seed's undefined low word taken from x itself (arbitrary mantissa bits)
// mode 1: rsqrt_masked  (r2 > 0 mask, subnormal rescaling)
// mode 2: rsqrt_unmasked with the seed's low word all ones (worst-case perturbation of the seed)
// mode 3: rsqrt_masked_sel (the branch-free complete form of the tile-repair loop)
__device__ __forceinline__ double all_ones_lo(double) { return __hiloint2double(0, -1); }
__device__ __forceinline__ float all_ones_lo(float) { return 0.0f; }  // the f32 seed has no undefined bits

template <class Real, int NEWTON>
__global__ void rsqrt_test_kernel(const Real* __restrict__ x, Real* __restrict__ y, long long n, int mode) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Real v = x[i];
  Real r;
  if (mode == 1) {
    r = rsqrt_masked<NEWTON>(v);
  } else if (mode == 3) {
    r = rsqrt_masked_sel<NEWTON>(v);
  } else if (mode == 2) {
    r = rsqrt_unmasked<NEWTON>(v, all_ones_lo(v));
  } else {
    r = rsqrt_unmasked<NEWTON>(v, v);
  }
  y[i] = r;
}

template <class Real>
cudaError_t run(int newton, int mode, long long n, const void* hx, void* hy) {
  Real *dx = nullptr, *dy = nullptr;
  cudaError_t e;
  if ((e = cudaMalloc(&dx, (size_t)n * sizeof(Real))) != cudaSuccess) return e;
  if ((e = cudaMalloc(&dy, (size_t)n * sizeof(Real))) != cudaSuccess) {
    cudaFree(dx);
    return e;
  }
  e = cudaMemcpy(dx, hx, (size_t)n * sizeof(Real), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    const unsigned blocks = (unsigned)((n + 255) / 256);
    if (newton == 0) rsqrt_test_kernel<Real, 0><<<blocks, 256>>>(dx, dy, n, mode);
    else if (newton == 1) rsqrt_test_kernel<Real, 1><<<blocks, 256>>>(dx, dy, n, mode);
    else rsqrt_test_kernel<Real, (sizeof(Real) == 8 ? 2 : 1)><<<blocks, 256>>>(dx, dy, n, mode);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(hy, dy, (size_t)n * sizeof(Real), cudaMemcpyDeviceToHost);
  cudaFree(dx);
  cudaFree(dy);
  return e;
}

}  // namespace
}  // namespace sctl_b200

extern "C" int sctl_b200_test_rsqrt(int dtype, int newton, int mode, int64_t n, const void* x, void* y) {
  if ((dtype != SCTL_B200_F32 && dtype != SCTL_B200_F64) || newton < 0 || newton > 2 || mode < 0 || mode > 3 || n < 0 || (n > 0 && (!x || !y)))
    return SCTL_B200_ERR_ARG;
  if (n == 0) return 0;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
    cudaGetLastError();
    return SCTL_B200_ERR_CUDA;
  }
  const cudaError_t e = dtype == SCTL_B200_F64 ? sctl_b200::run<double>(newton, mode, n, x, y) : sctl_b200::run<float>(newton, mode, n, x, y);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return SCTL_B200_ERR_CUDA;
  }
  return 0;
}
