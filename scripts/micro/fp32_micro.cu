// FP32-pipe microbenchmarks for sm_100a: does the packed FFMA2 (fma.rn.f32x2) free issue slots for co-issued MUFU / LDS work?
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o build/fp32_micro scripts/micro/fp32_micro.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

constexpr int NF = 24;  // scalar-FMA equivalents per "pair" (12 FFMA2)

// PACKED: 0 = scalar FFMA (3 distinct regs), 1 = FFMA2
template <int PACKED, int NMUFU, int NALU, int NLDS>
__global__ void __launch_bounds__(128, 4) k(float* out, const float* in, int iters) {
  __shared__ float4 sh[64];
  if (threadIdx.x < 64) sh[threadIdx.x] = make_float4(in[0], in[1], in[2], in[3]);
  __syncthreads();
  float2 x[12], y[4], z[4];
#pragma unroll
  for (int i = 0; i < 12; i++) x[i] = make_float2(in[i & 7] + threadIdx.x * 1e-6f, in[(i + 1) & 7]);
#pragma unroll
  for (int i = 0; i < 4; i++) { y[i] = make_float2(in[8 + i], in[9 + i]); z[i] = make_float2(in[12 + i], in[13 + i]); }
  float m32 = in[17];
  int alu = threadIdx.x;
  const unsigned shbase = (unsigned)__cvta_generic_to_shared(sh);
  for (int it = 0; it < iters; it++) {
#pragma unroll 4
    for (int rep = 0; rep < 4; rep++) {
#pragma unroll
      for (int j = 0; j < NF / 2; j++) {
        const int i = (j + rep * 5) % 12;
        if (PACKED) x[i] = __ffma2_rn(y[j & 3], z[(j >> 2) & 3], x[i]);
        else { x[i].x = fmaf(y[j & 3].x, z[(j >> 2) & 3].x, x[i].x); x[i].y = fmaf(y[j & 3].y, z[(j >> 2) & 3].y, x[i].y); }
      }
#pragma unroll
      for (int j = 0; j < NMUFU; j++) { float r; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(m32)); m32 = r; }
#pragma unroll
      for (int j = 0; j < NALU; j++) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(alu) : "r"(it), "r"(j)); }
#pragma unroll
      for (int j = 0; j < NLDS; j++) { float v0, v1, v2, v3; asm volatile("ld.volatile.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v0), "=f"(v1), "=f"(v2), "=f"(v3) : "r"(shbase + 16u * (unsigned)((j + rep) & 63))); }
    }
  }
  float s = m32 + alu;
#pragma unroll
  for (int i = 0; i < 12; i++) s += x[i].x + x[i].y;
  if (s == 123.456f) out[0] = s;
}

template <int PACKED, int NMUFU, int NALU, int NLDS>
void run(const char* name, float* d_out, float* d_in, int sms, double clk_ghz) {
  auto kern = k<PACKED, NMUFU, NALU, NLDS>;
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 128, 0));
  const int blocks = sms * occ;
  const int iters = 40000;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  kern<<<blocks, 128>>>(d_out, d_in, 200);
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < 3; r++) {
    CK(cudaEventRecord(e0));
    kern<<<blocks, 128>>>(d_out, d_in, iters);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  const double fma = (double)blocks * 128 * iters * 4.0 * NF;
  const double per_clk_sm = fma / (best * 1e-3) / (clk_ghz * 1e9) / sms;
  printf("%-44s occ=%2d ms=%8.3f  FP32 FMA lanes/clk/SM = %6.2f (%.1f%% of 128)  TFLOP/s=%.2f\n", name, occ, best, per_clk_sm, per_clk_sm / 128 * 100,
         fma * 2 / (best * 1e-3) * 1e-12);
}

int main(int argc, char** argv) {
  const double clk = argc > 1 ? atof(argv[1]) : 1.965;
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  float *d_out, *d_in;
  CK(cudaMalloc(&d_out, 64)); CK(cudaMalloc(&d_in, 32 * 4));
  std::vector<float> h(32);
  for (int i = 0; i < 32; i++) h[i] = 0.999f + 1e-4f * i;
  CK(cudaMemcpy(d_in, h.data(), 32 * 4, cudaMemcpyHostToDevice));
  printf("SMs=%d assumed clock %.3f GHz; one 'pair' = %d scalar-FMA equivalents\n", sms, clk, NF);
  run<0, 0, 0, 0>("scalar FFMA, 3 distinct regs", d_out, d_in, sms, clk);
  run<1, 0, 0, 0>("packed FFMA2", d_out, d_in, sms, clk);
  run<0, 1, 0, 0>("scalar FFMA + 1 MUFU /24", d_out, d_in, sms, clk);
  run<1, 1, 0, 0>("packed FFMA2 + 1 MUFU /24", d_out, d_in, sms, clk);
  run<0, 1, 2, 1>("scalar FFMA + 1 MUFU + 2 ALU + 1 LDS /24", d_out, d_in, sms, clk);
  run<1, 1, 2, 1>("packed FFMA2 + 1 MUFU + 2 ALU + 1 LDS /24", d_out, d_in, sms, clk);
  run<1, 2, 4, 2>("packed FFMA2 + 2 MUFU + 4 ALU + 2 LDS /24", d_out, d_in, sms, clk);
  run<1, 1, 8, 1>("packed FFMA2 + 1 MUFU + 8 ALU + 1 LDS /24", d_out, d_in, sms, clk);
  return 0;
}
