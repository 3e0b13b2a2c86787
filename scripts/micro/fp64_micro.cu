// FP64-pipe microbenchmarks for sm_100a: what does a DFMA-bound loop lose to co-issued MUFU / ALU / LDS work,
// to operand patterns and to occupancy?  Scratch tool (not part of the library): results are summarised in
// profiles/. Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o build/fp64_micro scripts/micro/fp64_micro.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

constexpr int NF = 22;  // FP64 instructions per "pair"

// MODE 0: x = fma(x, a, b) (a, b loop-invariant)   MODE 1: x_i = fma(y_j, z_k, x_i) (three distinct live registers)
// MODE 2: mix 3 DADD + 5 DMUL + 14 DFMA
template <int MODE, int NACC, int NMUFU64, int NMUFU32, int NALU, int NLDS, int MINB>
__global__ void __launch_bounds__(128, MINB) k(double* out, const double* in, int iters) {
  __shared__ double4 sh[64];
  if (threadIdx.x < 64) sh[threadIdx.x] = make_double4(in[0], in[1], in[2], in[3]);
  __syncthreads();
  double x[NACC], y[4], z[4];
#pragma unroll
  for (int i = 0; i < NACC; i++) x[i] = in[i & 7] + threadIdx.x * 1e-9;
#pragma unroll
  for (int i = 0; i < 4; i++) { y[i] = in[8 + i]; z[i] = in[12 + i]; }
  double m64 = in[16];
  float m32 = (float)in[17];
  int alu = threadIdx.x;
  int ldsidx = threadIdx.x & 31 ? 0 : 0;
  const unsigned shbase = (unsigned)__cvta_generic_to_shared(sh);
  for (int it = 0; it < iters; it++) {
#pragma unroll 4
    for (int rep = 0; rep < 4; rep++) {
#pragma unroll
      for (int j = 0; j < NF; j++) {
        const int i = (j + rep * NF) % NACC;
        if (MODE == 0) x[i] = fma(x[i], y[0], z[0]);
        else if (MODE == 1) x[i] = fma(y[j & 3], z[(j >> 2) & 3], x[i]);
        else {
          if (j < 3) x[i] = x[i] + y[j & 3];
          else if (j < 8) x[i] = x[i] * z[j & 3];
          else x[i] = fma(y[j & 3], z[(j >> 2) & 3], x[i]);
        }
      }
#pragma unroll
      for (int j = 0; j < NMUFU64; j++) { double r; asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(m64)); m64 = r; }
#pragma unroll
      for (int j = 0; j < NMUFU32; j++) { float r; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(m32)); m32 = r; }
#pragma unroll
      for (int j = 0; j < NALU; j++) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(alu) : "r"(it), "r"(j)); }
#pragma unroll
      for (int j = 0; j < NLDS; j++) { double v0, v1; asm volatile("ld.volatile.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v0), "=d"(v1) : "r"(shbase + 32u * (unsigned)((ldsidx + j + rep) & 63))); }
    }
  }
  double s = m64 + m32 + alu;
#pragma unroll
  for (int i = 0; i < NACC; i++) s += x[i];
  if (s == 123.456) out[0] = s;
}

template <int MODE, int NACC, int NMUFU64, int NMUFU32, int NALU, int NLDS, int MINB>
void run(const char* name, double* d_out, double* d_in, int sms, double clk_ghz) {
  auto kern = k<MODE, NACC, NMUFU64, NMUFU32, NALU, NLDS, MINB>;
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 128, 0));
  const int blocks = sms * occ;
  const int iters = 20000;
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
  const double fp64_instr = (double)blocks * 128 * iters * 4.0 * NF;
  const double per_clk_sm = fp64_instr / (best * 1e-3) / (clk_ghz * 1e9) / sms;
  printf("%-46s occ=%d warps/SMSP=%d  ms=%8.3f  FP64 lanes/clk/SM = %6.2f (%.1f%% of 64)  TFLOP/s(2/instr)=%.2f\n", name, occ, occ, best, per_clk_sm,
         per_clk_sm / 64 * 100, fp64_instr * 2 / (best * 1e-3) * 1e-12);
}

int main(int argc, char** argv) {
  const double clk = argc > 1 ? atof(argv[1]) : 1.965;
  int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  double *d_out, *d_in;
  CK(cudaMalloc(&d_out, 64)); CK(cudaMalloc(&d_in, 32 * 8));
  std::vector<double> h(32);
  for (int i = 0; i < 32; i++) h[i] = 0.999 + 1e-4 * i;
  CK(cudaMemcpy(d_in, h.data(), 32 * 8, cudaMemcpyHostToDevice));
  printf("SMs=%d assumed clock %.3f GHz\n", sms, clk);
  //        MODE NACC M64 M32 ALU LDS MINB
  run<0, 16, 0, 0, 0, 0, 4>("dfma x=fma(x,a,b) 16 acc, 4 warps/SMSP", d_out, d_in, sms, clk);
  run<0, 16, 0, 0, 0, 0, 8>("dfma x=fma(x,a,b) 16 acc, 8 warps/SMSP", d_out, d_in, sms, clk);
  run<0, 8, 0, 0, 0, 0, 16>("dfma x=fma(x,a,b) 8 acc, 16 warps/SMSP", d_out, d_in, sms, clk);
  run<1, 16, 0, 0, 0, 0, 4>("dfma 3 distinct regs, 16 acc, 4 w", d_out, d_in, sms, clk);
  run<1, 16, 0, 0, 0, 0, 8>("dfma 3 distinct regs, 16 acc, 8 w", d_out, d_in, sms, clk);
  run<1, 8, 0, 0, 0, 0, 4>("dfma 3 distinct regs, 8 acc, 4 w", d_out, d_in, sms, clk);
  run<1, 4, 0, 0, 0, 0, 4>("dfma 3 distinct regs, 4 acc, 4 w", d_out, d_in, sms, clk);
  run<2, 16, 0, 0, 0, 0, 4>("mix 3 dadd+5 dmul+14 dfma, 4 w", d_out, d_in, sms, clk);
  run<1, 16, 1, 0, 0, 0, 4>("+1 MUFU.RSQ64H /22", d_out, d_in, sms, clk);
  run<1, 16, 2, 0, 0, 0, 4>("+2 MUFU.RSQ64H /22", d_out, d_in, sms, clk);
  run<1, 16, 0, 1, 0, 0, 4>("+1 MUFU.RSQ(f32) /22", d_out, d_in, sms, clk);
  run<1, 16, 0, 2, 0, 0, 4>("+2 MUFU.RSQ(f32) /22", d_out, d_in, sms, clk);
  run<1, 16, 0, 0, 4, 0, 4>("+4 ALU /22", d_out, d_in, sms, clk);
  run<1, 16, 0, 0, 8, 0, 4>("+8 ALU /22", d_out, d_in, sms, clk);
  run<1, 16, 0, 0, 16, 0, 4>("+16 ALU /22", d_out, d_in, sms, clk);
  run<1, 16, 0, 0, 0, 1, 4>("+1 LDS.128 /22", d_out, d_in, sms, clk);
  run<1, 16, 1, 0, 4, 1, 4>("+1 RSQ64H +4 ALU +1 LDS /22 (like eval)", d_out, d_in, sms, clk);
  run<1, 16, 0, 1, 8, 1, 4>("+1 RSQ32 +8 ALU +1 LDS /22", d_out, d_in, sms, clk);
  run<1, 16, 1, 0, 4, 1, 8>("+1 RSQ64H +4 ALU +1 LDS /22, 8 w", d_out, d_in, sms, clk);
  return 0;
}
