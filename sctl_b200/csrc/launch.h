// Host-side launch tables shared between the translation units of libsctl_b200.so.
#pragma once
#include <cuda_runtime.h>

#include "eval_impl.cuh"
#include "kmat_impl.cuh"

namespace sctl_b200 {

constexpr int NUM_KERNELS = 8;
constexpr int NUM_RVAR = 2;    // register-blocking variants: [0] = small-Nt (R=1), [1] = throughput (R=6; 4 for Stokes FxT)
constexpr int NUM_NEWTON = 3;  // rsqrt refinement levels (f32 uses 0..1)

struct EvalEntry {
  int rec_len;  // Reals per packed source record
  int k0, k1, nd;
  int R[NUM_RVAR];
  cudaError_t (*pack)(const void* r_src, const void* n_src, const void* v_src, const void* src_wts, void* rec, long long Ns, long long Ns_pad,
                      cudaStream_t st);
  // launches grid (n_ttiles, p.n_chunks)
  cudaError_t (*eval[NUM_RVAR][NUM_NEWTON])(const EvalParams& p, long long n_ttiles, cudaStream_t st);
  // eval[.][newton] accumulates 2^x2_shift[newton] times the sum (X2 mode, ukernels.cuh): the caller divides p.scale by it
  int x2_shift[NUM_NEWTON];
  // resident CTAs per SM for that instantiation (queried once)
  int (*occupancy[NUM_RVAR][NUM_NEWTON])();
  cudaError_t (*reduce)(const double* partial, void* v_trg, long long Nt, long long nt_stride, int n_chunks, double scale, int accumulate,
                        cudaStream_t st);
  // one-launch path for small problems straight from the AoS arrays (small_impl.cuh)
  cudaError_t (*small[NUM_NEWTON])(const void* r_trg, const void* r_src, const void* n_src, const void* v_src, const void* src_wts, void* v_trg,
                                   long long Nt, long long Ns, double scale, int accumulate, int sms, cudaStream_t st);
  cudaError_t (*kmat[NUM_NEWTON])(const void* r_trg, const void* r_src, const void* n_src, void* M, long long Nt, long long Ns, double scale,
                                  int sources_per_cta, cudaStream_t st);
  int kmat_cta_targets[2];  // targets per CTA of the KernelMatrix kernels, [1] = with the target-normal contraction
  // batched small-block KernelMatrix; [1] = with the target-normal contraction (null when TrgDim is not a multiple of 3)
  cudaError_t (*kmat_batch[2][NUM_NEWTON])(const void* r_trg, const void* r_src, const void* n_src, const void* src_wts, const void* n_trg, void* M,
                                           const KmatBatchDesc& D, long long n_ctas, double scale, cudaStream_t st);
};

const EvalEntry* eval_table_f64();  // [NUM_KERNELS]
const EvalEntry* eval_table_f32();

void count_launch(int n = 1);

// eval_f64.cu / eval_f32.cu: U = Ufull contracted with the target normals (see contract_normal_kernel)
cudaError_t contract_normal_f64(const void* Ufull, const void* n_trg, void* U, long long Nt, int K1out, cudaStream_t st);
cudaError_t contract_normal_f32(const void* Ufull, const void* n_trg, void* U, long long Nt, int K1out, cudaStream_t st);

// peak.cu
cudaError_t run_peak_f64(double seconds, double* tflops, cudaStream_t st);
cudaError_t run_peak_f32(double seconds, double* tflops, cudaStream_t st);

}  // namespace sctl_b200
