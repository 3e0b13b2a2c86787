/* sctl_b200.h — C ABI of the B200-native direct particle-kernel summation library.
 *
 * This is the drop-in boundary for ONE hot path of SCTL (reference tree: /root/reference):
 *
 *   GenericKernel<uKernel>::Eval (static, type-erased, runtime digits)  include/sctl/generic-kernel.txx:46-74
 *   GenericKernel<uKernel>::Eval<Real,enable_openmp,digits>             include/sctl/generic-kernel.txx:76-189
 *   GenericKernel<uKernel>::KernelMatrix<Real,enable_openmp,digits>     include/sctl/generic-kernel.txx:191-307
 *   CoordDim/NormalDim/SrcDim/TrgDim, Name(), FLOPS(), uKerScaleFactor  include/sctl/generic-kernel.txx:22-36,
 *                                                                       include/sctl/kernel_functions.hpp:16-25 ...
 *   the per-rank body of ParticleFMM::EvalDirect                        include/sctl/fmm-wrapper.txx:490-562
 *
 * The reference is a header-only C++ template library; the binding a maintainer adds is the header
 * include/sctl_b200/generic-kernel-b200.hpp (explicit specialisations of the member templates above that
 * forward here) — see INTEGRATION.md.
 *
 * Conventions
 *  - Plain C types only. All arrays are contiguous AoS exactly as the reference's Vector<Real> holds them:
 *      r_trg[Nt*3] = {x0,y0,z0,x1,...}   r_src[Ns*3]   n_src[Ns*NormalDim]   v_src[Ns*SrcDim]   v_trg[Nt*TrgDim]
 *    n_src may be NULL iff NormalDim == 0 (generic-kernel.txx:97).
 *  - dX = x_trg - x_src (generic-kernel.txx:83); pairs with |dX|^2 == 0 contribute exactly 0 and every pair with
 *    |dX|^2 > 0 - subnormal values included - a finite rinv (the r2 > 0 mask of kernel_functions.hpp:28); the scale
 *    factor is applied once after the sum (generic-kernel.txx:184).
 *  - Every function returns 0 on success or a SCTL_B200_ERR_* code; sctl_b200_last_error() returns a
 *    thread-local message. There is NO CPU fallback: without a usable CUDA device every compute call fails
 *    with SCTL_B200_ERR_CUDA. The reference aborts on bad sizes (generic-kernel.txx:94-97); the C++ shim maps
 *    a non-zero return to SCTL_ERROR to preserve that.
 *  - Thread-safe and re-entrant: the reference's KernelMatrix is called from inside OpenMP parallel regions
 *    (boundary_integral.txx:949-986). Each host thread owns its stream and device workspace.
 *  - Results are NOT bit-identical to the CPU path (different summation order, different rsqrt refinement);
 *    the contract is rel-L2 <= 1e-12 (f64) / 1e-5 (f32) against the reference's CPU Eval.
 */
#ifndef SCTL_B200_H_
#define SCTL_B200_H_

#include <stdint.h>

#if defined(__GNUC__)
#define SCTL_B200_API __attribute__((visibility("default")))
#else
#define SCTL_B200_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* kernel ids: the eight GenericKernel typedefs of kernel_functions.hpp:207-214 */
enum {
  SCTL_B200_LAPLACE3D_FXU = 0,  /* kernel_functions.hpp:15-31   K 1x1, no normal, 1/(4 pi)  */
  SCTL_B200_LAPLACE3D_DXU = 1,  /* kernel_functions.hpp:33-51   K 1x1, normal,    1/(4 pi)  */
  SCTL_B200_LAPLACE3D_FXDU = 2, /* kernel_functions.hpp:53-72   K 1x3, no normal, -1/(4 pi) */
  SCTL_B200_STOKES3D_FXU = 3,   /* kernel_functions.hpp:74-95   K 3x3, no normal, 1/(8 pi)  */
  SCTL_B200_STOKES3D_DXU = 4,   /* kernel_functions.hpp:97-120  K 3x3, normal,    3/(4 pi)  */
  SCTL_B200_STOKES3D_FXT = 5,   /* kernel_functions.hpp:122-146 K 3x9, no normal, -3/(4 pi) */
  SCTL_B200_STOKES3D_FSXU = 6,  /* kernel_functions.hpp:148-172 K 4x3, no normal, 1/(8 pi)  */
  SCTL_B200_STOKES3D_FXUP = 7,  /* kernel_functions.hpp:174-198 K 3x4, no normal, 1/(8 pi)  */
  SCTL_B200_NUM_KERNELS = 8
};

/* dtype: Real of the reference's template parameter */
enum { SCTL_B200_F32 = 0, SCTL_B200_F64 = 1 };

/* flags */
enum {
  SCTL_B200_DEVICE_PTRS = 1u << 0, /* all pointers are device pointers on the current CUDA device */
  SCTL_B200_OVERWRITE = 1u << 1,   /* v_trg = scale*sum instead of v_trg += scale*sum (EvalDirect zeroes U first, fmm-wrapper.txx:501-502) */
  SCTL_B200_ALL_GPUS = 1u << 2     /* host pointers only: shard targets over every visible GPU of this process (fmm-wrapper.txx:504-510 analogue);
                                      each GPU uploads 1/G of the sources and NCCL replicates them over NVLink */
};

/* error codes */
enum {
  SCTL_B200_OK = 0,
  SCTL_B200_ERR_KERNEL = 1, /* unknown kernel id */
  SCTL_B200_ERR_DTYPE = 2,  /* unknown dtype */
  SCTL_B200_ERR_ARG = 3,    /* null pointer / negative size / bad flag combination */
  SCTL_B200_ERR_CUDA = 4,   /* CUDA runtime error or no usable device (no CPU fallback) */
  SCTL_B200_ERR_NOMEM = 5,  /* device allocation failed */
  SCTL_B200_ERR_NCCL = 6    /* NCCL could not be loaded, or an NCCL call failed */
};

/* Metadata — replaces CoordDim()/NormalDim()/SrcDim()/TrgDim() (generic-kernel.txx:22-36),
 * uKerScaleFactor<double>() and FLOPS() (kernel_functions.hpp). Any out pointer may be NULL. No GPU needed. */
SCTL_B200_API int sctl_b200_kernel_dims(int kernel_id, int* coord_dim, int* normal_dim, int* src_dim, int* trg_dim, double* scale, int* flops);
/* Name() of the micro-kernel, e.g. "Stokes3D-FxU" (kernel_functions.hpp:16-19); NULL for a bad id. */
SCTL_B200_API const char* sctl_b200_kernel_name(int kernel_id);

/* Potential evaluation — replaces GenericKernel::Eval (generic-kernel.txx:46-74 -> 76-189):
 *   v_trg[t*K1+k1] (+)= scale * sum_s sum_k0 K[k0][k1](x_t - x_s, n_s) * v_src[s*K0+k0]
 * digits: -1 or >= 16 = machine precision; 0..15 = requested significant digits (the GPU path may deliver
 * more, never fewer — the reference only ever relaxes the rsqrt refinement, intrin-wrapper.hpp:3026).
 * Host pointers (default): synchronous; inputs are copied to the device, result copied back.
 * SCTL_B200_DEVICE_PTRS: runs on the calling thread's library stream and synchronises before returning. */
SCTL_B200_API int sctl_b200_eval(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src, const void* n_src,
                   const void* v_src, void* v_trg, int digits, unsigned flags);

/* Same with device pointers, enqueued on a caller-supplied cudaStream_t (passed as void*; NULL = default
 * stream) WITHOUT synchronising: for callers that keep Vector storage device-resident. */
SCTL_B200_API int sctl_b200_eval_async(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* d_r_trg, const void* d_r_src,
                         const void* d_n_src, const void* d_v_src, void* d_v_trg, int digits, unsigned flags, void* cuda_stream);

/* sctl_b200_eval_async that allocates NOTHING. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, arithmetic and result as sctl_b200_eval_async; flags may hold SCTL_B200_OVERWRITE only. It always takes the
 * tiled path (pack -> eval_kernel -> reduce, results bit-identical to sctl_b200_eval_async whenever that picks the same plan) and
 * keeps the path's working memory - packed source records and per-chunk partial sums - in ONE static __device__ array of
 * SCTL_B200_NOALLOC_SCRATCH_BYTES (32 MiB) that every GPU loading the library holds until the process exits.
 * Limit: the packed records, ceil(Ns/256)*256 * record bytes, must fit SCTL_B200_NOALLOC_RECORD_BYTES (24 MiB); a record is
 * (3 + source payload) Reals rounded up to 16 bytes, at most 80 bytes (Stokes3D_DxU f64), so Ns <= SCTL_B200_NOALLOC_MAX_SRC
 * (314 368) is accepted for every kernel and dtype, Stokes3D_FxU f64 (48-byte records) up to 524 288. Larger source sets return
 * SCTL_B200_ERR_ARG with the limit in sctl_b200_last_error() and launch nothing. Nt is not limited: the partial sums use what the
 * records leave (>= 8 MiB) and the source range is split into fewer chunks when they would not fit.
 * Calls no cudaMalloc / cudaFree / stream-ordered allocator; the first call on each device resolves the array's address
 * (cudaGetSymbolAddress, which loads the module: make that call outside a capture).
 * Every call on a device shares the array: calls are safe when stream order (or events) serialises them; calls that can
 * OVERLAP on the GPU - different streams, several host threads - race and give undefined results. The array keeps the last
 * call's contents; no call reads bytes it has not written itself, so nothing needs clearing between calls. */
#define SCTL_B200_NOALLOC_SCRATCH_BYTES (32u << 20)
#define SCTL_B200_NOALLOC_RECORD_BYTES (24u << 20)
#define SCTL_B200_NOALLOC_MAX_SRC 314368
SCTL_B200_API int sctl_b200_eval_noalloc(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* d_r_trg, const void* d_r_src,
                           const void* d_n_src, const void* d_v_src, void* d_v_trg, int digits, unsigned flags, void* cuda_stream);

/* Dense operator — replaces GenericKernel::KernelMatrix (generic-kernel.txx:191-307):
 *   M[(s*K0+k0) * (Nt*K1) + (t*K1+k1)] = scale * K[k0][k1](x_t - x_s, n_s)       (row-major, every entry overwritten) */
SCTL_B200_API int sctl_b200_kernel_matrix(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* r_trg, const void* r_src,
                            const void* n_src, void* M, int digits, unsigned flags);
SCTL_B200_API int sctl_b200_kernel_matrix_async(int kernel_id, int dtype, int64_t Nt, int64_t Ns, const void* d_r_trg, const void* d_r_src,
                                  const void* d_n_src, void* d_M, int digits, unsigned flags, void* cuda_stream);

/* Batched small-block KernelMatrix — ONE launch for what the reference's BoundaryIntegralOp::SetupNear does with one
 * KernelMatrix call per element (boundary_integral.txx:946-1008). Block b couples targets [trg_dsp[b], trg_dsp[b+1]) of r_trg
 * (and n_trg) with sources [src_dsp[b], src_dsp[b+1]) of r_src / n_src / src_wts — the reference's near_elem_dsp /
 * elem_nds_dsp_far displacements, in POINTS — and is written row-major at M + M_dsp[b] (elements) as a
 * (Ns_b*K0) x (Nt_b*K1out) matrix. The three displacement arrays are always HOST pointers (trg_dsp, src_dsp: nbatch+1
 * entries; M_dsp: nbatch). Optional fused epilogues of that call site:
 *   src_wts != NULL : row (s,k0) is multiplied by src_wts[s]                                   (boundary_integral.txx:990-997)
 *   n_trg   != NULL : K1out = K1/3 and M[.][t*K1out+k] = sum_l K[.][k*3+l] * n_trg[t*3+l]      (boundary_integral.txx:968-984;
 *                     needs TrgDim % 3 == 0: Laplace3D_FxdU, Stokes3D_FxT); otherwise K1out = K1.
 * Entries of M not covered by any block are left untouched. */
SCTL_B200_API int sctl_b200_kernel_matrix_batch(int kernel_id, int dtype, int64_t nbatch, const int64_t* trg_dsp, const int64_t* src_dsp,
                                  const int64_t* M_dsp, const void* r_trg, const void* r_src, const void* n_src, const void* src_wts,
                                  const void* n_trg, void* M, int64_t M_len, int digits, unsigned flags);

/* Device-resident Vector storage (north_star: "device-resident Vector buffers"): raw device memory on the
 * current device, usable with SCTL_B200_DEVICE_PTRS / *_async. */
SCTL_B200_API int sctl_b200_buf_create(void** d_ptr, int dtype, int64_t n);
SCTL_B200_API int sctl_b200_buf_upload(void* d_ptr, const void* host, int dtype, int64_t n);
SCTL_B200_API int sctl_b200_buf_download(const void* d_ptr, void* host, int dtype, int64_t n);
SCTL_B200_API int sctl_b200_buf_zero(void* d_ptr, int dtype, int64_t n);                    /* Vector::SetZero (vector.txx:163) */
SCTL_B200_API int sctl_b200_buf_destroy(void* d_ptr);

/* Device-resident particle context — the GPU-aware counterpart of ParticleFMM's registry and EvalDirect
 * (fmm-wrapper.txx:297-479 and :490-562). Coordinates, normals, densities and the packed source records stay in
 * HBM between calls; records are re-packed only when a source's coordinates or density changed. Sources / targets
 * are declared with the dims the reference derives from the kernels passed to AddSrc / AddTrg
 * (fmm-wrapper.txx:306-309, :346). Inputs are COPIED at Set* time like the reference does (:447-448), from host
 * pointers or (SCTL_B200_DEVICE_PTRS) device pointers. eval_direct OVERWRITES U (:501-502) with the sum over every
 * source that has an S2T kernel for the target, in (src, trg) name order like the reference's std::map walk (:514).
 * Far-field FMM (PVFMM) is not part of this library: Eval == EvalDirect, as in a reference build without PVFMM (:481-489). */
SCTL_B200_API int sctl_b200_fmm_create(void** fmm, int dtype);
SCTL_B200_API int sctl_b200_fmm_destroy(void* fmm);
SCTL_B200_API int sctl_b200_fmm_set_accuracy(void* fmm, int digits);                                   /* SetAccuracy */
SCTL_B200_API int sctl_b200_fmm_add_src(void* fmm, const char* name, int src_dim, int normal_dim);     /* AddSrc  fmm-wrapper.txx:297 */
SCTL_B200_API int sctl_b200_fmm_add_trg(void* fmm, const char* name, int trg_dim);                     /* AddTrg  :337 */
SCTL_B200_API int sctl_b200_fmm_set_kernel_s2t(void* fmm, const char* src_name, const char* trg_name, int kernel_id); /* SetKernelS2T :371 */
SCTL_B200_API int sctl_b200_fmm_delete_src(void* fmm, const char* name);                               /* DeleteSrc :407 */
SCTL_B200_API int sctl_b200_fmm_delete_trg(void* fmm, const char* name);                               /* DeleteTrg :419 */
SCTL_B200_API int sctl_b200_fmm_set_src_coord(void* fmm, const char* name, int64_t Ns, const void* r_src, const void* n_src, unsigned flags); /* :444 */
SCTL_B200_API int sctl_b200_fmm_set_src_density(void* fmm, const char* name, int64_t n, const void* v_src, unsigned flags);                  /* :462 */
SCTL_B200_API int sctl_b200_fmm_set_trg_coord(void* fmm, const char* name, int64_t Nt, const void* r_trg, unsigned flags);                   /* :467 */
/* Fused glue of BoundaryIntegralOp::ComputeFarField (optional; n == 0 removes it again):
 *   weights : the density of source `name` is multiplied by wts[s] when records are packed   (boundary_integral.txx:1036-1051)
 *   normals : eval_direct returns U[t*K1'+k] = sum_l Usum[(t*K1'+k)*3+l] * n_trg[t*3+l], K1' = TrgDim/3   (:1060-1071) */
SCTL_B200_API int sctl_b200_fmm_set_src_weights(void* fmm, const char* name, int64_t n, const void* wts, unsigned flags);
SCTL_B200_API int sctl_b200_fmm_set_trg_normal(void* fmm, const char* name, int64_t n, const void* n_trg, unsigned flags);
/* Nt and the number of values per target that eval_direct returns (TrgDim, or TrgDim/3 with target normals) */
SCTL_B200_API int sctl_b200_fmm_trg_count(void* fmm, const char* name, int64_t* Nt, int* trg_dim);
SCTL_B200_API int sctl_b200_fmm_eval_direct(void* fmm, const char* trg_name, void* U, int64_t n, unsigned flags);                            /* :490 */

/* Device management. One process per GPU is the normal multi-GPU mode (each MPI / torchrun rank calls
 * sctl_b200_set_device(local_rank) once, as the reference's ranks each own a target block,
 * fmm-wrapper.txx:504-510). */
SCTL_B200_API int sctl_b200_device_count(int* n);
SCTL_B200_API int sctl_b200_set_device(int dev);
SCTL_B200_API int sctl_b200_get_device(int* dev);
SCTL_B200_API int sctl_b200_sm_count(int* n);
/* In-process multi-GPU driver (SCTL_B200_ALL_GPUS); plays the role of the communicator handed to ParticleFMM(comm)
 * (fmm-wrapper.txx:204, fmm-wrapper.hpp:45) for the target partition of EvalDirect (fmm-wrapper.txx:504-510). sctl_b200_init selects the devices the flag shards over (ndev == 0 or
 * dev_ids == NULL: every visible device, the default without a call) and creates their CUDA contexts up front, so the first
 * Eval does not pay for it; sctl_b200_finalize waits for those devices, releases the calling thread's streams, events and
 * staging blocks and forgets the selection. Both are optional: every entry point initialises lazily. */
SCTL_B200_API int sctl_b200_init(int ndev, const int* dev_ids);
SCTL_B200_API int sctl_b200_finalize(void);

/* NCCL target-sharding driver: one rank per GPU (MPI ranks, torchrun workers, or threads of one process), every rank
 * holding ITS block of the targets and ITS block of every source array - the multi-rank form of ParticleFMM::EvalDirect
 * (fmm-wrapper.txx:490-562). The reference sends every source block round an Isend/Irecv ring and evaluates after each
 * round (:537-558); here the blocks are replicated with one ncclAllGather per source array over NVLink (a group of
 * ncclBroadcasts when the blocks are ragged) and ONE Eval runs over all sources. Targets never move, so there is no
 * reduction and results stay with their owner (the reference re-partitions U back, :560-561).
 *   sctl_b200_partition       [begin, end) of rank's block of n items: the formula of fmm-wrapper.txx:507 (PartitionN).
 *   sctl_b200_comm_unique_id  rank 0 creates the 128-byte NCCL id and distributes it by any means it has (MPI_Bcast,
 *                             sctl::Comm::Bcast, torch.distributed, a file); id_bytes >= 128.
 *   sctl_b200_comm_create     collective over all ranks: ncclCommInitRank on the CURRENT device (call
 *                             sctl_b200_set_device first). nranks == 1 needs no id and never loads NCCL.
 *   sctl_b200_eval_sharded    v_trg[Nt_local*K1] (+)= scale * sum over ALL ranks' sources. Collective: every rank of the
 *                             communicator must call it with the same kernel_id / dtype / digits. Host pointers (default;
 *                             each rank uploads only its own blocks) or SCTL_B200_DEVICE_PTRS; SCTL_B200_OVERWRITE as for
 *                             sctl_b200_eval. src_counts: the number of sources of every rank (nranks entries), or NULL
 *                             to have them exchanged (one 8-byte all-gather and a stream synchronisation per call).
 *                             Synchronous. A communicator carries one call at a time (not thread-safe per handle).
 *   ..._async                 device pointers, enqueued on the caller's stream, no synchronisation when src_counts is given.
 * NCCL is loaded at the first communicator with nranks > 1 (dlopen of libnccl.so.2, or the file SCTL_B200_NCCL_LIB names);
 * it is not a link-time dependency. An error on one rank before the collective leaves the other ranks waiting in it - the
 * reference aborts all ranks in that case (SCTL_ASSERT); callers should treat a non-zero return as fatal. */
SCTL_B200_API int sctl_b200_partition(int64_t n, int rank, int nranks, int64_t* begin, int64_t* end);
SCTL_B200_API int sctl_b200_comm_unique_id(void* id, int64_t id_bytes);
SCTL_B200_API int sctl_b200_comm_create(void** comm, int nranks, int rank, const void* nccl_unique_id);
SCTL_B200_API int sctl_b200_comm_destroy(void* comm);
SCTL_B200_API int sctl_b200_comm_info(void* comm, int* rank, int* nranks, int* device);
SCTL_B200_API int sctl_b200_nccl_version(int* version);
SCTL_B200_API int sctl_b200_eval_sharded(void* comm, int kernel_id, int dtype, int64_t Nt_local, int64_t Ns_local, const void* r_trg,
                           const void* r_src, const void* n_src, const void* v_src, void* v_trg, int digits, unsigned flags,
                           const int64_t* src_counts);
SCTL_B200_API int sctl_b200_eval_sharded_async(void* comm, int kernel_id, int dtype, int64_t Nt_local, int64_t Ns_local, const void* d_r_trg,
                                 const void* d_r_src, const void* d_n_src, const void* d_v_src, void* d_v_trg, int digits, unsigned flags,
                                 const int64_t* src_counts, void* cuda_stream);
/* Multi-rank particle registry: after sctl_b200_fmm_set_comm every rank hands ITS particles to the Set* calls (which become
 * collective): sources are all-gathered when they are SET, not on every evaluation as the reference's ring does
 * (fmm-wrapper.txx:537-558) - a solver that sets coordinates once and a new density per iteration moves 8*SrcDim bytes per
 * source per iteration and every eval_direct is a local launch sequence. Targets stay with their owner unless the ranks hold
 * uneven numbers (more than 5 % above the mean): then they are re-partitioned by the reference's formula (:504-510) with one
 * all-gather of coordinates at SetTrgCoord and one of the potentials per eval_direct (:560-561). eval_direct returns the
 * potentials of the rank's OWN targets in both cases. comm == NULL detaches. The communicator must outlive the registry. */
SCTL_B200_API int sctl_b200_fmm_set_comm(void* fmm, void* comm);

/* Measurement helpers (bench.py): number of kernel launches this library has issued from this process, and a
 * register-resident DFMA / FFMA chain that measures the CUDA-core FP64 / FP32 peak used as roofline
 * denominator (MEASURED_PEAKS.json has no such figure). Runs for about `seconds`; reports TFLOP/s counting
 * one FMA as 2 flops. */
SCTL_B200_API int64_t sctl_b200_launch_count(void);
SCTL_B200_API int sctl_b200_measure_peak(int dtype, double seconds, double* tflops);

/* Timing of the most recent eval / kernel_matrix issued by this thread through the HOST-pointer entry points:
 * milliseconds spent in (h2d, pack, main kernel, reduce, d2h), measured with CUDA events. */
SCTL_B200_API int sctl_b200_last_timing(double ms[5]);

/* Test hooks (tests/test_gpu_rsqrt.py). sctl_b200_newton_level: the rsqrt refinement level (0 seed, 1 one Newton step, 2 full
 * precision) the kernels use for `digits`; -1 for a bad dtype. No GPU needed. sctl_b200_test_rsqrt: y[i] = the device rsqrt of
 * that level applied to x[i] (HOST pointers), mode 1 = the masked form of the one-launch /
 * KernelMatrix kernels and mode 3 the branch-free form of repaired tiles (both: x > 0 ? 1/sqrt(x) : 0,
 * approx_rsqrt<digits>(x, x > 0) of vec.txx:361-364), mode 0 / 2 = the unmasked form
 * of the tiled hot loop with an arbitrary / worst-case low word in the seed (x <= 0 or subnormal gives inf or NaN there). */
SCTL_B200_API int sctl_b200_newton_level(int dtype, int digits);
SCTL_B200_API int sctl_b200_test_rsqrt(int dtype, int newton, int mode, int64_t n, const void* x, void* y);

SCTL_B200_API const char* sctl_b200_last_error(void);
SCTL_B200_API const char* sctl_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SCTL_B200_H_ */
