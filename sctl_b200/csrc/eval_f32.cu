// FP32 instantiations of the direct-summation kernels (see eval_impl.cuh / kmat_impl.cuh).
#define SCTL_B200_REAL float
#define SCTL_B200_TABLE_FN eval_table_f32
#define SCTL_B200_CONTRACT_FN contract_normal_f32
#include "eval_tu.cuh"
