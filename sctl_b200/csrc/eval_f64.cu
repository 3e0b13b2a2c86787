// FP64 instantiations of the direct-summation kernels (see eval_impl.cuh / kmat_impl.cuh).
#define SCTL_B200_REAL double
#define SCTL_B200_TABLE_FN eval_table_f64
#define SCTL_B200_CONTRACT_FN contract_normal_f64
#include "eval_tu.cuh"
