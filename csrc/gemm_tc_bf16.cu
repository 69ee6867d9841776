// bf16 instantiations of the tcgen05 GEMM main kernel, compiled as their own translation unit so that nvcc works on
// the two dtypes in parallel (see the note above the host section of gemm_tc.cu).
#define AB_GEMM_BF16_TU 1
#include "gemm_tc.cu"
