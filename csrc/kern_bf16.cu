// all_reduce / reduce kernels for __nv_bfloat16 (see kern_reduce.inl).
#define PDCC_T __nv_bfloat16
#define PDCC_NAME reduce_table_bf16
#define PDCC_SYM_NAME sym_table_bf16
#define PDCC_HAS_WIDE 1
#include "kern_reduce.inl"
