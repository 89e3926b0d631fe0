// all_reduce / reduce kernels for __half (see kern_reduce.inl).
#define PDCC_T __half
#define PDCC_NAME reduce_table_f16
#define PDCC_SYM_NAME sym_table_f16
#define PDCC_HAS_WIDE 1
#include "kern_reduce.inl"
