// all_reduce / reduce kernels for int32_t (see kern_reduce.inl).
#define PDCC_T int32_t
#define PDCC_NAME reduce_table_i32
#define PDCC_SYM_NAME sym_table_i32
#define PDCC_HAS_WIDE 0
#include "kern_reduce.inl"
