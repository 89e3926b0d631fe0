// all_reduce / reduce kernels for int64_t (see kern_reduce.inl).
#define PDCC_T int64_t
#define PDCC_NAME reduce_table_i64
#define PDCC_SYM_NAME sym_table_i64
#define PDCC_HAS_WIDE 0
#include "kern_reduce.inl"
