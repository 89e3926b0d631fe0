// all_reduce / reduce kernels for float (see kern_reduce.inl).
#define PDCC_T float
#define PDCC_NAME reduce_table_f32
#define PDCC_SYM_NAME sym_table_f32
#define PDCC_HAS_WIDE 0
#include "kern_reduce.inl"
