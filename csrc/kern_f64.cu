// all_reduce / reduce kernels for double (see kern_reduce.inl).
#define PDCC_T double
#define PDCC_NAME reduce_table_f64
#define PDCC_SYM_NAME sym_table_f64
#define PDCC_HAS_WIDE 0
#include "kern_reduce.inl"
