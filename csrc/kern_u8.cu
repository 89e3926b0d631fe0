// all_reduce / reduce kernels for uint8_t (see kern_reduce.inl).
#define PDCC_T uint8_t
#define PDCC_NAME reduce_table_u8
#define PDCC_SYM_NAME sym_table_u8
#define PDCC_HAS_WIDE 0
#include "kern_reduce.inl"
