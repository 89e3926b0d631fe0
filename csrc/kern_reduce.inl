// Included by kern_<dtype>.cu with PDCC_T (element type), PDCC_NAME (table function) and
// PDCC_HAS_WIDE (16-bit floats: also build the fp32-output variant) defined. One
// translation unit per dtype keeps the nvcc jobs parallel.
#include "ll.cuh"

namespace pdcc {

PDCC_DECLARE_REDUCE_TABLE(PDCC_NAME) {
  using T = PDCC_T;
#define PDCC_PICK(W, R)                                                      \
  switch (ar_algo) {                                                         \
    case AR_ONESHOT_P2P:  return reduce_kernel<T, W, AR_ONESHOT_P2P, R>;     \
    case AR_TWOSHOT_P2P:  return reduce_kernel<T, W, AR_TWOSHOT_P2P, R>;     \
    case AR_ONESHOT_NVLS: return reduce_kernel<T, W, AR_ONESHOT_NVLS, R>;    \
    case AR_TWOSHOT_NVLS: return reduce_kernel<T, W, AR_TWOSHOT_NVLS, R>;    \
    case AR_LL:           return ll_reduce_kernel<T, W, R>;                  \
    default: return nullptr;                                                 \
  }
#if PDCC_HAS_WIDE
  if (wide) {
    if (rooted) return nullptr;
    PDCC_PICK(true, false)
  }
#else
  if (wide) return nullptr;
#endif
  if (rooted) { PDCC_PICK(false, true) }
  PDCC_PICK(false, false)
#undef PDCC_PICK
}

}  // namespace pdcc
