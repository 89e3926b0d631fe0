// Included by kern_<dtype>.cu with PDCC_T (element type), PDCC_NAME (table function) and
// PDCC_HAS_WIDE (16-bit floats: also build the fp32-output variant) defined. One
// translation unit per dtype keeps the nvcc jobs parallel.
#include "ll.cuh"
#include "sym.cuh"

namespace pdcc {

namespace {
// NVLS kernels only exist for the (type, op) pairs the switch implements (reduce.cuh MmFor).
template <typename T, bool W, int OP, int ALGO, bool R>
KernelFn nvls_or_null() {
  if constexpr (MmFor<T, OP>::kind == MM_NONE) return nullptr;
  else return reduce_kernel<T, W, OP, ALGO, R>;
}
template <typename T, bool W, int OP, bool R>
KernelFn pick_algo(int ar_algo) {
  switch (ar_algo) {
    case AR_ONESHOT_P2P:  return reduce_kernel<T, W, OP, AR_ONESHOT_P2P, R>;
    case AR_TWOSHOT_P2P:  return reduce_kernel<T, W, OP, AR_TWOSHOT_P2P, R>;
    case AR_ONESHOT_NVLS: return nvls_or_null<T, W, OP, AR_ONESHOT_NVLS, R>();
    case AR_TWOSHOT_NVLS: return nvls_or_null<T, W, OP, AR_TWOSHOT_NVLS, R>();
    case AR_LL:           return ll_reduce_kernel<T, W, OP, R>;
    default: return nullptr;
  }
}
template <typename T, bool W, bool R>
KernelFn pick_op(int redop, int ar_algo) {
  switch (redop) {
    case PDCC_SUM:     return pick_algo<T, W, PDCC_SUM, R>(ar_algo);
    case PDCC_PRODUCT: return W ? nullptr : pick_algo<T, false, PDCC_PRODUCT, R>(ar_algo);
    case PDCC_MIN:     return W ? nullptr : pick_algo<T, false, PDCC_MIN, R>(ar_algo);
    case PDCC_MAX:     return W ? nullptr : pick_algo<T, false, PDCC_MAX, R>(ar_algo);
    default: return nullptr;
  }
}
}  // namespace

namespace {
template <typename T, int OP>
SymKernelFn sym_pick(bool rooted) {
  if constexpr (MmFor<T, OP>::kind == MM_NONE) return nullptr;
  else return rooted ? sym_reduce_kernel<T, OP, true> : sym_reduce_kernel<T, OP, false>;
}
}  // namespace

PDCC_DECLARE_SYM_TABLE(PDCC_SYM_NAME) {
  using T = PDCC_T;
  switch (redop) {
    case PDCC_SUM: return sym_pick<T, PDCC_SUM>(rooted);
    case PDCC_MIN: return sym_pick<T, PDCC_MIN>(rooted);
    case PDCC_MAX: return sym_pick<T, PDCC_MAX>(rooted);
    default: return nullptr;  // PRODUCT: the switch cannot multiply
  }
}

PDCC_DECLARE_REDUCE_TABLE(PDCC_NAME) {
  using T = PDCC_T;
#if PDCC_HAS_WIDE
  // the fused bf16/f16 -> fp32 output only exists for the (scaled) sum of the DP average
  if (wide) return rooted ? nullptr : pick_op<T, true, false>(redop, ar_algo);
#else
  if (wide) return nullptr;
#endif
  return rooted ? pick_op<T, false, true>(redop, ar_algo) : pick_op<T, false, false>(redop, ar_algo);
}

}  // namespace pdcc
