// Low-latency (LL) variants of all six collectives for small messages.
//
// The bandwidth kernels in kernels.cuh publish data with "write, system fence, raise flag"
// and consume it with "see flag, system fence, read": several microseconds of fences per
// tile, irrelevant at megabytes, dominant at bytes. Here the flag travels INSIDE the data:
// every 8 payload bytes are shipped as one 16-byte store {d0, flag, d1, flag}; the two
// 8-byte halves of a 16-byte store are each written atomically, so a receiver that reads
// both flags equal to the expected value knows both data words are the new ones. No fence,
// no separate signal, no round trip: the sender stores straight from the user tensor into
// the receiver's LL slot and the receiver polls its own memory. Cost: 2x the bytes on the
// wire, which is why plan_collective() only picks LL below a few tens of KB.
//
// LL slots are a dedicated region of the symmetric block (common.h) that only ever holds
// such packets, so stale contents are old flags (always smaller), never payload that could
// alias the flag. The flag of ring tile g is g + 1; the ring/DONE discipline is the one of
// the bandwidth kernels, so LL and non-LL calls interleave freely on the same rings.
//
// One tile per CTA covers at most kLLTileBytes = 512 threads x 16 B: each thread owns ONE
// 16-byte unit per tile, which keeps the unit in registers between push and combine.
//
// Reference parity: the 4-byte messages of main.py:14,23,37,52,68,81 are exactly this
// regime; torch/NCCL serves it with NCCL's LL protocol.
#pragma once

#include "kernels.cuh"

namespace pdcc {

constexpr uint32_t kLLTileBytes = kThreads * 16;  // 8 KiB of payload per CTA per tile

__device__ __forceinline__ void st_ll(char* dst, const uint4& u, uint32_t flag) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "r"(u.x), "r"(flag),
               "r"(u.y), "r"(flag)
               : "memory");
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 16), "r"(u.z),
               "r"(flag), "r"(u.w), "r"(flag)
               : "memory");
}

// Polls one unit (two packets) in this rank's own LL slot until both carry `flag`.
__device__ __forceinline__ uint4 ld_ll(const Ctx& c, const char* src, uint32_t flag) {
  uint4 a, b;
  uint32_t spins = 0;
  for (;;) {
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w)
                 : "l"(src)
                 : "memory");
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
                 : "l"(src + 16)
                 : "memory");
    if (a.y == flag && a.w == flag && b.y == flag && b.w == flag) break;
    if ((++spins & 0x3ffu) == 0 && (globaltimer_ns() > c.deadline || *c.poisoned)) {
      if (atomicCAS(c.ctrl + kCtrlErr, 0u, 2u) == 0u)
        c.ctrl[kCtrlErrInfo] = (7u << 28) | (flag & 0xffffffu);
      *c.poisoned = 1;
      break;
    }
  }
  return make_uint4(a.x, a.z, b.x, b.z);
}

enum LLKind : int { LL_ALL_REDUCE, LL_REDUCE, LL_BROADCAST, LL_SCATTER, LL_GATHER, LL_ALL_GATHER };

// One LL tile of collective KIND. T/OP/WIDE matter for the two reducing kinds only.
template <int KIND, typename T, int OP, bool WIDE>
__device__ __forceinline__ void ll_tile(const KArgs& k, const Ctx& c, int t) {
  using Tr = Traits<T>;
  const uint32_t g = c.g0 + uint32_t(t);
  const uint32_t flag = g + 1u;
  const int me = c.rank, n = c.n, root = k.root;
  const uint32_t j = threadIdx.x;  // my unit in this tile
  const uint64_t off = (uint64_t(t) * gridDim.x + c.b) * uint64_t(kLLTileBytes);
  const uint64_t len = off < k.bytes ? (k.bytes - off < kLLTileBytes ? k.bytes - off : kLLTileBytes) : 0;
  const bool active = (uint64_t(j) << 4) < len;
  const size_t lo = c.ll_off(g);
  const size_t sub = (kLLSlotBytes / size_t(n)) & ~size_t(31);  // per-source region
  const size_t pkt = size_t(j) * 32;

  if (!active) return;
  if constexpr (KIND == LL_ALL_REDUCE || KIND == LL_REDUCE) {
    const uint4 own = load_user_unit(static_cast<const char*>(c.a->in) + off, len, j);
    if constexpr (KIND == LL_ALL_REDUCE) {
      for (int i = 1; i < n; ++i) {
        const int p = (me + i) % n;
        st_ll(c.peer(p, lo + size_t(me) * sub) + pkt, own, flag);
      }
    } else if (me != root) {
      st_ll(c.peer(root, lo + size_t(me) * sub) + pkt, own, flag);
      return;
    }
    Acc<T> acc;
    for (int p = 0; p < n; ++p) {  // rank order => bitwise identical everywhere
      const uint4 v = p == me ? own : ld_ll(c, c.mine(lo + size_t(p) * sub) + pkt, flag);
      if (p == 0) acc.init(v);
      else acc.template add<OP>(v);
    }
    if (k.scale != 1.0f) acc.scale(k.scale);
    const size_t nelem = len / sizeof(T);
    const size_t e = size_t(j) * Tr::K;
    const int valid = nelem - e < size_t(Tr::K) ? int(nelem - e) : Tr::K;
    store_unit<T, WIDE>(c.a->out, off / sizeof(T) + e, acc, valid);
  } else if constexpr (KIND == LL_BROADCAST) {
    if (me == root) {
      const uint4 own = load_user_unit(static_cast<const char*>(c.a->in) + off, len, j);
      for (int i = 1; i < n; ++i) {
        const int p = (me + i) % n;
        st_ll(c.peer(p, lo + size_t(me) * sub) + pkt, own, flag);
      }
    } else {
      const uint4 v = ld_ll(c, c.mine(lo + size_t(root) * sub) + pkt, flag);
      store_user_unit(static_cast<char*>(c.a->out) + off, len, j, v);
    }
  } else if constexpr (KIND == LL_SCATTER) {
    if (me == root) {
      for (int i = 1; i < n; ++i) {
        const int p = (me + i) % n;
        const uint4 v = load_user_unit(static_cast<const char*>(c.a->list[p]) + off, len, j);
        st_ll(c.peer(p, lo + size_t(me) * sub) + pkt, v, flag);
      }
      const uint4 v = load_user_unit(static_cast<const char*>(c.a->list[me]) + off, len, j);
      store_user_unit(static_cast<char*>(c.a->out) + off, len, j, v);
    } else {
      const uint4 v = ld_ll(c, c.mine(lo + size_t(root) * sub) + pkt, flag);
      store_user_unit(static_cast<char*>(c.a->out) + off, len, j, v);
    }
  } else if constexpr (KIND == LL_GATHER) {
    const uint4 own = load_user_unit(static_cast<const char*>(c.a->in) + off, len, j);
    if (me != root) {
      st_ll(c.peer(root, lo + size_t(me) * sub) + pkt, own, flag);
    } else {
      store_user_unit(static_cast<char*>(c.a->list[me]) + off, len, j, own);
      for (int i = 1; i < n; ++i) {
        const int p = (me + i) % n;
        const uint4 v = ld_ll(c, c.mine(lo + size_t(p) * sub) + pkt, flag);
        store_user_unit(static_cast<char*>(c.a->list[p]) + off, len, j, v);
      }
    }
  } else {  // LL_ALL_GATHER
    const uint4 own = load_user_unit(static_cast<const char*>(c.a->in) + off, len, j);
    for (int i = 1; i < n; ++i) {
      const int p = (me + i) % n;
      st_ll(c.peer(p, lo + size_t(me) * sub) + pkt, own, flag);
    }
    store_user_unit(static_cast<char*>(c.a->list[me]) + off, len, j, own);
    for (int i = 1; i < n; ++i) {
      const int p = (me + i) % n;
      const uint4 v = ld_ll(c, c.mine(lo + size_t(p) * sub) + pkt, flag);
      store_user_unit(static_cast<char*>(c.a->list[p]) + off, len, j, v);
    }
  }
}

template <int KIND, typename T, int OP, bool WIDE>
__device__ __forceinline__ void run_ll(const KArgs& k) {
  __shared__ Prologue sh;
  Ctx c;
  init_ctx(k, sh, c);
  const uint32_t all = (1u << c.n) - 1u;
  for (int t = 0; t < k.ntiles; ++t) {
    const uint32_t g = c.g0 + uint32_t(t);
    wait_done_cta(c, g - PDCC_RING_DEPTH + 1u, all);  // LL slot g % DEPTH is free again
    ll_tile<KIND, T, OP, WIDE>(k, c, t);
    post_done_cta(c, g + 1u, all);
  }
  if (threadIdx.x == 0) c.ctrl[kCtrlEpoch + c.b] = c.g0 + uint32_t(k.ntiles);
}

template <typename T, bool WIDE, int OP, bool ROOTED>
__global__ void __launch_bounds__(kThreads, 1) ll_reduce_kernel(const __grid_constant__ KArgs k) {
  run_ll<ROOTED ? LL_REDUCE : LL_ALL_REDUCE, T, OP, WIDE>(k);
}

template <int KIND>
__global__ void __launch_bounds__(kThreads, 1) ll_move_kernel(const __grid_constant__ KArgs k) {
  run_ll<KIND, uint8_t, PDCC_SUM, false>(k);
}

}  // namespace pdcc
