// K8 — device-side primitives: system-scope loads/stores, NVLS multimem ops, the
// per-CTA flag protocol (signal / wait with a wall-clock budget), and vector copy loops.
//
// Everything here is sm_100a PTX. Reference parity: the reference has no device code
// (SURVEY.md §2.3); these replace what NCCL's kernels do under torch's ProcessGroupNCCL.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include "common.h"

namespace pdcc {

enum Chan : int { CH_READY = 0, CH_MID = 1, CH_DONE = 2 };

// ---------------------------------------------------------------- memory ops ---------

__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() {
  asm volatile("fence.acq_rel.sys;" ::: "memory");
}

// Data written by a peer (into my slot) or living on a peer: never serve it from L1.
// Peer apertures are L1-cached / L2-bypassed on Blackwell (B300_MICROARCH.md, NVLink),
// and a ring slot is re-read at the same address every PDCC_RING_DEPTH tiles.
__device__ __forceinline__ uint4 ld_sys_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
// User tensors: read once, keep them out of L1.
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_v4(void* p, uint4 v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// NVLS: one store replicated by the switch into every bound device's block.
__device__ __forceinline__ void multimem_st_v4(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc),
               "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// NVLS: one load whose value is reduced inside the switch over every bound device.
// KIND encodes type+op; only the combinations ptxas accepts for sm_100a exist
// (no .mul at all, no min/max on .f32).
enum MmKind : int {
  MM_ADD_F32, MM_ADD_BF16, MM_ADD_F16, MM_ADD_F64,
  MM_MIN_BF16, MM_MAX_BF16, MM_MIN_F16, MM_MAX_F16,
  MM_ADD_S32, MM_MIN_S32, MM_MAX_S32, MM_MIN_S64, MM_MAX_S64, MM_ADD_U64,
  MM_NONE
};

template <int KIND>
__device__ __forceinline__ uint4 multimem_ld_reduce_v4(const void* mc) {
  uint4 v;
#define PDCC_MM_V4(INSTR)                                                                  \
  asm volatile(INSTR " {%0,%1,%2,%3}, [%4];"                                               \
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)                                \
               : "l"(mc)                                                                   \
               : "memory")
  if constexpr (KIND == MM_ADD_F32) {
    PDCC_MM_V4("multimem.ld_reduce.relaxed.sys.global.add.v4.f32");
  } else if constexpr (KIND == MM_ADD_BF16) {
    PDCC_MM_V4("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2");
  } else if constexpr (KIND == MM_ADD_F16) {
    PDCC_MM_V4("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2");
  } else if constexpr (KIND == MM_MIN_BF16) {
    PDCC_MM_V4("multimem.ld_reduce.relaxed.sys.global.min.v4.bf16x2");
  } else if constexpr (KIND == MM_MAX_BF16) {
    PDCC_MM_V4("multimem.ld_reduce.relaxed.sys.global.max.v4.bf16x2");
  } else if constexpr (KIND == MM_MIN_F16) {
    PDCC_MM_V4("multimem.ld_reduce.relaxed.sys.global.min.v4.f16x2");
  } else if constexpr (KIND == MM_MAX_F16) {
    PDCC_MM_V4("multimem.ld_reduce.relaxed.sys.global.max.v4.f16x2");
  } else if constexpr (KIND == MM_ADD_F64 || KIND == MM_MIN_S64 || KIND == MM_MAX_S64 ||
                       KIND == MM_ADD_U64) {
    // 64-bit forms have no .v4: two scalar loads cover the 16-byte unit.
    uint64_t a, b;
    const char* q = static_cast<const char*>(mc);
#define PDCC_MM_S64(INSTR)                                                                 \
  asm volatile(INSTR " %0, [%1];" : "=l"(a) : "l"(q) : "memory");                          \
  asm volatile(INSTR " %0, [%1];" : "=l"(b) : "l"(q + 8) : "memory")
    if constexpr (KIND == MM_ADD_F64) {
      PDCC_MM_S64("multimem.ld_reduce.relaxed.sys.global.add.f64");
    } else if constexpr (KIND == MM_MIN_S64) {
      PDCC_MM_S64("multimem.ld_reduce.relaxed.sys.global.min.s64");
    } else if constexpr (KIND == MM_MAX_S64) {
      PDCC_MM_S64("multimem.ld_reduce.relaxed.sys.global.max.s64");
    } else {
      PDCC_MM_S64("multimem.ld_reduce.relaxed.sys.global.add.u64");
    }
#undef PDCC_MM_S64
    v.x = static_cast<uint32_t>(a);
    v.y = static_cast<uint32_t>(a >> 32);
    v.z = static_cast<uint32_t>(b);
    v.w = static_cast<uint32_t>(b >> 32);
  } else {
    // 32-bit integer forms have no .v4 either: four scalar loads.
    const char* q = static_cast<const char*>(mc);
#define PDCC_MM_S32(INSTR)                                                                 \
  asm volatile(INSTR " %0, [%1];" : "=r"(v.x) : "l"(q) : "memory");                        \
  asm volatile(INSTR " %0, [%1];" : "=r"(v.y) : "l"(q + 4) : "memory");                    \
  asm volatile(INSTR " %0, [%1];" : "=r"(v.z) : "l"(q + 8) : "memory");                    \
  asm volatile(INSTR " %0, [%1];" : "=r"(v.w) : "l"(q + 12) : "memory")
    if constexpr (KIND == MM_ADD_S32) {
      PDCC_MM_S32("multimem.ld_reduce.relaxed.sys.global.add.s32");
    } else if constexpr (KIND == MM_MIN_S32) {
      PDCC_MM_S32("multimem.ld_reduce.relaxed.sys.global.min.s32");
    } else {
      static_assert(KIND == MM_MAX_S32, "unknown multimem kind");
      PDCC_MM_S32("multimem.ld_reduce.relaxed.sys.global.max.s32");
    }
#undef PDCC_MM_S32
  }
#undef PDCC_MM_V4
  return v;
}

// ---------------------------------------------------------------- block context ------

struct Ctx {
  const RankArgs* a;
  int rank, n;
  int b;             // CTA index == ring index
  uint32_t g0;       // tiles this ring had processed before this call
  uint32_t* my_pad;  // flags peers write into my block
  uint32_t* ctrl;
  uint64_t deadline;  // globaltimer value after which waits give up
  int* poisoned;      // smem: a wait already timed out in this CTA
  // Slot geometry of this call: the ring of CTA b is cut into `depth` slots of `slot_stride`
  // bytes; tile t of the call lives in slot t % depth.
  size_t ring_off;
  uint32_t slot_stride;
  int depth;
  // The worker group this thread belongs to (run_pipeline gives every pipeline stage its own
  // group): stage code strides its loops by (tid, nt). `bar` is the group's index + 1.
  int tid, nt, bar;

  __device__ __forceinline__ size_t slot_off(int t) const {
    return ring_off + size_t(t % depth) * slot_stride;
  }
  __device__ __forceinline__ size_t ll_off(uint32_t g) const {
    return kHeaderBytes + (size_t(b) * PDCC_RING_DEPTH + (g % PDCC_RING_DEPTH)) * kLLSlotBytes;
  }
  __device__ __forceinline__ char* peer(int p, size_t off) const { return a->comm.base[p] + off; }
  __device__ __forceinline__ char* mine(size_t off) const { return a->comm.base[rank] + off; }
  __device__ __forceinline__ char* mc(size_t off) const { return a->comm.mc + off; }
};

__device__ __forceinline__ int flag_index(int chan, int b, int src) {
  return (chan * PDCC_MAX_BLOCKS + b) * PDCC_MAX_RANKS + src;
}

__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// One store that the switch replicates into every rank's pad (the flag word of (chan, b,
// my rank) has exactly one writer, so a store is enough; no atomic needed).
__device__ __forceinline__ void multimem_st_u32(uint32_t* mc, uint32_t v) {
  asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}
// One-instruction arrive on a counter that exists in every rank's block (K8).
__device__ __forceinline__ void multimem_red_add_u32(uint32_t* mc, uint32_t v) {
  asm volatile("multimem.red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}

__device__ __forceinline__ void record_timeout(const Ctx& c, uint32_t code, uint32_t info) {
  if (atomicCAS(c.ctrl + kCtrlErr, 0u, code) == 0u) c.ctrl[kCtrlErrInfo] = info;
  *reinterpret_cast<volatile int*>(c.poisoned) = 1;
}

// Polls flag[chan][b][src] in MY pad until it reaches val (wrap-safe compare). Returns when
// it has, when the wall-clock budget is spent (error recorded, CTA poisoned) or when another
// group of this CTA has already been poisoned: a mismatched collective never hangs the GPU
// (SURVEY.md §5.2 / §7.3 item 9).
__device__ __forceinline__ void spin_until(const Ctx& c, int chan, int src, uint32_t val) {
  const uint32_t* f = c.my_pad + flag_index(chan, c.b, src);
  uint32_t spins = 0;
  while (static_cast<int32_t>(ld_relaxed_sys(f) - val) < 0) {
    if ((++spins & 0x3ffu) == 0) {
      if (*reinterpret_cast<volatile int*>(c.poisoned)) return;
      if (globaltimer_ns() > c.deadline) {
        record_timeout(c, 1u, (uint32_t(chan) << 28) | (uint32_t(src) << 24) | (val & 0xffffffu));
        return;
      }
    }
  }
}

// ---- CTA-scope hand-offs between the worker warps and the two control warps --------------
__device__ __forceinline__ uint32_t ld_acquire_cta_shared(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.cta.shared.u32 %0, [%1];"
               : "=r"(v)
               : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(p)))
               : "memory");
  return v;
}
__device__ __forceinline__ void st_release_cta_shared(uint32_t* p, uint32_t v) {
  asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(
                   static_cast<uint32_t>(__cvta_generic_to_shared(p))),
               "r"(v)
               : "memory");
}
__device__ __forceinline__ void red_release_cta_shared_add(uint32_t* p, uint32_t v) {
  asm volatile("red.release.cta.shared.add.u32 [%0], %1;" ::"r"(
                   static_cast<uint32_t>(__cvta_generic_to_shared(p))),
               "r"(v)
               : "memory");
}

// Raises flag[chan][b][my rank] = val on every rank in mask (relaxed stores; the caller has
// issued the fence that makes them the tail of a release pattern).
__device__ __forceinline__ void post_flags(const Ctx& c, int chan, uint32_t val, uint32_t mask) {
  for (int p = 0; p < c.n; ++p) {
    if ((mask >> p) & 1u) {
      uint32_t* pad = reinterpret_cast<uint32_t*>(c.a->comm.base[p] + kCtrlBytes);
      st_relaxed_sys(pad + flag_index(chan, c.b, c.rank), val);
    }
  }
}

// ---- CTA-wide forms used by the LL kernels (one unit per thread, no warp groups) ---------
// The slot guard is an acquire load per polled flag, DONE is a release store: the LL data
// path itself needs neither (the flag travels inside the data), these only order the reuse
// of an LL slot PDCC_RING_DEPTH tiles later.
__device__ __forceinline__ void wait_done_cta(const Ctx& c, uint32_t val, uint32_t mask) {
  const int t = threadIdx.x;
  if (t < c.n && ((mask >> t) & 1u) && !*c.poisoned) {
    const uint32_t* f = c.my_pad + flag_index(CH_DONE, c.b, t);
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(f) - val) < 0) {
      if ((++spins & 0x3ffu) == 0 && globaltimer_ns() > c.deadline) {
        record_timeout(c, 1u, (uint32_t(CH_DONE) << 28) | (uint32_t(t) << 24) | (val & 0xffffffu));
        break;
      }
    }
  }
  __syncthreads();
}
__device__ __forceinline__ void post_done_cta(const Ctx& c, uint32_t val, uint32_t mask) {
  __syncthreads();
  const int t = threadIdx.x;
  if (t < c.n && ((mask >> t) & 1u)) {
    uint32_t* pad = reinterpret_cast<uint32_t*>(c.a->comm.base[t] + kCtrlBytes);
    st_release_sys(pad + flag_index(CH_DONE, c.b, c.rank), val);
  }
}

// ---------------------------------------------------------------- copies -------------

enum Ld : int { LD_USER = 0, LD_SYS = 1 };

// LD_SYS marks a source that peers wrote (a slot of this rank's block). It used to be a
// strong system-scope load; those measured 2-3x slower than weak ones, and are not needed:
// every consumer runs after a flag wait whose fence.acq_rel.sys already invalidated L1,
// and slot loads never allocate in L1. Both kinds are the same instruction now; the tag
// stays for readability and for the 4-byte fallback path.
template <int LD>
__device__ __forceinline__ uint4 ld16(const void* p) {
  return ld_stream_v4(p);
}

__device__ __forceinline__ bool aligned16(const void* p) {
  return (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}

// The hot loops live in __noinline__ functions: each gets its own register allocation with
// only its few arguments live, instead of competing with everything the (heavily inlined,
// four-ops-in-one) kernel body keeps alive. The call costs nothing next to a tile.

// 16-byte vector copy of n16 units, UNROLL independent loads in flight per thread. Every
// pass issues all of its (predicated) loads before its first store, the last partial pass
// included: a range shorter than one pass still has n16 / nt loads in flight per thread
// instead of degrading to one load -> one store round trips.
template <int UNROLL>
__device__ __noinline__ void copy16(char* __restrict__ d, const char* __restrict__ s, size_t n16,
                                    int tid, int nt) {
  for (size_t i = tid; i < n16; i += size_t(nt) * UNROLL) {
    uint4 v[UNROLL];
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) {
      const size_t j = i + size_t(k) * nt;
      if (j < n16) v[k] = ld_stream_v4(s + (j << 4));
    }
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) {
      const size_t j = i + size_t(k) * nt;
      if (j < n16) st_v4(d + (j << 4), v[k]);
    }
  }
}

// CTA-wide copy of n bytes. Fast path: both sides 16-byte aligned. Anything else degrades
// to 4- or 1-byte words; that only happens for odd user views (slot addresses are aligned).
template <int LD, int UNROLL>
__device__ __forceinline__ void copy_bytes(const Ctx& c, void* dst, const void* src, size_t n) {
  const int tid = c.tid, nt = c.nt;
  char* d = static_cast<char*>(dst);
  const char* s = static_cast<const char*>(src);
  if (aligned16(d) && aligned16(s)) {
    const size_t n16 = n >> 4;
    copy16<UNROLL>(d, s, n16, tid, nt);
    for (size_t i = (n16 << 4) + tid; i < n; i += nt) d[i] = s[i];
  } else if (((reinterpret_cast<uintptr_t>(d) | reinterpret_cast<uintptr_t>(s)) & 3u) == 0) {
    const size_t n4 = n >> 2;
    const uint32_t* s4 = reinterpret_cast<const uint32_t*>(s);
    uint32_t* d4 = reinterpret_cast<uint32_t*>(d);
    for (size_t i = tid; i < n4; i += nt) d4[i] = s4[i];
    for (size_t i = (n4 << 2) + tid; i < n; i += nt) d[i] = s[i];
  } else {
    for (size_t i = tid; i < n; i += nt) d[i] = s[i];
  }
}

// CTA-wide push of n bytes from `src` (16-byte aligned) into the same slot offset of every
// rank in `mask` (P2P stores). n is rounded up to 16 (slots have room).
template <int UNROLL>
__device__ __noinline__ void push16_p2p(char* const* __restrict__ base, int nranks, uint32_t mask,
                                        size_t off, const char* __restrict__ s, size_t n16,
                                        int tid, int nt) {
  if (mask == 0) return;  // a group of one has nobody to push to
  if ((mask & (mask - 1u)) == 0) {  // one target: scatter, gather, reduce-scatter phase
    copy16<UNROLL>(base[__ffs(int(mask)) - 1] + off, s, n16, tid, nt);
    return;
  }
  // several targets: every loaded unit is stored to each of them
  constexpr int U2 = 4;
  for (size_t i = tid; i < n16; i += size_t(nt) * U2) {
    uint4 v[U2];
#pragma unroll
    for (int k = 0; k < U2; ++k) {
      const size_t j = i + size_t(k) * nt;
      if (j < n16) v[k] = ld_stream_v4(s + (j << 4));
    }
#pragma unroll
    for (int p = 0; p < PDCC_MAX_RANKS; ++p) {
      if (p < nranks && ((mask >> p) & 1u)) {
        char* d = base[p] + off;
#pragma unroll
        for (int k = 0; k < U2; ++k) {
          const size_t j = i + size_t(k) * nt;
          if (j < n16) st_v4(d + (j << 4), v[k]);
        }
      }
    }
  }
}

// Dealing loop of the two-shot / two-phase collectives: the tile [src, src + n16 units) is
// cut into parts of `sub16` units, part o belongs to rank owner[o] and goes to
// base[owner[o]] + dst_off + (position inside the part). ONE streaming loop over the whole
// tile (full passes of UNROLL loads in flight) instead of one short loop per owner; the
// part owned by this rank itself (skip) is not sent.
template <int UNROLL>
__device__ __noinline__ void deal16(char* const* __restrict__ base, int first, int nranks,
                                    int skip, size_t dst_off, size_t part_stride,
                                    const char* __restrict__ s, size_t n16, size_t sub16, int tid,
                                    int nt) {
  // j / sub16 as a multiply-high: exact while j * sub16 < 2^32 (a tile is at most a few
  // thousand units, a part a fraction of that), and free of the division's registers
  const uint32_t n32 = uint32_t(n16), s32 = uint32_t(sub16);
  const uint32_t magic = uint32_t(((uint64_t(1) << 32) + s32 - 1) / s32);
  for (uint32_t i = uint32_t(tid); i < n32; i += uint32_t(nt) * UNROLL) {
    uint4 v[UNROLL];
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) {
      const uint32_t j = i + uint32_t(k) * uint32_t(nt);
      if (j < n32) v[k] = ld_stream_v4(s + (size_t(j) << 4));
    }
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) {
      const uint32_t j = i + uint32_t(k) * uint32_t(nt);
      if (j < n32) {
        const uint32_t o = s32 == 1 ? j : __umulhi(j, magic);
        int r = first + int(o);
        if (r >= nranks) r -= nranks;  // first <= nranks and o < nranks
        if (r >= nranks) r -= nranks;
        if (r != skip)
          st_v4(base[r] + dst_off + size_t(o) * part_stride + (size_t(j - o * s32) << 4), v[k]);
      }
    }
  }
}

template <int LD, int UNROLL>
__device__ __forceinline__ void push_p2p(const Ctx& c, size_t off, const void* src, size_t n,
                                         uint32_t mask) {
  push16_p2p<UNROLL>(c.a->comm.base, c.n, mask, off, static_cast<const char*>(src), (n + 15) >> 4,
                     c.tid, c.nt);
}

// Same through the multicast alias: one multimem.st per 16 bytes reaches every rank.
template <int UNROLL>
__device__ __noinline__ void push16_mc(char* __restrict__ d, const char* __restrict__ s,
                                       size_t n16, int tid, int nt) {
  for (size_t i = tid; i < n16; i += size_t(nt) * UNROLL) {
    uint4 v[UNROLL];
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) {
      const size_t j = i + size_t(k) * nt;
      if (j < n16) v[k] = ld_stream_v4(s + (j << 4));
    }
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) {
      const size_t j = i + size_t(k) * nt;
      if (j < n16) multimem_st_v4(d + (j << 4), v[k]);
    }
  }
}

template <int LD, int UNROLL>
__device__ __forceinline__ void push_mc(const Ctx& c, size_t off, const void* src, size_t n) {
  push16_mc<UNROLL>(c.mc(off), static_cast<const char*>(src), (n + 15) >> 4, c.tid, c.nt);
}

}  // namespace pdcc
