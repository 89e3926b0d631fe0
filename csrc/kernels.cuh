// K1-K7, K9 — the collectives, written as policies over ONE tile pipeline.
//
// Execution model. A call is cut into tiles; CTA b of every rank owns
// ring b (PDCC_RING_DEPTH slots at the same offset in every rank's symmetric block) and
// processes tile t of the call as global tile g = g0 + t of that ring, where g0 is the
// ring's lifetime tile count. Per tile a rank runs up to three stages and raises one
// monotonic flag per stage on its peers:
//
//   IN   user tensor -> slot(s), local or pushed to peers      -> READY(g)
//   MID  wait READY; cross-GPU read/reduce/(push)              -> MID(g)   [two-shot only]
//   OUT  wait MID;   slot -> user tensor                        -> DONE(g)
//
// and before touching ring position g % DEPTH again every rank waits for DONE(g-DEPTH)
// from all ranks. Because the counters never reset and every collective obeys the same
// slot discipline, consecutive calls of *different* collectives pipeline into each other
// with no end-of-kernel barrier and no host-side epoch (so launches are graph-capturable).
// IN(t+1) is issued before MID(t) and OUT(t-1) after it, so the NVLink phase of one tile
// overlaps the HBM staging of its neighbours across the CTAs of a GPU.
//
// Reference parity: one kernel per torch.distributed call site of main.py (lines cited
// at each policy). The reference itself has no kernels.
#pragma once

#include "reduce.cuh"

namespace pdcc {

constexpr int kThreads = 512;

struct KArgs {
  RankArgs r;           // this rank's view (real multi-GPU launch)
  const RankArgs* emu;  // device array indexed by blockIdx.y (single-GPU emulation)
  uint64_t bytes;       // message bytes per tensor (per list element for list ops)
  uint64_t tile_bytes;  // payload bytes per CTA per tile
  uint32_t stride;      // gather / all_gather-push: bytes between per-rank sub-slots
  int ntiles;
  int root;
  int redop;
  float scale;
};

struct Geo {
  uint64_t bytes, tile_bytes;
  int nb, b;
  __device__ __forceinline__ void range(int t, uint64_t& off, uint64_t& len) const {
    off = (uint64_t(t) * nb + b) * tile_bytes;
    len = off < bytes ? (bytes - off < tile_bytes ? bytes - off : tile_bytes) : 0;
  }
};

// Shared kernel prologue: pick this (virtual) rank's args, read the ring's tile counter.
struct Prologue {
  uint32_t g0;
  int poison;
  RankArgs args;  // emulation only: this virtual rank's args
};

__device__ __forceinline__ void init_ctx(const KArgs& k, Prologue& sh, Ctx& c) {
  const RankArgs* a = &k.r;
  if (k.emu != nullptr) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(k.emu + blockIdx.y);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&sh.args);
    for (int i = threadIdx.x; i < int(sizeof(RankArgs) / 4); i += blockDim.x) dst[i] = src[i];
    __syncthreads();
    a = &sh.args;
  }
  c.a = a;
  c.rank = a->comm.rank;
  c.n = a->comm.nranks;
  c.b = blockIdx.x;
  c.my_pad = reinterpret_cast<uint32_t*>(a->comm.base[c.rank] + kCtrlBytes);
  c.ctrl = reinterpret_cast<uint32_t*>(a->comm.base[c.rank]);
  c.poisoned = &sh.poison;
  c.deadline = globaltimer_ns() + uint64_t(a->comm.timeout_ms) * 1000000ull;
  if (threadIdx.x == 0) {
    sh.g0 = c.ctrl[kCtrlEpoch + c.b];
    sh.poison = 0;
  }
  __syncthreads();
  c.g0 = sh.g0;
}

template <class P>
__device__ __forceinline__ void run_pipeline(const KArgs& k) {
  __shared__ Prologue sh;
  Ctx c;
  init_ctx(k, sh, c);

  const uint32_t all = (1u << c.n) - 1u;
  const int T = k.ntiles;
  P p(k, c);

  // Ring position of tile g is free once every rank finished tile g - DEPTH. Flags start
  // at 0 and the compare is wrap-safe, so the first DEPTH tiles pass immediately.
  auto guard = [&](uint32_t g) { wait<false>(c, CH_DONE, g - PDCC_RING_DEPTH + 1u, all); };

  guard(c.g0);
  p.in(0);
  signal(c, CH_READY, c.g0 + 1u, p.ready_to());
  for (int t = 0; t < T; ++t) {
    const uint32_t g = c.g0 + uint32_t(t);
    if (t + 1 < T) {
      guard(g + 1u);
      p.in(t + 1);
      signal(c, CH_READY, g + 2u, p.ready_to());
    }
    wait(c, CH_READY, g + 1u, p.ready_from());
    p.mid(t);
    if constexpr (P::kHasOut) {
      signal(c, CH_MID, g + 1u, all);
      if (t >= 1) {
        wait(c, CH_MID, g, all);
        p.out(t - 1);
        signal_relaxed(c, CH_DONE, g, all);
      }
    } else {
      signal_relaxed(c, CH_DONE, g + 1u, all);
    }
  }
  if constexpr (P::kHasOut) {
    wait(c, CH_MID, c.g0 + uint32_t(T), all);
    p.out(T - 1);
    signal_relaxed(c, CH_DONE, c.g0 + uint32_t(T), all);
  }
  if (threadIdx.x == 0) c.ctrl[kCtrlEpoch + c.b] = c.g0 + uint32_t(T);
}

// ---------------------------------------------------------------------------------------
// K1/K2/K3: all_reduce (main.py:23), fused scale+cast (README.md:5), reduce (main.py:14)
// ---------------------------------------------------------------------------------------
enum ArAlgo : int { AR_ONESHOT_P2P, AR_TWOSHOT_P2P, AR_ONESHOT_NVLS, AR_TWOSHOT_NVLS, AR_LL };

template <typename T, bool WIDE, int OP, int ALGO, bool ROOTED>
struct ReducePolicy {
  using Tr = Traits<T>;
  static constexpr bool kHasOut = (ALGO == AR_TWOSHOT_P2P || ALGO == AR_TWOSHOT_NVLS);
  static constexpr int kMm = MmFor<T, OP>::kind;

  const KArgs& k;
  const Ctx& c;
  Geo geo;
  uint32_t all;

  __device__ ReducePolicy(const KArgs& k_, const Ctx& c_) : k(k_), c(c_) {
    geo = Geo{k.bytes, k.tile_bytes, int(gridDim.x), c.b};
    all = (1u << c.n) - 1u;
  }
  __device__ __forceinline__ uint32_t ready_to() const { return ROOTED ? (1u << k.root) : all; }
  __device__ __forceinline__ uint32_t ready_from() const {
    return ROOTED ? (c.rank == k.root ? all : 0u) : all;
  }

  __device__ __forceinline__ void in(int t) {
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    copy_bytes<LD_USER, 4>(c.mine(c.slot_off(c.g0 + t)), static_cast<const char*>(c.a->in) + off,
                           len);
  }

  // P2P: U units per thread per pass, every peer's 16-byte load issued before the first
  // combine, so U x n loads are in flight per thread (NVLink round trips are ~2 us: one
  // load in flight per thread would cap a GPU at ~150 GB/s). Peers are combined in rank
  // order 0..n-1 on every rank => bitwise identical results.
  template <int NP, int U, bool TWOSHOT>
  __device__ __forceinline__ void mid_p2p(size_t so, size_t lo, size_t hi, size_t nelem,
                                          size_t elem0) {
    const int tid = threadIdx.x, nt = blockDim.x;
    const bool scaled = k.scale != 1.0f;
    for (size_t j0 = lo + tid; j0 < hi; j0 += size_t(nt) * U) {
      uint4 v[U][NP];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t j = j0 + size_t(u) * nt;
        if (j < hi) {
#pragma unroll
          for (int p = 0; p < NP; ++p)
            if (p < c.n) v[u][p] = ld_sys_v4(c.peer(p, so) + (j << 4));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const size_t j = j0 + size_t(u) * nt;
        if (j < hi) {
          Acc<T> acc;
          acc.init(v[u][0]);
#pragma unroll
          for (int p = 1; p < NP; ++p)
            if (p < c.n) acc.template add<OP>(v[u][p]);
          if (scaled) acc.scale(k.scale);
          if constexpr (TWOSHOT) {
            const uint4 res = acc.pack();
#pragma unroll
            for (int p = 0; p < NP; ++p)
              if (p < c.n) st_v4(c.peer(p, so) + (j << 4), res);
          } else {
            const size_t e = j * Tr::K;
            const int valid = nelem - e < size_t(Tr::K) ? int(nelem - e) : Tr::K;
            store_unit<T, WIDE>(c.a->out, elem0 + e, acc, valid);
          }
        }
      }
    }
  }

  template <bool TWOSHOT>
  __device__ __forceinline__ void mid_p2p_dispatch(size_t so, size_t lo, size_t hi, size_t nelem,
                                                   size_t elem0) {
    if (c.n <= 2) mid_p2p<2, 8, TWOSHOT>(so, lo, hi, nelem, elem0);
    else if (c.n <= 4) mid_p2p<4, 4, TWOSHOT>(so, lo, hi, nelem, elem0);
    else mid_p2p<8, 2, TWOSHOT>(so, lo, hi, nelem, elem0);
  }

  // NVLS: the switch reduces; 8 multimem.ld_reduce in flight per thread.
  template <bool TWOSHOT>
  __device__ __forceinline__ void mid_nvls(size_t so, size_t lo, size_t hi, size_t nelem,
                                           size_t elem0) {
    if constexpr (kMm != MM_NONE) {
      constexpr int U = 8;
      const int tid = threadIdx.x, nt = blockDim.x;
      const bool scaled = k.scale != 1.0f;
      for (size_t j0 = lo + tid; j0 < hi; j0 += size_t(nt) * U) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const size_t j = j0 + size_t(u) * nt;
          if (j < hi) v[u] = multimem_ld_reduce_v4<kMm>(c.mc(so) + (j << 4));
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const size_t j = j0 + size_t(u) * nt;
          if (j < hi) {
            if constexpr (TWOSHOT) {
              uint4 r = v[u];
              if (scaled) {
                Acc<T> acc;
                acc.init(r);
                acc.scale(k.scale);
                r = acc.pack();
              }
              multimem_st_v4(c.mc(so) + (j << 4), r);
            } else {
              Acc<T> acc;
              acc.init(v[u]);
              if (scaled) acc.scale(k.scale);
              const size_t e = j * Tr::K;
              const int valid = nelem - e < size_t(Tr::K) ? int(nelem - e) : Tr::K;
              store_unit<T, WIDE>(c.a->out, elem0 + e, acc, valid);
            }
          }
        }
      }
    }
  }

  __device__ __forceinline__ void mid(int t) {
    if (ROOTED && c.rank != k.root) return;
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    const size_t so = c.slot_off(c.g0 + t);
    const size_t n16 = (len + 15) >> 4;
    const size_t nelem = len / sizeof(T);
    const size_t elem0 = off / sizeof(T);
    if constexpr (ALGO == AR_ONESHOT_P2P) {
      mid_p2p_dispatch<false>(so, 0, n16, nelem, elem0);
    } else if constexpr (ALGO == AR_ONESHOT_NVLS) {
      mid_nvls<false>(so, 0, n16, nelem, elem0);
    } else {
      // two-shot: this rank owns units [lo, hi) of the tile
      const size_t sub16 = (k.tile_bytes / c.n) >> 4;
      const size_t lo = size_t(c.rank) * sub16;
      size_t hi = lo + sub16;
      if (hi > n16) hi = n16;
      if (lo < hi) {
        if constexpr (ALGO == AR_TWOSHOT_P2P) mid_p2p_dispatch<true>(so, lo, hi, nelem, elem0);
        else mid_nvls<true>(so, lo, hi, nelem, elem0);
      }
    }
  }

  __device__ __forceinline__ void out(int t) {
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    const char* slot = c.mine(c.slot_off(c.g0 + t));
    if constexpr (!WIDE) {
      copy_bytes<LD_SYS, 4>(static_cast<char*>(c.a->out) + off, slot, len);
    } else {
      const int tid = threadIdx.x, nt = blockDim.x;
      const size_t n16 = (len + 15) >> 4;
      const size_t nelem = len / sizeof(T);
      const size_t elem0 = off / sizeof(T);
      for (size_t j = tid; j < n16; j += nt) {
        Acc<T> acc;
        acc.init(ld_sys_v4(slot + (j << 4)));
        const size_t e = j * Tr::K;
        const int valid = nelem - e < size_t(Tr::K) ? int(nelem - e) : Tr::K;
        store_unit<T, true>(c.a->out, elem0 + e, acc, valid);
      }
    }
  }
};

template <typename T, bool WIDE, int ALGO, bool ROOTED>
__global__ void __launch_bounds__(kThreads, 1) reduce_kernel(const __grid_constant__ KArgs k) {
  switch (k.redop) {
    case PDCC_SUM:
      run_pipeline<ReducePolicy<T, WIDE, PDCC_SUM, ALGO, ROOTED>>(k);
      break;
    case PDCC_PRODUCT:
      run_pipeline<ReducePolicy<T, WIDE, PDCC_PRODUCT, ALGO, ROOTED>>(k);
      break;
    case PDCC_MIN:
      run_pipeline<ReducePolicy<T, WIDE, PDCC_MIN, ALGO, ROOTED>>(k);
      break;
    default:
      run_pipeline<ReducePolicy<T, WIDE, PDCC_MAX, ALGO, ROOTED>>(k);
      break;
  }
}

// ---------------------------------------------------------------------------------------
// K4-K7 + K9: byte movers. List packing/unpacking is part of the kernel (pointer table in
// RankArgs::list), never a separate launch.
// ---------------------------------------------------------------------------------------
enum MoveKind : int {
  MV_BROADCAST_P2P,   // main.py:81  root stores the tile into every peer's slot
  MV_BROADCAST_NVLS,  // main.py:81  root multimem.st once, the switch replicates
  MV_SCATTER,         // main.py:37  root pushes list[j] into rank j's slot
  MV_GATHER,          // main.py:52  rank r pushes into sub-slot r of the root's slot
  MV_ALLGATHER_PULL,  // main.py:68  stage own shard, pull every peer's into list[p]
  MV_ALLGATHER_NVLS,  // main.py:68  multimem.st own shard into sub-slot r everywhere
  MV_BARRIER
};

template <int KIND>
struct MovePolicy {
  static constexpr bool kHasOut = false;
  const KArgs& k;
  const Ctx& c;
  Geo geo;
  uint32_t all;

  __device__ MovePolicy(const KArgs& k_, const Ctx& c_) : k(k_), c(c_) {
    geo = Geo{k.bytes, k.tile_bytes, int(gridDim.x), c.b};
    all = (1u << c.n) - 1u;
  }
  static constexpr bool kRootSends = KIND == MV_BROADCAST_P2P || KIND == MV_BROADCAST_NVLS ||
                                     KIND == MV_SCATTER;
  __device__ __forceinline__ uint32_t ready_to() const {
    if constexpr (kRootSends) return c.rank == k.root ? all : 0u;
    else if constexpr (KIND == MV_GATHER) return 1u << k.root;
    else return all;
  }
  __device__ __forceinline__ uint32_t ready_from() const {
    if constexpr (kRootSends) return c.rank == k.root ? 0u : (1u << k.root);
    else if constexpr (KIND == MV_GATHER) return c.rank == k.root ? all : 0u;
    else return all;
  }

  // Push `len` bytes of a user tensor to slot offset `dst_off` of the ranks in `mask`
  // (or of everyone through the multicast alias). A 16-byte aligned, 16-byte multiple
  // source streams straight from the user tensor; anything else is staged through
  // `stage_off` of this rank's own slot first.
  template <bool MC>
  __device__ __forceinline__ void push(const char* src, uint64_t len, size_t dst_off,
                                       size_t stage_off, uint32_t mask) {
    if (aligned16(src) && (len & 15u) == 0) {
      if constexpr (MC) push_mc<LD_USER, 8>(c, dst_off, src, len);
      else push_p2p<LD_USER, 8>(c, dst_off, src, len, mask);
    } else {
      copy_bytes<LD_USER, 4>(c.mine(stage_off), src, len);
      __syncthreads();
      if constexpr (MC) push_mc<LD_SYS, 8>(c, dst_off, c.mine(stage_off), len);
      else push_p2p<LD_SYS, 8>(c, dst_off, c.mine(stage_off), len, mask);
      __syncthreads();
    }
  }

  __device__ __forceinline__ void in(int t) {
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    const size_t so = c.slot_off(c.g0 + t);
    if constexpr (KIND == MV_BROADCAST_P2P || KIND == MV_BROADCAST_NVLS) {
      if (c.rank != k.root) return;
      const char* src = static_cast<const char*>(c.a->in) + off;
      if constexpr (KIND == MV_BROADCAST_NVLS) push<true>(src, len, so, so, all);
      else push<false>(src, len, so, so, all & ~(1u << k.root));
    } else if constexpr (KIND == MV_SCATTER) {
      if (c.rank != k.root) return;
      for (int j = 0; j < c.n; ++j) {
        const char* src = static_cast<const char*>(c.a->list[j]) + off;
        if (j == c.rank) copy_bytes<LD_USER, 4>(static_cast<char*>(c.a->out) + off, src, len);
        else push<false>(src, len, so, so, 1u << j);
      }
    } else if constexpr (KIND == MV_GATHER) {
      const char* src = static_cast<const char*>(c.a->in) + off;
      if (c.rank == k.root) {
        copy_bytes<LD_USER, 4>(static_cast<char*>(c.a->list[c.rank]) + off, src, len);
      } else {
        push<false>(src, len, so + size_t(c.rank) * k.stride, so, 1u << k.root);
      }
    } else if constexpr (KIND == MV_ALLGATHER_PULL) {
      const char* src = static_cast<const char*>(c.a->in) + off;
      copy_bytes<LD_USER, 4>(c.mine(so), src, len);
      copy_bytes<LD_USER, 4>(static_cast<char*>(c.a->list[c.rank]) + off, src, len);
    } else if constexpr (KIND == MV_ALLGATHER_NVLS) {
      const char* src = static_cast<const char*>(c.a->in) + off;
      const size_t sub = so + size_t(c.rank) * k.stride;
      push<true>(src, len, sub, sub, all);
    }
  }

  __device__ __forceinline__ void mid(int t) {
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    const size_t so = c.slot_off(c.g0 + t);
    if constexpr (KIND == MV_BROADCAST_P2P || KIND == MV_BROADCAST_NVLS || KIND == MV_SCATTER) {
      if (c.rank == k.root) return;
      copy_bytes<LD_SYS, 4>(static_cast<char*>(c.a->out) + off, c.mine(so), len);
    } else if constexpr (KIND == MV_GATHER) {
      if (c.rank != k.root) return;
      for (int j = 0; j < c.n; ++j)
        if (j != c.rank)
          copy_bytes<LD_SYS, 4>(static_cast<char*>(c.a->list[j]) + off,
                                c.mine(so + size_t(j) * k.stride), len);
    } else if constexpr (KIND == MV_ALLGATHER_PULL) {
      for (int i = 1; i < c.n; ++i) {
        const int p = (c.rank + i) % c.n;  // stagger peers so links load evenly
        copy_bytes<LD_SYS, 8>(static_cast<char*>(c.a->list[p]) + off, c.peer(p, so), len);
      }
    } else if constexpr (KIND == MV_ALLGATHER_NVLS) {
      for (int j = 0; j < c.n; ++j)
        copy_bytes<LD_SYS, 4>(static_cast<char*>(c.a->list[j]) + off,
                              c.mine(so + size_t(j) * k.stride), len);
    }
  }
  __device__ __forceinline__ void out(int) {}
};

template <int KIND>
__global__ void __launch_bounds__(kThreads, 1) move_kernel(const __grid_constant__ KArgs k) {
  run_pipeline<MovePolicy<KIND>>(k);
}

// per-dtype launch tables, one translation unit each (kern_<dtype>.cu)
typedef void (*KernelFn)(KArgs);
KernelFn reduce_kernel_for(int dtype, bool wide, int ar_algo, bool rooted);
#define PDCC_DECLARE_REDUCE_TABLE(NAME) KernelFn NAME(bool wide, int ar_algo, bool rooted)
PDCC_DECLARE_REDUCE_TABLE(reduce_table_f32);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_bf16);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_f16);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_f64);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_i32);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_i64);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_u8);

}  // namespace pdcc
