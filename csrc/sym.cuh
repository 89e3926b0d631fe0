// Zero-copy collectives on symmetric memory.
//
// A tensor allocated with pdcc_sym_alloc sits at the same offset of a region that every
// rank has mapped (peer pointers) and that is bound to an NVSwitch multicast object (one
// address that reaches every copy). The kernels below therefore work on the user's tensor
// directly: no staging slot, no copy-in, no copy-out, no tile flags.
//
//   all_reduce  rank r owns 1/n of the tensor: v = multimem.ld_reduce(in)  (the switch adds
//               the n copies), optional scale, multimem.st(out, v) (the switch replicates).
//               Link bytes per GPU and direction: S (1 + 1/n); HBM: read S, write S.
//   reduce      the same load; the result goes to the root's copy with plain peer stores.
//   all_gather  multimem.st of my shard into list[rank] on every GPU.
//   broadcast   small: the root multimem.st's its tensor. Large: the root deals 1/(n-1) of
//               every tile to each other rank with peer stores into THEIR tensor, and each
//               of them re-sends its part with one multimem.st (one GPU cannot feed the
//               multicast engine at link rate, n-1 GPUs can).
//
// Synchronisation is two barriers per call and CTA (all ranks have entered: their tensor may
// be read / overwritten; all ranks are done: every store has landed), each ONE instruction
// on the sending side: multimem.red.add.u32 on an arrival counter that exists in every
// rank's block (K8). The counters only ever grow; their expected value is derived from a
// per-CTA call count kept in the block, so launches carry no host-side epoch.
//
// Contract (same as NCCL's registered buffers / NVSHMEM): in one call either every rank
// passes a tensor from the SAME symmetric allocation at the same offset, or none does.
//
// Reference parity: main.py:23 (all_reduce), :14 (reduce), :68 (all_gather), :81
// (broadcast) on tensors the library allocated; the reference has no counterpart of the
// allocator itself.
#pragma once

#include "kernels.cuh"

namespace pdcc {

struct SymArgs {
  DevComm comm;                    // the communicator's block: pads and ctrl words
  char* in_local;                  // my copy of the input tensor
  char* in_mc;                     // multicast alias of the input tensor
  char* out_local;                 // my copy of the output tensor (== in for in-place)
  char* out_mc;                    // multicast alias of the output tensor
  char* out_peer[PDCC_MAX_RANKS];  // every rank's copy of the output tensor
  uint64_t bytes;                  // message bytes (multiple of 16)
  uint64_t tile_bytes;             // broadcast two-phase: bytes per CTA per tile
  int root;
  float scale;
};

struct SymCtx {
  int rank, n, b;
  uint32_t* ctrl;
  uint32_t* my_pad;
  uint32_t* mc_pad;
  uint32_t epoch;  // zero-copy calls this CTA index has made before this one
  uint64_t deadline;
};

__device__ __forceinline__ void sym_init(const SymArgs& a, SymCtx& s) {
  s.rank = a.comm.rank;
  s.n = a.comm.nranks;
  s.b = blockIdx.x;
  s.ctrl = reinterpret_cast<uint32_t*>(a.comm.base[s.rank]);
  s.my_pad = reinterpret_cast<uint32_t*>(a.comm.base[s.rank] + kCtrlBytes);
  s.mc_pad = reinterpret_cast<uint32_t*>(a.comm.mc + kCtrlBytes);
  s.epoch = s.ctrl[kCtrlSymEpoch + s.b];
  s.deadline = globaltimer_ns() + uint64_t(a.comm.timeout_ms) * 1000000ull;
}

// All threads of the CTA call. phase 0 = "everyone has entered", 1 = "everyone is done".
// Thread 0: fence (publishes the CTA's writes, ordered before it by the barrier), ONE
// multimem.red.add that bumps the counter of CTA index b on every rank, then waits until
// its own copy has been bumped by all n ranks for this call, acquire-loads it and releases
// the CTA. A wait that outlives the budget records the error and gives up (no hang).
__device__ __forceinline__ void sym_barrier(const SymCtx& s, int phase) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const size_t word = kSymArriveWord + 2 * size_t(s.b) + size_t(phase);
    fence_acq_rel_sys();
    multimem_red_add_u32(s.mc_pad + word, 1u);
    const uint32_t target = (s.epoch + 1u) * uint32_t(s.n);
    const uint32_t* f = s.my_pad + word;
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_relaxed_sys(f) - target) < 0) {
      if ((++spins & 0x3ffu) == 0 && globaltimer_ns() > s.deadline) {
        if (atomicCAS(s.ctrl + kCtrlErr, 0u, 3u) == 0u)
          s.ctrl[kCtrlErrInfo] = (uint32_t(8 + phase) << 28) | (target & 0xffffffu);
        break;
      }
    }
    (void)ld_acquire_sys(f);
  }
  __syncthreads();
}

__device__ __forceinline__ void sym_finish(const SymCtx& s) {
  if (threadIdx.x == 0) s.ctrl[kCtrlSymEpoch + s.b] = s.epoch + 1u;
}

// ---------------------------------------------------------------------------------------
// all_reduce / reduce
// ---------------------------------------------------------------------------------------
template <typename T, int KMM, bool ROOTED>
__device__ __noinline__ void sym_reduce_units(const char* __restrict__ in_mc,
                                              char* __restrict__ out, size_t lo, size_t hi,
                                              float scale, int tid, int nt) {
  constexpr int U = 8;
  const bool scaled = scale != 1.0f;
  for (size_t j0 = lo + tid; j0 < hi; j0 += size_t(nt) * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = j0 + size_t(u) * nt;
      if (j < hi) v[u] = multimem_ld_reduce_v4<KMM>(in_mc + (j << 4));
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = j0 + size_t(u) * nt;
      if (j < hi) {
        uint4 r = v[u];
        if (scaled) {
          Acc<T> acc;
          acc.init(r);
          acc.scale(scale);
          r = acc.pack();
        }
        if constexpr (ROOTED) st_v4(out + (j << 4), r);      // the root's copy (peer store)
        else multimem_st_v4(out + (j << 4), r);              // every rank's copy
      }
    }
  }
}

template <typename T, int OP, bool ROOTED>
__global__ void __launch_bounds__(kThreads, 1) sym_reduce_kernel(const __grid_constant__ SymArgs a) {
  constexpr int kMm = MmFor<T, OP>::kind;
  if constexpr (kMm != MM_NONE) {
    SymCtx s;
    sym_init(a, s);
    sym_barrier(s, 0);
    // rank r owns units [r*per, (r+1)*per); CTA b a contiguous part of that
    const size_t units = size_t(a.bytes >> 4);
    const size_t per = (units + s.n - 1) / s.n;
    const size_t rlo = per * size_t(s.rank) < units ? per * size_t(s.rank) : units;
    const size_t rhi = rlo + per < units ? rlo + per : units;
    const size_t cper = (rhi - rlo + gridDim.x - 1) / gridDim.x;
    const size_t lo = rlo + cper * blockIdx.x < rhi ? rlo + cper * blockIdx.x : rhi;
    const size_t hi = lo + cper < rhi ? lo + cper : rhi;
    char* out = ROOTED ? a.out_peer[a.root] : a.out_mc;
    if (lo < hi) sym_reduce_units<T, kMm, ROOTED>(a.in_mc, out, lo, hi, a.scale, threadIdx.x, blockDim.x);
    sym_barrier(s, 1);
    sym_finish(s);
  }
}

#ifdef PDCC_SYM_MOVERS  // the type-less kernels are compiled once, in launch.cu
// ---------------------------------------------------------------------------------------
// all_gather: out_mc is the multicast alias of list[rank]
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1) sym_all_gather_kernel(const __grid_constant__ SymArgs a) {
  SymCtx s;
  sym_init(a, s);
  sym_barrier(s, 0);
  const size_t units = size_t(a.bytes >> 4);
  const size_t cper = (units + gridDim.x - 1) / gridDim.x;
  const size_t lo = cper * blockIdx.x < units ? cper * blockIdx.x : units;
  const size_t hi = lo + cper < units ? lo + cper : units;
  if (lo < hi) {
    if (s.n >= 4) {
      push16_mc<8>(a.out_mc + (lo << 4), a.in_local + (lo << 4), hi - lo, threadIdx.x, blockDim.x);
    } else {
      // two or three ranks: the multicast echo of my own shard costs 1/n of the ingress
      // (measured at 2 GPUs, 256 MiB: 405 us through the switch, 242 us with peer stores),
      // so store to every copy of list[rank] directly, mine included
      push16_p2p<8>(a.out_peer, s.n, (1u << s.n) - 1u, lo << 4, a.in_local + (lo << 4), hi - lo,
                    threadIdx.x, blockDim.x);
    }
  }
  sym_barrier(s, 1);
  sym_finish(s);
}

// ---------------------------------------------------------------------------------------
// broadcast, in place on a symmetric tensor
// ---------------------------------------------------------------------------------------
// Direct form: the root multicasts everything (one GPU feeds the switch: about half the
// link rate, but a single hop).
__global__ void __launch_bounds__(kThreads, 1) sym_broadcast_kernel(const __grid_constant__ SymArgs a) {
  SymCtx s;
  sym_init(a, s);
  sym_barrier(s, 0);
  if (s.rank == a.root) {
    const size_t units = size_t(a.bytes >> 4);
    const size_t cper = (units + gridDim.x - 1) / gridDim.x;
    const size_t lo = cper * blockIdx.x < units ? cper * blockIdx.x : units;
    const size_t hi = lo + cper < units ? lo + cper : units;
    if (lo < hi)
      push16_mc<8>(a.out_mc + (lo << 4), a.in_local + (lo << 4), hi - lo, threadIdx.x, blockDim.x);
  }
  sym_barrier(s, 1);
  sym_finish(s);
}

// Two-phase form. The CTA is cut into two halves that take alternate tiles, so that while
// one half waits in its fence the other keeps storing. Tile t of CTA b covers bytes
// [(t*grid + b) * tile, +tile) and is cut into n-1 parts, one per non-root rank.
//   root      part j of the tile -> rank j's tensor (peer stores); fence; progress[h] = k
//   rank j    waits for progress[h] >= k, then multimem.st's its part from its own tensor
// progress[h] is a per-(CTA, half) counter in every rank's pad region that only the root of
// the call writes; every rank keeps the running count in its ctrl words.
__global__ void __launch_bounds__(kThreads, 1) sym_broadcast2_kernel(const __grid_constant__ SymArgs a) {
  SymCtx s;
  sym_init(a, s);
  sym_barrier(s, 0);
  const int half = int(blockDim.x) >> 1;
  const int h = int(threadIdx.x) >= half ? 1 : 0;
  const int tid = int(threadIdx.x) - h * half;
  const int nt = h ? int(blockDim.x) - half : half;
  const int bar = 1 + h;
  const uint64_t step = uint64_t(gridDim.x) * a.tile_bytes;
  const int ntiles = int((a.bytes + step - 1) / step);
  const int own = s.n - 1;
  const uint32_t base = s.ctrl[kCtrlSymProg + 2 * s.b + h];
  const size_t sub = size_t(a.tile_bytes / own);  // tile_bytes is a multiple of 16 * own
  uint32_t k = 0;
  for (int t = h; t < ntiles; t += 2) {
    ++k;
    const uint64_t off = (uint64_t(t) * gridDim.x + blockIdx.x) * a.tile_bytes;
    const uint64_t len = off < a.bytes ? (a.bytes - off < a.tile_bytes ? a.bytes - off : a.tile_bytes) : 0;
    if (s.rank == a.root) {
      for (int i = 1; i < s.n; ++i) {
        const int j = (s.rank + i) % s.n;
        const int o = (j - a.root - 1 + s.n) % s.n;
        const uint64_t plo = uint64_t(o) * sub;
        if (len > plo) {
          const uint64_t plen = len - plo < sub ? len - plo : sub;
          char* d = a.out_peer[j] + off + plo;
          const char* src = a.in_local + off + plo;
          const size_t n16 = size_t(plen >> 4);
          for (size_t u = tid; u < n16; u += nt) st_v4(d + (u << 4), ld_stream_v4(src + (u << 4)));
        }
      }
      asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(nt) : "memory");
      if (tid == 0) {
        fence_acq_rel_sys();
        for (int j = 0; j < s.n; ++j) {
          if (j == a.root) continue;
          uint32_t* pad = reinterpret_cast<uint32_t*>(a.comm.base[j] + kCtrlBytes);
          st_relaxed_sys(pad + kSymProgWord + 2 * size_t(s.b) + size_t(h), base + k);
        }
      }
    } else {
      if (tid == 0) {
        const uint32_t* f = s.my_pad + kSymProgWord + 2 * size_t(s.b) + size_t(h);
        uint32_t spins = 0;
        while (static_cast<int32_t>(ld_relaxed_sys(f) - (base + k)) < 0) {
          if ((++spins & 0x3ffu) == 0 && globaltimer_ns() > s.deadline) {
            if (atomicCAS(s.ctrl + kCtrlErr, 0u, 3u) == 0u)
              s.ctrl[kCtrlErrInfo] = (10u << 28) | ((base + k) & 0xffffffu);
            break;
          }
        }
        (void)ld_acquire_sys(f);
      }
      asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(nt) : "memory");
      const int o = (s.rank - a.root - 1 + s.n) % s.n;
      const uint64_t plo = uint64_t(o) * sub;
      if (len > plo) {
        const uint64_t plen = len - plo < sub ? len - plo : sub;
        push16_mc<8>(a.out_mc + off + plo, a.out_local + off + plo, size_t(plen >> 4), tid, nt);
      }
    }
  }
  if (tid == 0) s.ctrl[kCtrlSymProg + 2 * s.b + h] = base + k;
  sym_barrier(s, 1);
  sym_finish(s);
}

#endif  // PDCC_SYM_MOVERS

typedef void (*SymKernelFn)(SymArgs);
SymKernelFn sym_reduce_kernel_for(int dtype, int redop, bool rooted);
#define PDCC_DECLARE_SYM_TABLE(NAME) SymKernelFn NAME(int redop, bool rooted)
PDCC_DECLARE_SYM_TABLE(sym_table_f32);
PDCC_DECLARE_SYM_TABLE(sym_table_bf16);
PDCC_DECLARE_SYM_TABLE(sym_table_f16);
PDCC_DECLARE_SYM_TABLE(sym_table_f64);
PDCC_DECLARE_SYM_TABLE(sym_table_i32);
PDCC_DECLARE_SYM_TABLE(sym_table_i64);
PDCC_DECLARE_SYM_TABLE(sym_table_u8);

}  // namespace pdcc
