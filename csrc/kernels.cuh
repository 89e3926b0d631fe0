// K1-K7, K9 — the collectives, written as policies over ONE tile pipeline.
//
// Execution model. A call is cut into tiles; CTA b of every rank owns
// ring b (the same offset in every rank's symmetric block, cut into `depth` slots for this
// call) and processes tile t of the call as global tile g = g0 + t of that ring, where g0
// is the ring's lifetime tile count. A tile passes through up to three stages and every
// stage raises one monotonic flag on the peers that consume it:
//
//   IN   user tensor -> slot(s): pushed to peers, or copied locally   -> READY(g)
//   MID  wait READY; combine / re-send / copy-out                     -> MID(g)  [3-stage]
//   OUT  wait MID;   slot -> user tensor                              -> DONE(g)
//
// and before touching slot t % depth again a rank waits for DONE(t - depth) from all
// ranks. Counters never reset and every collective obeys the same slot discipline, so
// consecutive calls of different collectives follow each other with no end-of-kernel
// barrier and no host-side epoch (launches are CUDA-graph capturable).
//
// run_pipeline() gives every stage that has work on a rank its OWN group of worker warps and
// lets the groups run their tile loops independently: the only coupling between stages — on
// this rank and across ranks alike — is the flags. Two control warps keep every flag wait and
// every system-scope fence off the data path (admitter / publisher, see run_pipeline below).
// Tiles are small (64 KiB per CTA for large messages) and rings are deep (8-16 slots), so a
// 256 MiB message is 28 tiles per CTA and IN runs up to `depth` tiles ahead of OUT.
//
// Reference parity: one kernel per torch.distributed call site of main.py (lines cited
// at each policy). The reference itself has no kernels.
#pragma once

#include "reduce.cuh"

namespace pdcc {

constexpr int kThreads = 512;     // LL kernels: one 16-byte unit per thread
constexpr int kMaxThreads = 512;   // pipeline kernels: 2 control warps + 14 worker warps (128 registers each)

struct KArgs {
  RankArgs r;           // this rank's view (real multi-GPU launch)
  const RankArgs* emu;  // device array indexed by blockIdx.y (single-GPU emulation)
  uint64_t bytes;       // message bytes per tensor (per list element for list ops)
  uint64_t tile_bytes;  // payload bytes per CTA per tile
  uint32_t stride;      // gather / all_gather-push: bytes between per-rank sub-slots
  uint32_t slot_stride; // bytes between the slots of a ring (this call's geometry)
  int depth;            // slots per ring
  int ntiles;
  int root;
  float scale;
  int trace;            // PDCC_TRACE: CTA 0 records a timeline into its ctrl words
};

__device__ __forceinline__ void trace_event(const Ctx& c, uint32_t kind, uint32_t tile, uint64_t t0,
                                            uint64_t t1, uint32_t aux0, uint32_t aux1) {
  const uint32_t i = atomicAdd(c.ctrl + kCtrlTrace, 1u);
  if (i < uint32_t(kCtrlTraceCap)) {
    uint32_t* e = c.ctrl + kCtrlTrace + 4 + 8 * i;
    e[0] = kind; e[1] = tile;
    e[2] = uint32_t(t0); e[3] = uint32_t(t0 >> 32);
    e[4] = uint32_t(t1); e[5] = uint32_t(t1 >> 32);
    e[6] = aux0; e[7] = aux1;
  }
}

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
  RankArgs args;  // this (virtual) rank's args: shared memory => LDS, not generic loads
};

__device__ __forceinline__ void init_ctx(const KArgs& k, Prologue& sh, Ctx& c) {
  {
    uint32_t* dst = reinterpret_cast<uint32_t*>(&sh.args);
    constexpr int kWords = int(sizeof(RankArgs) / 4);
    if (k.emu != nullptr) {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(k.emu + blockIdx.y);
      for (int i = threadIdx.x; i < kWords; i += blockDim.x) dst[i] = src[i];
    } else {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(&k.r);
      for (int i = threadIdx.x; i < kWords; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
  }
  const RankArgs* a = &sh.args;
  c.a = a;
  c.rank = a->comm.rank;
  c.n = a->comm.nranks;
  c.b = blockIdx.x;
  c.tid = threadIdx.x;
  c.nt = blockDim.x;
  c.bar = 0;
  c.ring_off = kDataOffset + size_t(c.b) * a->comm.ring_bytes;
  c.slot_stride = k.slot_stride;
  c.depth = k.depth > 0 ? k.depth : 1;
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

// Warps per stage from the policy's relative weights. A stage with no work on this rank
// (weight 0) gets no warps at all: its flag is raised on behalf of the stage before it.
__device__ __forceinline__ void split_warps(int nwarps, int nstages, const int* w, int* a) {
  int tot = 0;
  for (int i = 0; i < nstages; ++i) tot += w[i] > 0 ? w[i] : 0;
  int used = 0, last = -1;
  for (int i = 0; i < 3; ++i) {
    if (i >= nstages || w[i] <= 0) a[i] = 0;
    else {
      a[i] = nwarps * w[i] / tot;
      if (a[i] < 1) a[i] = 1;
      used += a[i];
      last = i;
    }
  }
  if (last >= 0 && nwarps > used) a[last] += nwarps - used;
}

// ---------------------------------------------------------------------------------------
// The tile pipeline.
//
// Dependencies, per tile t (g = g0 + t):
//   IN(t)   after DONE(t - depth) from all ranks   (first `depth` tiles: after every rank
//           finished everything that came before this call on this ring)
//   MID(t)  after READY(g+1) from the ranks that feed me
//   OUT(t)  after MID(g+1) from all ranks
// IN(t) of every rank precedes READY, which precedes every MID(t) that writes into peers'
// slots, so the guard taken by IN also covers the slot writes of MID.
//
// Publishing a tile costs a system-scope fence, and a MEMBAR.SYS issued while the SM has
// NVLink stores in flight takes 10-25 us to return (profiles/membar_probe_n2.txt): it waits
// for every outstanding store of the SM to be acknowledged. Nothing that moves data may sit
// behind it. So a CTA is
//
//   warp 0      admitter   polls the flags every stage is waiting for (24 lanes = 3 stages x
//                          8 ranks, one pass reads them all), acquires, and admits tiles to
//                          the worker warps through shared memory. Never blocks.
//   warp 1      publisher  watches the workers' completion counters; whenever some stage has
//                          finished more tiles it issues ONE fence and raises every flag
//                          that became due (all stages, all channels) behind it. The fence
//                          latency delays flags, never data: ring depth absorbs it.
//   warps 2..   workers    one group per stage that has work on this rank; a warp runs
//                          "wait until tile t is admitted; do my share of it; count myself
//                          done" and goes straight on to tile t+1.
//
// Who raises which flag: a stage's flag follows the completion of that stage's workers; the
// flag of a stage with no work on this rank follows the nearest earlier stage that has work
// ("my MID(t) is done" is vacuously true the moment my IN(t) is; a rank without an OUT stage
// is finished with slot t as soon as its MID(t) is).
//
// Memory model: remote data -> peer's fence.sys + flag store (release) -> admitter's
// ld.acquire.sys -> st.release.cta (shared) -> worker's ld.acquire.cta: the worker's loads
// are causally after the peer's stores. Worker stores -> __syncwarp -> red.release.cta ->
// publisher's ld.acquire.cta -> fence.acq_rel.sys -> flag store: a release pattern that is
// cumulative over every worker's stores.
// ---------------------------------------------------------------------------------------
struct Roles {
  bool on[3];  // which stages have work on this rank
};

struct PipeShared {
  uint32_t admitted[3];   // tiles the workers of stage s may process (monotonic)
  uint32_t progress[16];  // per worker warp: tiles it has finished (a stage is as far as its
                          // slowest warp)
};

template <class P, int STAGE>
__device__ __noinline__ void worker_loop(const KArgs* kp, Ctx c, PipeShared* ps, int slot) {
  const KArgs& k = *kp;
  const int T = k.ntiles;
  const int lane = int(threadIdx.x) & 31;
  P p(k, c);
  uint32_t have = 0;
  const bool tr = k.trace && c.b == 0 && lane == 0;
  for (int t = 0; t < T; ++t) {
    uint64_t t0 = 0, t1 = 0;
    if (tr) t0 = globaltimer_ns();
    if (lane == 0) {
      while (have <= uint32_t(t)) {
        have = ld_acquire_cta_shared(&ps->admitted[STAGE]);
        if (have <= uint32_t(t)) __nanosleep(40);
      }
    }
    if (tr) t1 = globaltimer_ns();
    __syncwarp();
    if constexpr (STAGE == 0) p.in(t);
    else if constexpr (STAGE == 1) p.mid(t);
    else p.out(t);
    __syncwarp();
    if (lane == 0) st_release_cta_shared(&ps->progress[slot], uint32_t(t) + 1u);
    if (tr) trace_event(c, uint32_t(STAGE) | (uint32_t(slot) << 8), uint32_t(t), t0, globaltimer_ns(),
                        uint32_t(t1 - t0), 0u);
  }
}

template <class P>
__device__ __noinline__ void admitter_loop(const KArgs* kp, Ctx c, PipeShared* ps, Roles roles) {
  const KArgs& k = *kp;
  constexpr int NS = P::kHasOut ? 3 : 2;
  const uint32_t all = (1u << c.n) - 1u;
  const int T = k.ntiles;
  const int lane = int(threadIdx.x) & 31;
  const int s = lane >> 3, r = lane & 7;  // lane = (stage, source rank)
  P p(k, c);
  const bool stage_on = s < NS && roles.on[s];
  const uint32_t mask = s == 1 ? p.ready_from() : all;
  const int chan = s == 0 ? CH_DONE : s == 1 ? CH_READY : CH_MID;
  const bool active = stage_on && r < c.n && ((mask >> r) & 1u);
  const uint32_t* flag = c.my_pad + flag_index(chan, c.b, r < c.n ? r : 0);
  const unsigned grp = 0xffu << (8 * (s < 3 ? s : 0));
  uint32_t tw = 0, spins = 0;
  for (;;) {
    const bool pending = stage_on && tw < uint32_t(T);
    if (!__any_sync(0xffffffffu, pending)) break;
    bool ok = true;
    if (pending && active) {
      const uint32_t need = s == 0 ? (tw < uint32_t(c.depth) ? c.g0 : c.g0 + tw - uint32_t(c.depth) + 1u)
                                   : c.g0 + tw + 1u;
      ok = static_cast<int32_t>(ld_relaxed_sys(flag) - need) >= 0;
    }
    const unsigned vote = __ballot_sync(0xffffffffu, ok);
    const bool go = pending && (vote & grp) == grp;
    if (go && active) (void)ld_acquire_sys(flag);  // monotonic flag: still >= need
    __syncwarp();
    if (go) {
      ++tw;
      if (r == 0) {
        st_release_cta_shared(&ps->admitted[s], tw);
        if (k.trace && c.b == 0) {
          const uint64_t now = globaltimer_ns();
          trace_event(c, 32u + uint32_t(s), tw - 1u, now, now, 0u, 0u);
        }
      }
    }
    if ((++spins & 0x7ffu) == 0) {
      const bool late = globaltimer_ns() > c.deadline;
      if (__any_sync(0xffffffffu, late)) {
        // Give up waiting: record who we were waiting for, then admit everything so that
        // the workers run to the end (on garbage) and the publisher still raises our flags;
        // pdcc_comm_check reports the error. A mismatched collective never hangs the GPU.
        if (pending && active && !ok) {
          if (atomicCAS(c.ctrl + kCtrlErr, 0u, 1u) == 0u)
            c.ctrl[kCtrlErrInfo] = (uint32_t(chan) << 28) | (uint32_t(r) << 24) | ((c.g0 + tw + 1u) & 0xffffffu);
        }
        if (stage_on && r == 0) st_release_cta_shared(&ps->admitted[s], uint32_t(T));
        break;
      }
    }
  }
}

template <class P>
__device__ __noinline__ void publisher_loop(const KArgs* kp, Ctx c, PipeShared* ps, Roles roles,
                                            int nw0, int nw1, int nw2) {
  if ((threadIdx.x & 31) != 0) return;
  const KArgs& k = *kp;
  constexpr bool kOut = P::kHasOut;
  const uint32_t all = (1u << c.n) - 1u;
  const uint32_t T = uint32_t(k.ntiles);
  P p(k, c);
  const uint32_t to = p.ready_to();
  const bool in_on = roles.on[0], mid_on = roles.on[1], out_on = kOut && roles.on[2];
  uint32_t pub0 = 0, pub1 = 0, pub2 = 0;
  for (;;) {
    uint32_t d0 = T, d1 = T, d2 = T;  // tiles every worker warp of the stage has finished
    for (int i = 0; i < nw0; ++i) d0 = min(d0, ld_acquire_cta_shared(&ps->progress[i]));
    for (int i = 0; i < nw1; ++i) d1 = min(d1, ld_acquire_cta_shared(&ps->progress[nw0 + i]));
    if (out_on)
      for (int i = 0; i < nw2; ++i) d2 = min(d2, ld_acquire_cta_shared(&ps->progress[nw0 + nw1 + i]));
    const bool n0 = in_on && d0 > pub0, n1 = mid_on && d1 > pub1, n2 = out_on && d2 > pub2;
    if (!(n0 || n1 || n2)) {
      if ((!in_on || pub0 == T) && (!mid_on || pub1 == T) && (!out_on || pub2 == T)) break;
      __nanosleep(100);
      continue;
    }
    uint64_t tf = 0;
    if (k.trace && c.b == 0) tf = globaltimer_ns();
    fence_acq_rel_sys();  // ONE fence for everything that completed since the last round
    if (k.trace && c.b == 0)
      trace_event(c, 16u, 0u, tf, globaltimer_ns(), d0 | (d1 << 10) | (d2 << 20), 0u);
    if (n0) {
      const uint32_t v = c.g0 + d0;
      post_flags(c, CH_READY, v, to);
      if (kOut && !mid_on) post_flags(c, CH_MID, v, all);
      if (!mid_on && !out_on) post_flags(c, CH_DONE, v, all);
      pub0 = d0;
    }
    if (n1) {
      const uint32_t v = c.g0 + d1;
      if (kOut) post_flags(c, CH_MID, v, all);
      if (!out_on) post_flags(c, CH_DONE, v, all);
      pub1 = d1;
    }
    if (n2) {
      post_flags(c, CH_DONE, c.g0 + d2, all);
      pub2 = d2;
    }
  }
  c.ctrl[kCtrlEpoch + c.b] = c.g0 + T;
}

constexpr int kControlWarps = 2;

template <class P>
__device__ __forceinline__ void run_pipeline(const KArgs& k) {
  __shared__ Prologue sh;
  __shared__ PipeShared ps;
  Ctx c;
  if (threadIdx.x < 3) ps.admitted[threadIdx.x] = 0;
  if (threadIdx.x < 16) ps.progress[threadIdx.x] = 0;
  init_ctx(k, sh, c);  // contains the barriers that publish the zeroed counters

  constexpr int NS = P::kHasOut ? 3 : 2;
  int w[3] = {0, 0, 0};
  {
    P p(k, c);
    w[0] = p.w_in();
    w[1] = p.w_mid();
    if constexpr (P::kHasOut) w[2] = p.w_out();
  }
  // A rank with no work at all still takes part in the flag chain through its IN stage.
  if (w[0] <= 0 && w[1] <= 0 && w[2] <= 0) w[0] = 1;
  int a[3];
  split_warps((int(blockDim.x) >> 5) - kControlWarps, NS, w, a);
  const Roles roles{{w[0] > 0, w[1] > 0, w[2] > 0}};
  const int warp = int(threadIdx.x) >> 5;
  if (warp == 0) {
    admitter_loop<P>(&k, c, &ps, roles);
    return;
  }
  if (warp == 1) {
    publisher_loop<P>(&k, c, &ps, roles, a[0], a[1], a[2]);
    return;
  }
  const int ww = warp - kControlWarps;
  int stage, first;
  if (ww < a[0]) { stage = 0; first = 0; }
  else if (ww < a[0] + a[1]) { stage = 1; first = a[0]; }
  else if (ww < a[0] + a[1] + a[2]) { stage = 2; first = a[0] + a[1]; }
  else return;
  c.tid = int(threadIdx.x) - (first + kControlWarps) * 32;
  c.nt = a[stage] * 32;
  c.bar = 1 + stage;
  if (stage == 0) worker_loop<P, 0>(&k, c, &ps, ww);
  else if (stage == 1) worker_loop<P, 1>(&k, c, &ps, ww);
  else if constexpr (P::kHasOut) worker_loop<P, 2>(&k, c, &ps, ww);
}

// ---------------------------------------------------------------------------------------
// K1/K2/K3: all_reduce (main.py:23), fused scale+cast (README.md:5), reduce (main.py:14)
// ---------------------------------------------------------------------------------------
enum ArAlgo : int { AR_ONESHOT_P2P, AR_TWOSHOT_P2P, AR_ONESHOT_NVLS, AR_TWOSHOT_NVLS, AR_LL };

// Byte-exact 16-byte unit access on a user tensor range (ragged tail / odd views).
__device__ __forceinline__ uint4 load_user_unit(const char* base, uint64_t len, size_t j) {
  const char* p = base + (uint64_t(j) << 4);
  const uint64_t left = len - (uint64_t(j) << 4);
  if (left >= 16 && aligned16(p)) return ld_stream_v4(p);
  union {
    uint4 v;
    unsigned char b[16];
  } u;
  u.v = make_uint4(0, 0, 0, 0);
  const int nb = left < 16 ? int(left) : 16;
  for (int i = 0; i < nb; ++i) u.b[i] = static_cast<unsigned char>(p[i]);
  return u.v;
}

__device__ __forceinline__ void store_user_unit(char* base, uint64_t len, size_t j,
                                                const uint4& v) {
  char* p = base + (uint64_t(j) << 4);
  const uint64_t left = len - (uint64_t(j) << 4);
  if (left >= 16 && aligned16(p)) {
    st_v4(p, v);
    return;
  }
  union {
    uint4 v;
    unsigned char b[16];
  } u;
  u.v = v;
  const int nb = left < 16 ? int(left) : 16;
  for (int i = 0; i < nb; ++i) p[i] = static_cast<char>(u.b[i]);
}

// Push `len` bytes of a user tensor into slot offset `dst_off` of the ranks in `mask` (P2P
// stores) or of everyone (MC: one multimem.st per 16 bytes). A 16-byte aligned source
// whose length is a multiple of 16 streams through the unrolled vector loops; a ragged
// tail or an odd view is assembled byte-exactly, one unit per thread, and stored as whole
// 16-byte units all the same (slots have room). No staging and no barrier either way, so
// the warp groups of run_pipeline() never have to synchronise inside a stage.
template <bool MC>
__device__ __forceinline__ void push_user(const Ctx& c, const char* src, uint64_t len,
                                          size_t dst_off, size_t /*stage_off*/, uint32_t mask) {
  if (len == 0 || (!MC && mask == 0)) return;
  if (aligned16(src) && (len & 15u) == 0) {
    if constexpr (MC) push_mc<LD_USER, 8>(c, dst_off, src, len);
    else push_p2p<LD_USER, 8>(c, dst_off, src, len, mask);
    return;
  }
  const size_t n16 = size_t((len + 15) >> 4);
  for (size_t j = c.tid; j < n16; j += c.nt) {
    const uint4 v = load_user_unit(src, len, j);
    if constexpr (MC) {
      multimem_st_v4(c.mc(dst_off) + (j << 4), v);
    } else {
      for (int p = 0; p < c.n; ++p)
        if ((mask >> p) & 1u) st_v4(c.peer(p, dst_off) + (j << 4), v);
    }
  }
}

// Hot loop of every P2P reduction: whole passes of nt*U units, no bounds checks, every load
// unconditional (ranks q >= n alias my own operand and are skipped in the combine), all
// loads local. Operand q comes from the user tensor (q == rank) or from A[q] of my slot;
// combined in rank order 0..n-1 on every rank => bitwise identical results everywhere.
//   out   : user tensor position of unit 0, or nullptr when this rank does not keep it
//   base  : peers' block bases (shared memory); push_mask: ranks whose B[me] gets the result
template <typename T, int OP, bool WIDE, bool TWOSHOT, int NP, int U>
__device__ __noinline__ void combine_main(const char* __restrict__ user,
                                          const char* __restrict__ slot_a, size_t a_stride,
                                          size_t nmain, void* __restrict__ out, float scale,
                                          int rank, int n, char* const* __restrict__ base,
                                          size_t b_off, uint32_t push_mask, int tid, int nt) {
  using Tr = Traits<T>;
  const size_t pass = size_t(nt) * U;
  const bool scaled = scale != 1.0f;
  for (size_t j0 = tid; j0 < nmain; j0 += pass) {
    uint4 v[U][NP];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = j0 + size_t(u) * nt;
#pragma unroll
      for (int q = 0; q < NP; ++q) {
        const char* sq = (q == rank || q >= n) ? user : slot_a + size_t(q) * a_stride;
        v[u][q] = ld_stream_v4(sq + (j << 4));
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = j0 + size_t(u) * nt;
      Acc<T> acc;
      acc.init(v[u][0]);
#pragma unroll
      for (int q = 1; q < NP; ++q)
        if (q < n) acc.template add<OP>(v[u][q]);
      if (scaled) acc.scale(scale);
      if constexpr (TWOSHOT) {
        const uint4 res = acc.pack();
#pragma unroll
        for (int p = 0; p < NP; ++p)
          if ((push_mask >> p) & 1u) st_v4(base[p] + b_off + (j << 4), res);
        if constexpr (WIDE) acc.init(res);  // every rank must see the same rounded value
      }
      if (out != nullptr) {
        if constexpr (WIDE) {
          float* o = static_cast<float*>(out) + j * Tr::K;
#pragma unroll
          for (int i = 0; i < Tr::K; i += 4) {
            uint4 w;
            float* f = reinterpret_cast<float*>(&w);
            f[0] = float(acc.v[i]); f[1] = float(acc.v[i + 1]);
            f[2] = float(acc.v[i + 2]); f[3] = float(acc.v[i + 3]);
            st_v4(o + i, w);
          }
        } else {
          st_v4(static_cast<T*>(out) + j * Tr::K, acc.pack());
        }
      }
    }
  }
}

// Hot loop of the NVLS reductions over units [lo, hi) of a slot: 16 multimem.ld_reduce in
// flight per thread, then either the replicated store (all_reduce two-shot), a plain store
// into the root's slot (rooted two-shot) or the vector store to the user tensor (one-shot).
template <typename T, int KMM, bool TWOSHOT, bool ROOTED, bool WIDE>
__device__ __noinline__ void nvls_units(char* __restrict__ mc, char* __restrict__ root_slot,
                                        void* __restrict__ out, size_t lo, size_t hi, float scale,
                                        int tid, int nt) {
  using Tr = Traits<T>;
  // 16 reductions in flight per thread: this loop usually runs on one half of the CTA
  // (the other half stages the next tile), and a multimem.ld_reduce is a full round trip
  // through the switch, so the in-flight bytes per SM have to come from the unroll.
  constexpr int U = 16;
  const bool scaled = scale != 1.0f;
  for (size_t j0 = lo + tid; j0 < hi; j0 += size_t(nt) * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = j0 + size_t(u) * nt;
      if (j < hi) v[u] = multimem_ld_reduce_v4<KMM>(mc + (j << 4));  // never a redundant switch read
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = j0 + size_t(u) * nt;
      if (j < hi) {
        uint4 r = v[u];
        Acc<T> acc;
        if (scaled || (!TWOSHOT && WIDE)) {
          acc.init(r);
          if (scaled) acc.scale(scale);
          r = acc.pack();
        }
        if constexpr (TWOSHOT) {
          if constexpr (ROOTED) st_v4(root_slot + (j << 4), r);
          else multimem_st_v4(mc + (j << 4), r);
        } else if constexpr (WIDE) {
          float* o = static_cast<float*>(out) + j * Tr::K;
#pragma unroll
          for (int i = 0; i < Tr::K; i += 4) {
            uint4 w;
            float* f = reinterpret_cast<float*>(&w);
            f[0] = float(acc.v[i]); f[1] = float(acc.v[i + 1]);
            f[2] = float(acc.v[i + 2]); f[3] = float(acc.v[i + 3]);
            st_v4(o + i, w);
          }
        } else {
          st_v4(static_cast<T*>(out) + j * Tr::K, r);
        }
      }
    }
  }
}

// Slot layouts (offsets from the tile's slot, `tile` = KArgs::tile_bytes, `sub` = tile/owners):
//   one-shot P2P  A[q] at q*stride            rank q's whole tile        (tile <= slot/n)
//   two-shot P2P  A[q] at q*sub               rank q's partial of MY chunk
//                 B[o] at (n + o)*sub         reduced chunk o from its owner
//                 (owners: all n ranks, or the n-1 non-root ranks of a rooted reduce)
//   NVLS          the tile itself at 0        (multicast-addressable copy-in)
//
// The P2P variants are write-based: every byte that crosses NVLink is a posted store issued
// straight from the user tensor (IN) or from registers (MID); all loads are local. Remote
// loads cost a round trip per 16 bytes and measured 2-3x slower than stores on B200
// (profiles/sweep_n2_pull_baseline.jsonl vs sweep_n2_push.jsonl), so nothing here pulls.
template <typename T, bool WIDE, int OP, int ALGO, bool ROOTED>
struct ReducePolicy {
  using Tr = Traits<T>;
  static constexpr bool kTwoShot = (ALGO == AR_TWOSHOT_P2P || ALGO == AR_TWOSHOT_NVLS);
  static constexpr bool kNvls = (ALGO == AR_ONESHOT_NVLS || ALGO == AR_TWOSHOT_NVLS);
  static constexpr bool kHasOut = kTwoShot;
  static constexpr int kMm = MmFor<T, OP>::kind;

  const KArgs& k;
  const Ctx& c;
  Geo geo;
  uint32_t all;

  __device__ ReducePolicy(const KArgs& k_, const Ctx& c_) : k(k_), c(c_) {
    geo = Geo{k.bytes, k.tile_bytes, int(gridDim.x), c.b};
    all = (1u << c.n) - 1u;
  }
  // one-shot rooted reduce: only the root consumes, so only the root is signalled
  __device__ __forceinline__ uint32_t ready_to() const {
    return (ROOTED && !kTwoShot) ? (1u << k.root) : all;
  }
  __device__ __forceinline__ uint32_t ready_from() const {
    if constexpr (ROOTED && !kTwoShot) {
      if (c.rank != k.root) return 0u;
      // one-shot push: the root has no IN stage of its own, so it raises no READY
      return ALGO == AR_ONESHOT_P2P ? all & ~(1u << k.root) : all;
    } else {
      return all;
    }
  }
  // Relative weight of each stage on THIS rank (uniform across the CTA); run_pipeline()
  // sizes the warp groups from these. 0 = the stage has no work here.
  __device__ __forceinline__ int w_in() const {
    return (ROOTED && ALGO == AR_ONESHOT_P2P && c.rank == k.root) ? 0 : 1;
  }
  // The NVLS stages are lopsided: IN is a local copy (fast), MID is a round trip through
  // the switch per 16 bytes (latency-bound: it needs most of the warps to keep enough
  // multimem.ld_reduce in flight; profiles/trace_v4_n8.jsonl: 520 us of MID work against
  // 139 us of IN with equal groups), OUT a local copy again.
  __device__ __forceinline__ int w_mid() const {
    if constexpr (kNvls) {
      if constexpr (ROOTED && ALGO == AR_ONESHOT_NVLS) return c.rank == k.root ? 4 : 0;
      return 4;
    } else if constexpr (!ROOTED) {
      return 1;
    } else if constexpr (ALGO == AR_TWOSHOT_P2P) {
      return c.rank != k.root ? 1 : 0;  // owners
    } else {
      return c.rank == k.root ? 1 : 0;
    }
  }
  __device__ __forceinline__ int w_out() const {
    if (ROOTED && c.rank != k.root) return 0;
    return kNvls ? 2 : 1;
  }
  // Two-shot P2P ownership. all_reduce: every rank owns one chunk of the tile. Rooted
  // reduce: only the n-1 NON-root ranks own chunks, so the root receives each reduced
  // chunk exactly once and nothing else: S bytes in, S bytes out on every GPU, instead of
  // 2S(n-1)/n into the root when it also collects partials for a chunk of its own.
  static constexpr bool kRootOwnsNothing = ROOTED && ALGO == AR_TWOSHOT_P2P;
  __device__ __forceinline__ int nown() const { return kRootOwnsNothing ? c.n - 1 : c.n; }
  __device__ __forceinline__ int own_index(int r) const {
    return kRootOwnsNothing ? (r - k.root - 1 + c.n) % c.n : r;
  }
  __device__ __forceinline__ bool is_owner(int r) const { return !kRootOwnsNothing || r != k.root; }
  __device__ __forceinline__ size_t sub_bytes() const { return size_t(k.tile_bytes / nown()); }
  // bytes of chunk q inside a tile of `len` valid bytes
  __device__ __forceinline__ uint64_t chunk_len(uint64_t len, int q) const {
    const uint64_t lo = uint64_t(q) * sub_bytes();
    if (len <= lo) return 0;
    return len - lo < sub_bytes() ? len - lo : sub_bytes();
  }
  // A[p] (partials from rank p) sits at p*sub, B[o] (reduced chunk o) behind the n A's
  __device__ __forceinline__ size_t b_off(size_t so, int o) const {
    return so + (size_t(c.n) + size_t(o)) * sub_bytes();
  }

  __device__ __forceinline__ void in(int t) {
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    const size_t so = c.slot_off(t);
    const char* src = static_cast<const char*>(c.a->in) + off;
    if constexpr (kNvls) {
      copy_bytes<LD_USER, 8>(c, c.mine(so), src, len);
    } else if constexpr (ALGO == AR_ONESHOT_P2P) {
      const size_t mine_off = so + size_t(c.rank) * k.stride;
      if constexpr (ROOTED) {
        if (c.rank != k.root) push_user<false>(c, src, len, mine_off, mine_off, 1u << k.root);
      } else {
        push_user<false>(c, src, len, mine_off, mine_off, all & ~(1u << c.rank));
      }
    } else {  // two-shot P2P: my partial of each chunk goes to the chunk's owner
      const size_t sub = sub_bytes();
      if (aligned16(src) && (len & 15u) == 0) {
        // one streaming loop over the tile; chunk o lands in A[me] of its owner's slot
        deal16<8>(c.a->comm.base, kRootOwnsNothing ? k.root + 1 : 0, c.n, c.rank,
                  so + size_t(c.rank) * sub, 0, src, size_t(len >> 4), sub >> 4, c.tid, c.nt);
        return;
      }
      for (int i = 1; i < c.n; ++i) {
        const int p = (c.rank + i) % c.n;
        if (!is_owner(p)) continue;
        const int o = own_index(p);
        push_user<false>(c, src + size_t(o) * sub, chunk_len(len, o), so + size_t(c.rank) * sub,
                         b_off(so, o), 1u << p);
      }
    }
  }

  // Combine `nunits` 16-byte units: rank q's operand comes from the user tensor (q == me)
  // or from A[q] of my own slot, in rank order 0..n-1 on every rank (bitwise identical
  // results). The bulk runs a branch-free loop with U x n vector loads issued before the
  // first combine; a ragged last unit or an odd user view falls to a compact scalar-safe
  // loop (one unit at a time) so the fast loop stays in registers.
  //   write_out : the result goes to the user tensor (always for one-shot; owner's own
  //               chunk for two-shot; only on the root when ROOTED)
  //   push_mask : two-shot: ranks whose B[me] receives the packed result
  template <int NP, int U>
  __device__ __forceinline__ void combine_units(const char* user, uint64_t user_len, size_t a_off,
                                                size_t a_stride, size_t nunits, size_t elem0,
                                                bool write_out, size_t b_off, uint32_t push_mask) {
    const int tid = c.tid, nt = c.nt;
    const bool scaled = k.scale != 1.0f;
    const size_t nelem = user_len / sizeof(T);
    using Out = typename std::conditional<WIDE, float, T>::type;
    Out* outp = static_cast<Out*>(c.a->out) + elem0;
    const bool fast = aligned16(user) && (!write_out || aligned16(outp));
    const size_t nfast = fast ? size_t(user_len >> 4) : 0;

    // Operand/peer addresses are recomputed from two bases instead of being held in
    // 2 x NP pointer registers: all loads are local (HBM latency, not NVLink), so register
    // room is better spent on the U x NP vectors in flight.
    const char* slot_a = c.mine(a_off);

    const size_t pass = size_t(nt) * U;
    const size_t nmain = nfast / pass * pass;
    if (nmain)
      combine_main<T, OP, WIDE, kTwoShot, NP, U>(user, slot_a, a_stride, nmain,
                                                 write_out ? static_cast<void*>(outp) : nullptr,
                                                 k.scale, c.rank, c.n, c.a->comm.base, b_off,
                                                 push_mask, tid, nt);

    // remainder, ragged tail, odd views: one unit per thread per pass, byte-exact on the
    // user side (load_user_unit / store_unit take their vector path when they can)
    for (size_t j = nmain + tid; j < nunits; j += nt) {
      Acc<T> acc;
      for (int q = 0; q < c.n; ++q) {
        const uint4 v = q == c.rank
                            ? load_user_unit(user, user_len, j)
                            : ld_stream_v4(c.mine(a_off + size_t(q) * a_stride) + (j << 4));
        if (q == 0) acc.init(v);
        else acc.template add<OP>(v);
      }
      if (scaled) acc.scale(k.scale);
      if constexpr (kTwoShot) {
        const uint4 res = acc.pack();
        for (int p = 0; p < c.n; ++p)
          if ((push_mask >> p) & 1u) st_v4(c.peer(p, b_off) + (j << 4), res);
        if constexpr (WIDE) acc.init(res);
      }
      if (write_out) {
        const size_t e = j * Tr::K;
        const int valid = nelem - e < size_t(Tr::K) ? int(nelem - e) : Tr::K;
        store_unit<T, WIDE>(c.a->out, elem0 + e, acc, valid);
      }
    }
  }

  template <typename... A>
  __device__ __forceinline__ void combine_dispatch(A... a) {
    if (c.n <= 2) combine_units<2, 4>(a...);
    else if (c.n <= 4) combine_units<4, 2>(a...);
    else combine_units<8, 1>(a...);
  }

  // NVLS: the switch reduces (nvls_units, below the class).
  template <bool TWOSHOT>
  __device__ __forceinline__ void mid_nvls(size_t so, size_t lo, size_t hi, size_t nelem,
                                           size_t elem0) {
    if constexpr (kMm != MM_NONE) {
      using Out = typename std::conditional<WIDE, float, T>::type;
      Out* outp = static_cast<Out*>(c.a->out) + elem0;
      // full, aligned units go through the unrolled loop; the ragged last unit (one-shot
      // only: two-shot works on whole units of the slot) through the scalar-safe path
      const bool vec_out = TWOSHOT || aligned16(outp);
      size_t hi_main = hi;
      if (!TWOSHOT) {
        const size_t full = nelem / Tr::K;  // units that are entirely inside the tensor
        hi_main = vec_out ? (full < hi ? full : hi) : lo;
        if (hi_main < lo) hi_main = lo;
      }
      nvls_units<T, kMm, TWOSHOT, ROOTED, WIDE>(
          c.mc(so), ROOTED && TWOSHOT ? c.peer(k.root, so) : nullptr, outp, lo, hi_main, k.scale,
          c.tid, c.nt);
      if constexpr (!TWOSHOT) {
        for (size_t j = hi_main + c.tid; j < hi; j += c.nt) {
          Acc<T> acc;
          acc.init(multimem_ld_reduce_v4<kMm>(c.mc(so) + (j << 4)));
          if (k.scale != 1.0f) acc.scale(k.scale);
          const size_t e = j * Tr::K;
          const int valid = nelem - e < size_t(Tr::K) ? int(nelem - e) : Tr::K;
          store_unit<T, WIDE>(c.a->out, elem0 + e, acc, valid);
        }
      }
    }
  }

  __device__ __forceinline__ void mid(int t) {
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    const size_t so = c.slot_off(t);
    const size_t n16 = (len + 15) >> 4;
    const size_t nelem = len / sizeof(T);
    const size_t elem0 = off / sizeof(T);
    const char* src = static_cast<const char*>(c.a->in) + off;
    if constexpr (ALGO == AR_ONESHOT_P2P) {
      if (ROOTED && c.rank != k.root) return;
      combine_dispatch(src, len, so, size_t(k.stride), n16, elem0, true, size_t(0), 0u);
    } else if constexpr (ALGO == AR_ONESHOT_NVLS) {
      if (ROOTED && c.rank != k.root) return;
      mid_nvls<false>(so, 0, n16, nelem, elem0);
    } else if constexpr (ALGO == AR_TWOSHOT_P2P) {
      // I own one chunk of the tile: combine the partials in A, keep my part of the result
      // (all_reduce) and hand the reduced chunk to whoever needs it (B[o] on their side).
      if (!is_owner(c.rank)) return;
      const int o = own_index(c.rank);
      const size_t sub = sub_bytes();
      const uint64_t clen = chunk_len(len, o);
      if (clen == 0) return;
      const uint32_t others = all & ~(1u << c.rank);
      const uint32_t push_mask = ROOTED ? (1u << k.root) : others;
      combine_dispatch(src + size_t(o) * sub, clen, so, sub, size_t((clen + 15) >> 4),
                       elem0 + size_t(o) * sub / sizeof(T), !ROOTED, b_off(so, o), push_mask);
    } else {
      // NVLS two-shot: this rank owns units [lo, hi) of the tile
      const size_t sub16 = (k.tile_bytes / c.n) >> 4;
      const size_t lo = size_t(c.rank) * sub16;
      size_t hi = lo + sub16;
      if (hi > n16) hi = n16;
      if (lo < hi) mid_nvls<true>(so, lo, hi, nelem, elem0);
    }
  }

  // slot bytes -> user tensor, widening 16-bit floats to fp32 when WIDE
  __device__ __forceinline__ void copy_out(const char* slot, uint64_t byte_off, uint64_t len) {
    if constexpr (!WIDE) {
      copy_bytes<LD_SYS, 8>(c, static_cast<char*>(c.a->out) + byte_off, slot, len);
    } else {
      const int tid = c.tid, nt = c.nt;
      const size_t n16 = (len + 15) >> 4;
      const size_t nelem = len / sizeof(T);
      const size_t elem0 = byte_off / sizeof(T);
      for (size_t j = tid; j < n16; j += nt) {
        Acc<T> acc;
        acc.init(ld_stream_v4(slot + (j << 4)));
        const size_t e = j * Tr::K;
        const int valid = nelem - e < size_t(Tr::K) ? int(nelem - e) : Tr::K;
        store_unit<T, true>(c.a->out, elem0 + e, acc, valid);
      }
    }
  }

  __device__ __forceinline__ void out(int t) {
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    const size_t so = c.slot_off(t);
    if constexpr (ALGO == AR_TWOSHOT_P2P) {
      if (ROOTED && c.rank != k.root) return;
      const size_t sub = sub_bytes();
      for (int i = 1; i < c.n; ++i) {
        const int j = (c.rank + i) % c.n;  // every owner but me
        const int o = own_index(j);
        const uint64_t clen = chunk_len(len, o);
        if (clen) copy_out(c.mine(b_off(so, o)), off + size_t(o) * sub, clen);
      }
    } else {
      if (ROOTED && c.rank != k.root) return;
      copy_out(c.mine(so), off, len);
    }
  }
};

// The ReduceOp is a template parameter (chosen on the host, launch.cu) so that every kernel
// carries exactly one combine: no four-way switch around the pipeline, no register spills.
template <typename T, bool WIDE, int OP, int ALGO, bool ROOTED>
__global__ void __launch_bounds__(kMaxThreads, 1) reduce_kernel(const __grid_constant__ KArgs k) {
  run_pipeline<ReducePolicy<T, WIDE, OP, ALGO, ROOTED>>(k);
}

// ---------------------------------------------------------------------------------------
// K4-K7 + K9: byte movers. List packing/unpacking is part of the kernel (pointer table in
// RankArgs::list), never a separate launch.
// ---------------------------------------------------------------------------------------
enum MoveKind : int {
  MV_BROADCAST_P2P,   // main.py:81  root stores the tile into every peer's slot
  MV_BROADCAST_NVLS,  // main.py:81  root multimem.st once, the switch replicates
  MV_BROADCAST_2P_P2P,   // main.py:81  root scatters shards, every rank re-sends its shard
  MV_BROADCAST_2P_NVLS,  // main.py:81  same, second phase as one multimem.st per shard
  MV_SCATTER,         // main.py:37  root pushes list[j] into rank j's slot
  MV_GATHER,          // main.py:52  rank r pushes into sub-slot r of the root's slot
  MV_ALLGATHER_PULL,  // main.py:68  stage own shard, pull every peer's into list[p]
  MV_ALLGATHER_PUSH,  // main.py:68  store own shard into sub-slot r of every peer
  MV_ALLGATHER_NVLS,  // main.py:68  multimem.st own shard into sub-slot r everywhere
  MV_BARRIER
};

template <int KIND>
struct MovePolicy {
  // Two-phase broadcast (scatter + all-gather) is the only mover with a third stage: a
  // single GPU cannot feed n-1 peers (or the multicast engine) at link rate, n GPUs each
  // re-sending 1/n can.
  static constexpr bool kTwoPhase = KIND == MV_BROADCAST_2P_P2P || KIND == MV_BROADCAST_2P_NVLS;
  static constexpr bool kHasOut = kTwoPhase;
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
  // two-phase broadcast: the n-1 non-root ranks each own one shard of the tile
  __device__ __forceinline__ int own_index(int r) const { return (r - k.root - 1 + c.n) % c.n; }
  __device__ __forceinline__ size_t sub_bytes() const { return size_t(k.tile_bytes / (c.n - 1)); }
  __device__ __forceinline__ uint64_t chunk_len(uint64_t len, int q) const {
    const uint64_t lo = uint64_t(q) * sub_bytes();
    if (len <= lo) return 0;
    return len - lo < sub_bytes() ? len - lo : sub_bytes();
  }
  __device__ __forceinline__ uint32_t ready_to() const {
    if constexpr (kTwoPhase) return c.rank == k.root ? all : 0u;  // root included
    else if constexpr (kRootSends) return c.rank == k.root ? all : 0u;
    else if constexpr (KIND == MV_GATHER) return 1u << k.root;
    else return all;
  }
  __device__ __forceinline__ uint32_t ready_from() const {
    if constexpr (kTwoPhase) return 1u << k.root;
    else if constexpr (kRootSends) return c.rank == k.root ? 0u : (1u << k.root);
    else if constexpr (KIND == MV_GATHER) return c.rank == k.root ? all : 0u;
    else return all;
  }

  // relative weight of each stage on THIS rank (0 = no work here), see run_pipeline()
  __device__ __forceinline__ int w_in() const {
    if constexpr (kTwoPhase || kRootSends) return c.rank == k.root ? 1 : 0;
    else return 1;  // barrier: the IN group only raises READY
  }
  __device__ __forceinline__ int w_mid() const {
    if constexpr (kTwoPhase || kRootSends) return c.rank != k.root ? 1 : 0;
    else if constexpr (KIND == MV_GATHER) return c.rank == k.root ? 3 : 0;
    else if constexpr (KIND == MV_ALLGATHER_PULL) return 3;
    else if constexpr (KIND == MV_ALLGATHER_NVLS) return 2;
    else return 1;  // barrier: the MID group waits for everybody's READY
  }
  __device__ __forceinline__ int w_out() const { return (kTwoPhase && c.rank != k.root) ? 1 : 0; }

  template <bool MC>
  __device__ __forceinline__ void push(const char* src, uint64_t len, size_t dst_off,
                                       size_t stage_off, uint32_t mask) {
    push_user<MC>(c, src, len, dst_off, stage_off, mask);
  }

  __device__ __forceinline__ void in(int t) {
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    const size_t so = c.slot_off(t);
    if constexpr (KIND == MV_BROADCAST_P2P || KIND == MV_BROADCAST_NVLS) {
      if (c.rank != k.root) return;
      const char* src = static_cast<const char*>(c.a->in) + off;
      if constexpr (KIND == MV_BROADCAST_NVLS) push<true>(src, len, so, so, all);
      else push<false>(src, len, so, so, all & ~(1u << k.root));
    } else if constexpr (kTwoPhase) {
      // phase 1: the root deals one shard of the tile to each other rank (S bytes out)
      if (c.rank != k.root) return;
      const char* src = static_cast<const char*>(c.a->in) + off;
      const size_t sub = sub_bytes();
      if (aligned16(src) && (len & 15u) == 0) {
        // one streaming loop over the tile; shard o lands at the same offset of its owner's slot
        deal16<8>(c.a->comm.base, k.root + 1, c.n, k.root, so, sub, src, size_t(len >> 4), sub >> 4,
                  c.tid, c.nt);
        return;
      }
      for (int i = 1; i < c.n; ++i) {
        const int j = (c.rank + i) % c.n;
        const int o = own_index(j);
        push<false>(src + size_t(o) * sub, chunk_len(len, o), so + size_t(o) * sub,
                    so + size_t(o) * sub, 1u << j);
      }
    } else if constexpr (KIND == MV_SCATTER) {
      if (c.rank != k.root) return;
      for (int j = 0; j < c.n; ++j) {
        const char* src = static_cast<const char*>(c.a->list[j]) + off;
        if (j == c.rank) copy_bytes<LD_USER, 8>(c, static_cast<char*>(c.a->out) + off, src, len);
        else push<false>(src, len, so, so, 1u << j);
      }
    } else if constexpr (KIND == MV_GATHER) {
      const char* src = static_cast<const char*>(c.a->in) + off;
      if (c.rank == k.root) {
        copy_bytes<LD_USER, 8>(c, static_cast<char*>(c.a->list[c.rank]) + off, src, len);
      } else {
        push<false>(src, len, so + size_t(c.rank) * k.stride, so, 1u << k.root);
      }
    } else if constexpr (KIND == MV_ALLGATHER_PULL) {
      const char* src = static_cast<const char*>(c.a->in) + off;
      copy_bytes<LD_USER, 8>(c, c.mine(so), src, len);
      copy_bytes<LD_USER, 8>(c, static_cast<char*>(c.a->list[c.rank]) + off, src, len);
    } else if constexpr (KIND == MV_ALLGATHER_NVLS) {
      const char* src = static_cast<const char*>(c.a->in) + off;
      const size_t sub = so + size_t(c.rank) * k.stride;
      push<true>(src, len, sub, sub, all);
    } else if constexpr (KIND == MV_ALLGATHER_PUSH) {
      const char* src = static_cast<const char*>(c.a->in) + off;
      const size_t sub = so + size_t(c.rank) * k.stride;
      push<false>(src, len, sub, sub, all & ~(1u << c.rank));
      copy_bytes<LD_USER, 8>(c, static_cast<char*>(c.a->list[c.rank]) + off, src, len);
    }
  }

  __device__ __forceinline__ void mid(int t) {
    uint64_t off, len;
    geo.range(t, off, len);
    if (len == 0) return;
    const size_t so = c.slot_off(t);
    if constexpr (KIND == MV_BROADCAST_P2P || KIND == MV_BROADCAST_NVLS || KIND == MV_SCATTER) {
      if (c.rank == k.root) return;
      copy_bytes<LD_SYS, 8>(c, static_cast<char*>(c.a->out) + off, c.mine(so), len);
    } else if constexpr (kTwoPhase) {
      // phase 2: every non-root rank re-sends its shard to the other non-root ranks, so
      // each of them also receives exactly S bytes and sends at most S.
      if (c.rank == k.root) return;
      const size_t sub = sub_bytes();
      const int o = own_index(c.rank);
      const uint64_t clen = chunk_len(len, o);
      if (clen == 0) return;
      const size_t mine_off = so + size_t(o) * sub;
      const uint32_t others = all & ~(1u << c.rank) & ~(1u << k.root);
      if constexpr (KIND == MV_BROADCAST_2P_NVLS) push_mc<LD_SYS, 8>(c, mine_off, c.mine(mine_off), clen);
      else push_p2p<LD_SYS, 8>(c, mine_off, c.mine(mine_off), clen, others);
    } else if constexpr (KIND == MV_GATHER) {
      if (c.rank != k.root) return;
      for (int j = 0; j < c.n; ++j)
        if (j != c.rank)
          copy_bytes<LD_SYS, 8>(c, static_cast<char*>(c.a->list[j]) + off,
                                c.mine(so + size_t(j) * k.stride), len);
    } else if constexpr (KIND == MV_ALLGATHER_PULL) {
      for (int i = 1; i < c.n; ++i) {
        const int p = (c.rank + i) % c.n;  // stagger peers so links load evenly
        copy_bytes<LD_SYS, 8>(c, static_cast<char*>(c.a->list[p]) + off, c.peer(p, so), len);
      }
    } else if constexpr (KIND == MV_ALLGATHER_NVLS) {
      for (int j = 0; j < c.n; ++j)
        copy_bytes<LD_SYS, 8>(c, static_cast<char*>(c.a->list[j]) + off,
                              c.mine(so + size_t(j) * k.stride), len);
    } else if constexpr (KIND == MV_ALLGATHER_PUSH) {
      for (int i = 1; i < c.n; ++i) {
        const int j = (c.rank + i) % c.n;
        copy_bytes<LD_SYS, 8>(c, static_cast<char*>(c.a->list[j]) + off,
                              c.mine(so + size_t(j) * k.stride), len);
      }
    }
  }
  __device__ __forceinline__ void out(int t) {
    if constexpr (kTwoPhase) {
      if (c.rank == k.root) return;
      uint64_t off, len;
      geo.range(t, off, len);
      if (len == 0) return;
      copy_bytes<LD_SYS, 8>(c, static_cast<char*>(c.a->out) + off, c.mine(c.slot_off(t)), len);
    }
  }
};

template <int KIND>
__global__ void __launch_bounds__(kMaxThreads, 1) move_kernel(const __grid_constant__ KArgs k) {
  run_pipeline<MovePolicy<KIND>>(k);
}

// per-dtype launch tables, one translation unit each (kern_<dtype>.cu)
typedef void (*KernelFn)(KArgs);
KernelFn reduce_kernel_for(int dtype, bool wide, int redop, int ar_algo, bool rooted);
#define PDCC_DECLARE_REDUCE_TABLE(NAME) KernelFn NAME(bool wide, int redop, int ar_algo, bool rooted)
PDCC_DECLARE_REDUCE_TABLE(reduce_table_f32);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_bf16);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_f16);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_f64);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_i32);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_i64);
PDCC_DECLARE_REDUCE_TABLE(reduce_table_u8);

}  // namespace pdcc
