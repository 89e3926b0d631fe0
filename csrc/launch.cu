// Host side of the kernels: size -> (algorithm, grid, tile) planning, argument checks,
// kernel selection and the launch itself (plain launch for a real rank, cooperative
// launch with grid.y = nranks for the single-GPU emulation).
//
// The thresholds in plan_collective() are the measured size->variant table of
// SURVEY.md §7.1 step 9; profiles/ holds the sweeps they were read from.
#include <cuda_runtime.h>

#include <string.h>

#include <atomic>
#include <mutex>

#include "ll.cuh"
#define PDCC_SYM_MOVERS
#include "sym.cuh"
#include "tma_copy.cuh"

namespace pdcc {

std::atomic<uint64_t> g_launches{0};

size_t dtype_size(int dtype) {
  switch (dtype) {
    case PDCC_F32: case PDCC_I32: return 4;
    case PDCC_BF16: case PDCC_F16: case PDCC_I16: return 2;
    case PDCC_F64: case PDCC_I64: return 8;
    case PDCC_U8: case PDCC_I8: return 1;
    default: return 0;
  }
}

// Can the switch do (dtype, op)?  Mirrors MmFor<> in reduce.cuh.
static bool nvls_reducible(int dtype, int redop) {
  if (redop == PDCC_PRODUCT) return false;
  switch (dtype) {
    case PDCC_F32: case PDCC_F64: return redop == PDCC_SUM;
    case PDCC_BF16: case PDCC_F16: case PDCC_I32: case PDCC_I64: return true;
    default: return false;
  }
}

KernelFn reduce_kernel_for(int dtype, bool wide, int redop, int ar_algo, bool rooted) {
  switch (dtype) {
    case PDCC_F32: return reduce_table_f32(wide, redop, ar_algo, rooted);
    case PDCC_BF16: return reduce_table_bf16(wide, redop, ar_algo, rooted);
    case PDCC_F16: return reduce_table_f16(wide, redop, ar_algo, rooted);
    case PDCC_F64: return reduce_table_f64(wide, redop, ar_algo, rooted);
    case PDCC_I32: return reduce_table_i32(wide, redop, ar_algo, rooted);
    case PDCC_I64: return reduce_table_i64(wide, redop, ar_algo, rooted);
    case PDCC_U8: return reduce_table_u8(wide, redop, ar_algo, rooted);
    default: return nullptr;
  }
}

static KernelFn move_kernel_for(int kind) {
  switch (kind) {
    case MV_BROADCAST_P2P: return move_kernel<MV_BROADCAST_P2P>;
    case MV_BROADCAST_NVLS: return move_kernel<MV_BROADCAST_NVLS>;
    case MV_BROADCAST_2P_P2P: return move_kernel<MV_BROADCAST_2P_P2P>;
    case MV_BROADCAST_2P_NVLS: return move_kernel<MV_BROADCAST_2P_NVLS>;
    case MV_SCATTER: return move_kernel<MV_SCATTER>;
    case MV_GATHER: return move_kernel<MV_GATHER>;
    case MV_ALLGATHER_PULL: return move_kernel<MV_ALLGATHER_PULL>;
    case MV_ALLGATHER_NVLS: return move_kernel<MV_ALLGATHER_NVLS>;
    case MV_ALLGATHER_PUSH: return move_kernel<MV_ALLGATHER_PUSH>;
    case MV_BARRIER: return move_kernel<MV_BARRIER>;
    default: return nullptr;
  }
}

static KernelFn ll_move_kernel_for(int coll) {
  switch (coll) {
    case PDCC_BROADCAST: return ll_move_kernel<LL_BROADCAST>;
    case PDCC_SCATTER: return ll_move_kernel<LL_SCATTER>;
    case PDCC_GATHER: return ll_move_kernel<LL_GATHER>;
    case PDCC_ALL_GATHER: return ll_move_kernel<LL_ALL_GATHER>;
    default: return nullptr;
  }
}

static uint64_t round_up(uint64_t v, uint64_t m) { return (v + m - 1) / m * m; }
static uint64_t ceil_div(uint64_t a, uint64_t b) { return (a + b - 1) / b; }
__device__ __forceinline__ uint64_t round_up_dev(uint64_t v, uint64_t m) { return (v + m - 1) / m * m; }
__device__ __forceinline__ uint64_t ceil_div_dev(uint64_t a, uint64_t b) { return (a + b - 1) / b; }

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

// Planner knobs. Read from the environment ONCE (first use) and cached: the planner runs on
// every collective call and getenv is a linear scan of the environment. pdcc_reload_env()
// re-reads them (benchmarks A/B a knob inside one job).
struct PlanEnv {
  int ll_max;             // PDCC_LL_MAX            bytes; <0 = default (512 KiB / n)
  int oneshot_p2p_max;    // PDCC_ONESHOT_P2P_MAX   all_reduce one-shot up to this size (n > 2)
  int nvls_min_ranks;     // PDCC_NVLS_MIN_RANKS    all_reduce / reduce use the switch above this group size
  int rooted_twoshot_min; // PDCC_ROOTED_TWOSHOT_MIN reduce/broadcast: two-phase from this size
  int max_ctas;           // PDCC_MAX_CTAS          hard cap on CTAs per call (0 = none)
  int default_ctas;       // PDCC_DEFAULT_CTAS      cap below 8 MiB (0 = 128; >0 also caps big ones)
  int tile_bytes;         // PDCC_TILE_BYTES        payload per CTA per tile (0 = table)
  int depth;              // PDCC_DEPTH             max slots per ring (0 = 16)
  int threads;            // PDCC_THREADS           threads per CTA of the pipeline kernels (0 = auto)
  int copy_tma;           // PDCC_COPY_TMA          local copies through the TMA engine (default 1)
  int trace;              // PDCC_TRACE             CTA 0 records a timeline (pdcc_comm_read_trace)
  int copy_chunk_kb;      // PDCC_COPY_CHUNK_KB     TMA local copy: KiB per bulk copy
  int copy_stages;        // PDCC_COPY_STAGES       .. stages of the shared-memory ring
  int copy_ctas_per_sm;   // PDCC_COPY_CTAS_PER_SM  .. CTAs per SM
};
static PlanEnv g_plan_env;
static std::atomic<int> g_plan_env_state{0};  // 0 = not read, 1 = valid

static void read_plan_env(PlanEnv* e) {
  e->ll_max = env_int("PDCC_LL_MAX", -1);
  e->oneshot_p2p_max = env_int("PDCC_ONESHOT_P2P_MAX", 1 << 20);
  e->nvls_min_ranks = env_int("PDCC_NVLS_MIN_RANKS", 3);
  e->rooted_twoshot_min = env_int("PDCC_ROOTED_TWOSHOT_MIN", 2 << 20);
  e->max_ctas = env_int("PDCC_MAX_CTAS", 0);
  e->default_ctas = env_int("PDCC_DEFAULT_CTAS", 0);
  e->tile_bytes = env_int("PDCC_TILE_BYTES", 0);
  e->depth = env_int("PDCC_DEPTH", 0);
  e->threads = env_int("PDCC_THREADS", 0);
  e->copy_tma = env_int("PDCC_COPY_TMA", 1);
  e->trace = env_int("PDCC_TRACE", 0);
  e->copy_chunk_kb = env_int("PDCC_COPY_CHUNK_KB", int(kTmaChunk >> 10));
  e->copy_stages = env_int("PDCC_COPY_STAGES", kTmaStages);
  e->copy_ctas_per_sm = env_int("PDCC_COPY_CTAS_PER_SM", 1);
}

static const PlanEnv& plan_env() {
  if (g_plan_env_state.load(std::memory_order_acquire) == 0) {
    static std::mutex mu;
    std::lock_guard<std::mutex> g(mu);
    if (g_plan_env_state.load(std::memory_order_relaxed) == 0) {
      read_plan_env(&g_plan_env);
      g_plan_env_state.store(1, std::memory_order_release);
    }
  }
  return g_plan_env;
}

// Payload bytes per CTA per tile when the caller does not force one (PDCC_TILE_BYTES).
// Measured table (profiles/): see DESIGN.md §4 "tiles".
static uint64_t default_tile(int coll, int algo, int n, uint64_t bytes) {
  // 64 KiB per CTA per tile: with 148 CTAs a 256 MiB message is 28 tiles per CTA, the
  // pipeline fill is a few percent, and 16 slots of it fit the default 2 MiB ring for the
  // NVLS and mover layouts (profiles/sweep_r2_*.jsonl: 32-64 KiB best at 4 and 8 GPUs,
  // 128 KiB and more lose 5-15 %).
  (void)coll; (void)algo; (void)n;
  if (bytes >= (uint64_t(8) << 20)) return 64 << 10;
  return 32 << 10;
}

int plan_collective(const DevComm& comm, int ring_limit, bool has_mc, const pdcc_op_t& op,
                    Plan* plan) {
  const int n = comm.nranks;
  const size_t esz = dtype_size(op.dtype);
  if (esz == 0) {
    set_error("unsupported dtype %d", op.dtype);
    return PDCC_EINVAL;
  }
  if (op.coll < 0 || op.coll >= PDCC_NUM_COLLS) {
    set_error("unknown collective %d", op.coll);
    return PDCC_EINVAL;
  }
  const uint64_t bytes = op.count * esz;
  const bool is_reduce = op.coll == PDCC_ALL_REDUCE || op.coll == PDCC_REDUCE;
  const int redop = op.redop == PDCC_AVG ? PDCC_SUM : op.redop;
  if (is_reduce && !(redop == PDCC_SUM || redop == PDCC_PRODUCT || redop == PDCC_MIN ||
                     redop == PDCC_MAX)) {
    set_error("unsupported reduce op %d", op.redop);
    return PDCC_EINVAL;
  }

  // ---- algorithm -----------------------------------------------------------------
  int algo = op.algo;
  const bool mc_ok = has_mc && comm.mc != nullptr;
  // Small messages: flag-in-data push (ll.cuh). Above the threshold the doubled wire
  // bytes cost more than the fences they save. Thresholds from profiles/ sweeps.
  if (algo == PDCC_ALGO_AUTO && op.coll != PDCC_BARRIER && n > 1 &&
      bytes <= uint64_t(plan_env().ll_max >= 0 ? plan_env().ll_max : (512 << 10) / n))
    algo = PDCC_ALGO_LL;
  if (algo == PDCC_ALGO_LL) {
    if (op.coll == PDCC_BARRIER) {
      set_error("barrier has no LL variant");
      return PDCC_EINVAL;
    }
    int nb = op.nblocks > 0 ? op.nblocks : int(ceil_div(bytes ? bytes : 1, kLLTileBytes));
    int limit = ring_limit > 0 && ring_limit < kLLRings ? ring_limit : kLLRings;
    if (nb > limit) nb = limit;
    plan->algo = algo;
    plan->nblocks = nb;
    plan->threads = kThreads;
    plan->tile_bytes = kLLTileBytes;
    plan->ntiles = int(ceil_div(bytes ? bytes : 1, uint64_t(kLLTileBytes) * nb));
    plan->depth = PDCC_RING_DEPTH;
    plan->slot_stride = uint32_t(kLLSlotBytes);
    return PDCC_OK;
  }
  if (op.coll == PDCC_ALL_REDUCE) {
    const bool nvls = mc_ok && nvls_reducible(op.dtype, redop);
    if (algo == PDCC_ALGO_AUTO) {
      // One-shot moves N x bytes into every GPU but needs a single flag round; two-shot
      // moves ~2 x bytes and needs two. Crossover measured on 8xB200 (profiles/).
      // Link bytes per GPU and direction: one-shot push (n-1)S, two-shot push 2S(n-1)/n,
      // NVLS two-shot S(1 + 1/n) (the switch reads all of my data once; I receive my
      // shard and everyone's broadcast). At n = 2 one-shot moves no more than two-shot and
      // has one flag round less, so it wins at every size; the switch wins from n = 4 up.
      if (n <= 2) {
        algo = PDCC_ALGO_ONESHOT_P2P;
      } else if (bytes <= uint64_t(plan_env().oneshot_p2p_max)) {
        // one multimem.ld_reduce per unit instead of n-1 pushed copies: measured at
        // 8 GPUs, 256 KiB: 17.4 us vs 20.7 us (profiles/sweep_n8_v1.jsonl, graph-timed)
        algo = (nvls && n > plan_env().nvls_min_ranks) ? PDCC_ALGO_ONESHOT_NVLS
                                                               : PDCC_ALGO_ONESHOT_P2P;
      } else {
        algo = (nvls && n > plan_env().nvls_min_ranks) ? PDCC_ALGO_TWOSHOT_NVLS
                                                               : PDCC_ALGO_TWOSHOT_P2P;
      }
    }
    if ((algo == PDCC_ALGO_ONESHOT_NVLS || algo == PDCC_ALGO_TWOSHOT_NVLS) && !nvls) {
      set_error("NVLS all_reduce not available for dtype %d op %d (multicast %s)", op.dtype,
                redop, mc_ok ? "mapped" : "absent");
      return PDCC_EUNSUPPORTED;
    }
    if (algo < PDCC_ALGO_ONESHOT_P2P || algo > PDCC_ALGO_TWOSHOT_NVLS) {
      set_error("algo %d is not an all_reduce algorithm", algo);
      return PDCC_EINVAL;
    }
  } else if (op.coll == PDCC_REDUCE) {
    const bool nvls = mc_ok && nvls_reducible(op.dtype, redop);
    if (algo == PDCC_ALGO_ONESHOT_P2P) algo = PDCC_ALGO_P2P;
    if (algo == PDCC_ALGO_ONESHOT_NVLS) algo = PDCC_ALGO_NVLS;
    if (algo == PDCC_ALGO_AUTO) {
      // Root ingress: NVLS S (the switch reduces), one-shot push (n-1)S, two-phase push
      // (reduce-scatter among all ranks, then the owners hand their chunk to the root)
      // 2S(n-1)/n. Two-phase only pays from n = 3 and once the extra flag round is amortised.
      // A single GPU pulling everything through multimem.ld_reduce tops out near half the
      // link rate (profiles/). Large messages: two-shot push where only the non-root ranks
      // own chunks: S bytes in and out of every GPU, all of it posted stores.
      const uint64_t big = uint64_t(plan_env().rooted_twoshot_min);
      // Large: every rank multimem.ld_reduce's 1/n of each tile and hands it to the root
      // (measured 4 GPUs 256 MiB: 560 us against 617 us for the push variant and 433 us for
      // NCCL; 8 GPUs: 595 against 684). Without the switch: the push variant.
      if (n > 2 && bytes >= big) algo = (nvls && n > plan_env().nvls_min_ranks) ? PDCC_ALGO_TWOSHOT_NVLS
                                                                               : PDCC_ALGO_TWOSHOT_P2P;
      else if (nvls && n >= 4) algo = PDCC_ALGO_NVLS;
      else algo = PDCC_ALGO_P2P;
    }
    if (algo == PDCC_ALGO_TWOSHOT_NVLS && !nvls) {
      set_error("NVLS reduce not available for dtype %d op %d", op.dtype, redop);
      return PDCC_EUNSUPPORTED;
    }
    if (algo == PDCC_ALGO_NVLS && !nvls) {
      set_error("NVLS reduce not available for dtype %d op %d", op.dtype, redop);
      return PDCC_EUNSUPPORTED;
    }
    if (algo != PDCC_ALGO_P2P && algo != PDCC_ALGO_NVLS && algo != PDCC_ALGO_TWOSHOT_P2P &&
        algo != PDCC_ALGO_TWOSHOT_NVLS) {
      set_error("algo %d is not a reduce algorithm", algo);
      return PDCC_EINVAL;
    }
  } else if (op.coll == PDCC_BROADCAST || op.coll == PDCC_ALL_GATHER) {
    if (algo == PDCC_ALGO_AUTO) {
      if (op.coll == PDCC_BROADCAST) {
        // One GPU cannot feed the multicast engine (or n-1 peers) at link rate; from a few
        // MB up the root deals shards to the other ranks and they re-send them among
        // themselves (two-phase, unicast stores, S bytes per GPU and direction).
        const uint64_t big = uint64_t(plan_env().rooted_twoshot_min);
        if (n <= 2) algo = PDCC_ALGO_P2P;
        else if (bytes >= big) algo = PDCC_ALGO_TWOSHOT_P2P;  // 4 GPUs 256 MiB: 477-494 us, NVLS 511
        else algo = mc_ok ? PDCC_ALGO_NVLS : PDCC_ALGO_P2P;
      } else {
        // Posted stores beat remote loads on NVLink (profiles/): push my shard into every
        // peer's slot. From n = 4 one multimem.st replaces the n-1 stores (the echo of my
        // own shard costs 1/n extra ingress, which is what makes it lose at n = 2).
        algo = (mc_ok && n >= 4) ? PDCC_ALGO_NVLS : PDCC_ALGO_P2P;
      }
    }
    if ((algo == PDCC_ALGO_NVLS || algo == PDCC_ALGO_TWOSHOT_NVLS) && !mc_ok) {
      set_error("multicast not mapped for this comm");
      return PDCC_EUNSUPPORTED;
    }
    if (algo != PDCC_ALGO_P2P && algo != PDCC_ALGO_NVLS &&
        !(algo == PDCC_ALGO_PULL && op.coll == PDCC_ALL_GATHER) &&
        !((algo == PDCC_ALGO_TWOSHOT_P2P || algo == PDCC_ALGO_TWOSHOT_NVLS) &&
          op.coll == PDCC_BROADCAST)) {
      set_error("algo %d is not valid for collective %d", algo, op.coll);
      return PDCC_EINVAL;
    }
  } else {
    algo = PDCC_ALGO_P2P;
  }

  // the root-owns-nothing two-phase variants need at least one other rank to own chunks
  if (n < 2 && (op.coll == PDCC_REDUCE || op.coll == PDCC_BROADCAST) &&
      (algo == PDCC_ALGO_TWOSHOT_P2P || algo == PDCC_ALGO_TWOSHOT_NVLS))
    algo = PDCC_ALGO_P2P;

  // ---- grid + tile + ring geometry -------------------------------------------------
  const bool reduce_coll = op.coll == PDCC_ALL_REDUCE || op.coll == PDCC_REDUCE;
  const bool oneshot_push = reduce_coll && (algo == PDCC_ALGO_ONESHOT_P2P || algo == PDCC_ALGO_P2P);
  const bool twoshot_push = reduce_coll && algo == PDCC_ALGO_TWOSHOT_P2P;
  const bool sub_slots = op.coll == PDCC_GATHER || oneshot_push ||
                         (op.coll == PDCC_ALL_GATHER && algo != PDCC_ALGO_PULL);
  const bool bcast_2p = op.coll == PDCC_BROADCAST &&
                        (algo == PDCC_ALGO_TWOSHOT_P2P || algo == PDCC_ALGO_TWOSHOT_NVLS);
  // chunk owners: every rank, or every rank but the root (rooted two-shot push, two-phase
  // broadcast), see ReducePolicy / MovePolicy
  const uint64_t nown = ((twoshot_push && op.coll == PDCC_REDUCE) || bcast_2p) ? uint64_t(n - 1)
                                                                               : uint64_t(n);
  uint64_t align = 16;
  if (twoshot_push || bcast_2p || (reduce_coll && algo == PDCC_ALGO_TWOSHOT_NVLS))
    align = 16 * (nown ? nown : 1);
  // payload one CTA may place in one slot of `slot` bytes for this collective
  auto payload_cap = [&](uint64_t slot) -> uint64_t {
    uint64_t cap = slot;
    if (sub_slots) cap = (slot / n) & ~uint64_t(15);                       // one sub-slot per source rank
    if (twoshot_push) cap = slot * nown / (uint64_t(n) + nown);           // n partials + reduced chunks
    return cap / align * align;
  };

  const PlanEnv& pe = plan_env();
  if (ring_limit <= 0 || ring_limit > PDCC_MAX_BLOCKS) ring_limit = PDCC_MAX_BLOCKS;
  int max_blocks = pe.max_ctas > 0 ? pe.max_ctas : (pe.default_ctas > 0 ? pe.default_ctas : 128);
  if (max_blocks > ring_limit) max_blocks = ring_limit;

  int nb = op.nblocks;
  if (nb <= 0) {
    const uint64_t per_cta = 32 << 10;  // below this a CTA is latency- not bandwidth-bound
    nb = int(ceil_div(bytes ? bytes : 1, per_cta));
    // big messages: every SM takes part (the stage groups share each SM's warps)
    if (bytes >= (uint64_t(8) << 20) && pe.max_ctas <= 0 && pe.default_ctas <= 0) max_blocks = ring_limit;
    if (nb > max_blocks) nb = max_blocks;
    if (nb < 1) nb = 1;
  }
  if (nb > ring_limit) {
    set_error("nblocks %d exceeds the %d provisioned rings", nb, ring_limit);
    return PDCC_EINVAL;
  }

  // Tile: small enough that a message is tens of tiles per CTA (short pipeline fill) and
  // that `depth` of them fit in the ring; big enough that the per-tile fence + flag of a
  // stage group (a few microseconds, overlapped by the other groups) stays cheap.
  const uint64_t per_cta = ceil_div(bytes ? bytes : 1, uint64_t(nb));
  uint64_t want = pe.tile_bytes > 0 ? uint64_t(pe.tile_bytes) : default_tile(op.coll, algo, n, bytes);
  uint64_t tile = per_cta < want ? per_cta : want;
  tile = round_up(tile, align);
  // Depth before tile size: the publisher's fence delays every flag by 10-25 us, so a ring
  // needs about eight tiles in flight to keep its IN stage busy (8 GPUs, gather with 64 KiB
  // tiles and depth 2: 0.45 ms; round 1 with depth 4: 0.35 ms). Layouts that put n sub-slots
  // in a slot (gather, all_gather, one-shot) therefore take smaller tiles instead of fewer.
  const int max_depth = pe.depth > 0 ? pe.depth : 16;
  int depth = 1;
  uint64_t slot = 0;
  if (pe.tile_bytes <= 0) {
    for (int want_depth = 8; want_depth >= 1; want_depth >>= 1) {
      if (want_depth > max_depth) continue;
      const uint64_t sl = (uint64_t(comm.ring_bytes) / uint64_t(want_depth)) & ~uint64_t(255);
      const uint64_t cap = payload_cap(sl);
      const uint64_t floor_tile = uint64_t(want_depth >= 4 ? (8 << 10) : 1);
      if (cap >= tile || cap >= floor_tile || want_depth == 1) {
        if (cap < tile) tile = cap;
        break;
      }
    }
  }
  for (int d = max_depth; d >= 1; d >>= 1) {
    const uint64_t sl = (uint64_t(comm.ring_bytes) / uint64_t(d)) & ~uint64_t(255);
    if (payload_cap(sl) >= tile || d == 1) {
      depth = d;
      slot = sl;
      break;
    }
  }
  if (payload_cap(slot) < tile) tile = payload_cap(slot);
  if (tile < align) {
    set_error("ring of %u bytes too small for %d ranks", comm.ring_bytes, n);
    return PDCC_EINVAL;
  }
  if (op.coll == PDCC_BARRIER) tile = align;

  plan->algo = algo;
  plan->nblocks = nb;
  // three-stage collectives share an SM between three warp groups: give them more warps
  const bool three = (reduce_coll && (algo == PDCC_ALGO_TWOSHOT_P2P || algo == PDCC_ALGO_TWOSHOT_NVLS)) || bcast_2p;
  plan->threads = pe.threads > 0 ? pe.threads : (three ? kMaxThreads : kThreads);
  if (plan->threads > kMaxThreads) plan->threads = kMaxThreads;
  plan->threads = plan->threads / 32 * 32;
  if (plan->threads < 96) plan->threads = 96;
  plan->tile_bytes = tile;
  plan->ntiles = int(ceil_div(bytes ? bytes : 1, tile * uint64_t(nb)));
  if (plan->ntiles < 1) plan->ntiles = 1;
  plan->depth = depth;
  plan->slot_stride = uint32_t(slot);
  return PDCC_OK;
}

int launch_collective(const RankArgs* ranks, int nlocal, const RankArgs* dev_ranks,
                      const pdcc_op_t& op, const Plan& plan, void* stream) {
  const DevComm& comm = ranks[0].comm;
  const int n = comm.nranks;
  const size_t esz = dtype_size(op.dtype);
  const int redop = op.redop == PDCC_AVG ? PDCC_SUM : op.redop;

  KArgs k{};
  k.r = ranks[0];
  k.emu = nullptr;
  k.bytes = op.count * esz;
  k.tile_bytes = plan.tile_bytes;
  k.stride = uint32_t(round_up(plan.tile_bytes, 16));
  k.ntiles = plan.ntiles;
  k.depth = plan.depth;
  k.slot_stride = plan.slot_stride;
  k.root = op.root;
  k.trace = plan_env().trace;
  k.scale = op.scale == 0.0f ? 1.0f : op.scale;
  if (op.redop == PDCC_AVG) k.scale *= 1.0f / float(n);

  if (op.root < 0 || op.root >= n) {
    set_error("root %d out of range for %d ranks", op.root, n);
    return PDCC_EINVAL;
  }

  KernelFn fn = nullptr;
  switch (op.coll) {
    case PDCC_ALL_REDUCE: {
      const bool wide = op.out_dtype != op.dtype;
      if (wide && !(op.out_dtype == PDCC_F32 && (op.dtype == PDCC_BF16 || op.dtype == PDCC_F16))) {
        set_error("fused cast supports bf16/f16 -> f32 only");
        return PDCC_EINVAL;
      }
      fn = reduce_kernel_for(op.dtype, wide, redop,
                             plan.algo == PDCC_ALGO_LL ? AR_LL : plan.algo - PDCC_ALGO_ONESHOT_P2P,
                             false);
      break;
    }
    case PDCC_REDUCE:
      if (op.out_dtype != op.dtype) {
        set_error("reduce has no fused cast");
        return PDCC_EINVAL;
      }
      fn = reduce_kernel_for(op.dtype, false, redop,
                             plan.algo == PDCC_ALGO_LL            ? AR_LL
                             : plan.algo == PDCC_ALGO_NVLS        ? AR_ONESHOT_NVLS
                             : plan.algo == PDCC_ALGO_TWOSHOT_P2P ? AR_TWOSHOT_P2P
                             : plan.algo == PDCC_ALGO_TWOSHOT_NVLS ? AR_TWOSHOT_NVLS
                                                                  : AR_ONESHOT_P2P,
                             true);
      break;
    case PDCC_BROADCAST:
      fn = plan.algo == PDCC_ALGO_LL ? ll_move_kernel_for(op.coll)
           : move_kernel_for(plan.algo == PDCC_ALGO_NVLS           ? MV_BROADCAST_NVLS
                             : plan.algo == PDCC_ALGO_TWOSHOT_NVLS ? MV_BROADCAST_2P_NVLS
                             : plan.algo == PDCC_ALGO_TWOSHOT_P2P  ? MV_BROADCAST_2P_P2P
                                                                   : MV_BROADCAST_P2P);
      break;
    case PDCC_SCATTER:
      fn = plan.algo == PDCC_ALGO_LL ? ll_move_kernel_for(op.coll) : move_kernel_for(MV_SCATTER);
      break;
    case PDCC_GATHER:
      fn = plan.algo == PDCC_ALGO_LL ? ll_move_kernel_for(op.coll) : move_kernel_for(MV_GATHER);
      break;
    case PDCC_ALL_GATHER:
      fn = plan.algo == PDCC_ALGO_LL ? ll_move_kernel_for(op.coll)
           : move_kernel_for(plan.algo == PDCC_ALGO_NVLS   ? MV_ALLGATHER_NVLS
                             : plan.algo == PDCC_ALGO_PULL ? MV_ALLGATHER_PULL
                                                           : MV_ALLGATHER_PUSH);
      break;
    case PDCC_BARRIER: fn = move_kernel_for(MV_BARRIER); break;
    default: break;
  }
  if (!fn) {
    set_error("no kernel for collective %d dtype %d algo %d", op.coll, op.dtype, plan.algo);
    return PDCC_EUNSUPPORTED;
  }

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t err;
  if (nlocal == 1) {
    fn<<<dim3(plan.nblocks, 1, 1), dim3(plan.threads, 1, 1), 0, st>>>(k);
    err = cudaGetLastError();
  } else {
    // Emulation: the per-rank args live in device memory (dev_ranks); all
    // nlocal x nblocks CTAs wait on each other, so they must be co-resident, which is
    // exactly what a cooperative launch guarantees (or refuses with an error).
    k.emu = dev_ranks;
    void* params[] = {&k};
    err = cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(fn),
                                      dim3(plan.nblocks, nlocal, 1), dim3(plan.threads, 1, 1),
                                      params, 0, st);
  }
  if (err != cudaSuccess) {
    set_error("kernel launch failed: %s", cudaGetErrorString(err));
    return PDCC_ECUDA;
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return PDCC_OK;
}

// ---------------------------------------------------------------------------------------
// Per-rank arguments of the single-GPU emulation, written to device memory by a kernel that
// receives them BY VALUE: the launch copies them at enqueue time, so neither a later
// overwrite of the host array nor a replay of a captured launch can read stale host memory.
// ---------------------------------------------------------------------------------------
struct RankArgsPack {
  RankArgs r[PDCC_MAX_RANKS];
};
__global__ void rank_args_kernel(const __grid_constant__ RankArgsPack pack, RankArgs* dst, int n) {
  constexpr int kWords = int(sizeof(RankArgs) / 4);
  const uint32_t* src = reinterpret_cast<const uint32_t*>(&pack);
  uint32_t* d = reinterpret_cast<uint32_t*>(dst);
  for (int i = threadIdx.x; i < n * kWords; i += blockDim.x) d[i] = src[i];
}

int upload_rank_args(const RankArgs* host, int n, RankArgs* dev, void* stream) {
  if (n < 1 || n > PDCC_MAX_RANKS || !host || !dev) return PDCC_EINVAL;
  RankArgsPack pack;
  memset(&pack, 0, sizeof(pack));
  for (int r = 0; r < n; ++r) pack.r[r] = host[r];
  rank_args_kernel<<<1, 128, 0, static_cast<cudaStream_t>(stream)>>>(pack, dev, n);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("uploading per-rank arguments: %s", cudaGetErrorString(e));
    return PDCC_ECUDA;
  }
  return PDCC_OK;
}

// ---------------------------------------------------------------------------------------
// Zero-copy collectives on symmetric memory (sym.cuh)
// ---------------------------------------------------------------------------------------
SymKernelFn sym_reduce_kernel_for(int dtype, int redop, bool rooted) {
  switch (dtype) {
    case PDCC_F32: return sym_table_f32(redop, rooted);
    case PDCC_BF16: return sym_table_bf16(redop, rooted);
    case PDCC_F16: return sym_table_f16(redop, rooted);
    case PDCC_F64: return sym_table_f64(redop, rooted);
    case PDCC_I32: return sym_table_i32(redop, rooted);
    case PDCC_I64: return sym_table_i64(redop, rooted);
    case PDCC_U8: return sym_table_u8(redop, rooted);
    default: return nullptr;
  }
}

uint64_t sym_min_bytes(int nranks) {
  const PlanEnv& pe = plan_env();
  return uint64_t(pe.ll_max >= 0 ? pe.ll_max : (512 << 10) / (nranks > 0 ? nranks : 1));
}

bool sym_supported(int coll, int dtype, int redop) {
  if (redop == PDCC_AVG) redop = PDCC_SUM;
  switch (coll) {
    case PDCC_ALL_REDUCE: return sym_reduce_kernel_for(dtype, redop, false) != nullptr;
    case PDCC_REDUCE: return sym_reduce_kernel_for(dtype, redop, true) != nullptr;
    case PDCC_BROADCAST: case PDCC_ALL_GATHER: return true;
    default: return false;
  }
}

int launch_sym(const DevComm& comm, int ring_limit, const pdcc_op_t& op, const SymTensor& in,
               const SymTensor& out, void* stream) {
  const int n = comm.nranks;
  const uint64_t bytes = op.count * dtype_size(op.dtype);
  if (comm.mc == nullptr || (bytes & 15u) != 0 || bytes == 0) {
    set_error("zero-copy path needs a multicast mapping and a multiple of 16 bytes");
    return PDCC_EINVAL;
  }
  if (ring_limit <= 0 || ring_limit > PDCC_MAX_BLOCKS) ring_limit = PDCC_MAX_BLOCKS;
  SymArgs a;
  memset(&a, 0, sizeof(a));
  a.comm = comm;
  a.in_local = in.local;
  a.in_mc = in.mc;
  a.out_local = out.local;
  a.out_mc = out.mc;
  for (int p = 0; p < PDCC_MAX_RANKS; ++p) a.out_peer[p] = out.peer[p];
  a.bytes = bytes;
  a.root = op.root;
  a.scale = op.scale == 0.0f ? 1.0f : op.scale;
  if (op.redop == PDCC_AVG) a.scale *= 1.0f / float(n);
  const PlanEnv& pe = plan_env();

  SymKernelFn fn = nullptr;
  uint64_t work = bytes;  // bytes one rank moves: sizes the grid
  switch (op.coll) {
    case PDCC_ALL_REDUCE:
      fn = sym_reduce_kernel_for(op.dtype, op.redop == PDCC_AVG ? PDCC_SUM : op.redop, false);
      work = bytes / n;
      break;
    case PDCC_REDUCE:
      fn = sym_reduce_kernel_for(op.dtype, op.redop == PDCC_AVG ? PDCC_SUM : op.redop, true);
      work = bytes / n;
      break;
    case PDCC_ALL_GATHER: fn = sym_all_gather_kernel; break;
    case PDCC_BROADCAST: {
      const uint64_t big = uint64_t(pe.rooted_twoshot_min);
      if (n > 2 && bytes >= big && op.algo != PDCC_ALGO_NVLS) {
        fn = sym_broadcast2_kernel;
        const uint64_t align = 16 * uint64_t(n - 1);
        uint64_t tile = pe.tile_bytes > 0 ? uint64_t(pe.tile_bytes) : (uint64_t(128) << 10);
        a.tile_bytes = (tile + align - 1) / align * align;
      } else {
        fn = sym_broadcast_kernel;
      }
      break;
    }
    default: break;
  }
  if (!fn) {
    set_error("no zero-copy kernel for collective %d dtype %d op %d", op.coll, op.dtype, op.redop);
    return PDCC_EUNSUPPORTED;
  }
  int nb = op.nblocks > 0 ? op.nblocks : int(ceil_div(work, uint64_t(16) << 10));
  int cap = pe.max_ctas > 0 ? pe.max_ctas : ring_limit;
  if (cap > ring_limit) cap = ring_limit;
  if (nb > cap) nb = cap;
  if (nb < 1) nb = 1;
  if (fn == sym_broadcast2_kernel) {
    const uint64_t tiles = ceil_div(bytes, a.tile_bytes);
    if (uint64_t(nb) > tiles) nb = int(tiles);
  }
  fn<<<dim3(nb), dim3(kThreads), 0, static_cast<cudaStream_t>(stream)>>>(a);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) {
    set_error("zero-copy kernel launch failed: %s", cudaGetErrorString(err));
    return PDCC_ECUDA;
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return PDCC_OK;
}

bool reduce_supported(int dtype, int redop) {
  return reduce_kernel_for(dtype, false, redop == PDCC_AVG ? PDCC_SUM : redop, AR_ONESHOT_P2P,
                           false) != nullptr;
}

// ---------------------------------------------------------------------------------------
// K9 local helpers: the world-size-1 degenerate collectives (a list op over one rank is a
// copy) and the DP demo's scale+cast when there is nobody to reduce with.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 2) local_copy_kernel(char* dst, const char* src,
                                                              uint64_t bytes) {
  // each CTA takes a contiguous span so its loads stream
  const uint64_t per = round_up_dev(ceil_div_dev(bytes, gridDim.x), 16);
  const uint64_t lo = per * blockIdx.x;
  if (lo >= bytes) return;
  const uint64_t len = bytes - lo < per ? bytes - lo : per;
  Ctx c{};
  c.tid = threadIdx.x;
  c.nt = blockDim.x;
  c.bar = 0;
  copy_bytes<LD_USER, 8>(c, dst + lo, src + lo, len);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) scale_cast_kernel(void* dst, const T* src,
                                                              uint64_t count, float scale,
                                                              bool wide) {
  using Tr = Traits<T>;
  const uint64_t n16 = (count + Tr::K - 1) / Tr::K;
  const bool vec = aligned16(src);
  for (uint64_t j = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; j < n16;
       j += uint64_t(gridDim.x) * blockDim.x) {
    const uint64_t e = j * Tr::K;
    const int valid = count - e < uint64_t(Tr::K) ? int(count - e) : Tr::K;
    Acc<T> acc;
    if (vec && valid == Tr::K) {
      acc.init(ld_stream_v4(src + e));
    } else {
      for (int i = 0; i < Tr::K; ++i) acc.v[i] = i < valid ? Tr::up(src[e + i]) : typename Tr::C(0);
    }
    acc.scale(scale);
    if (wide) store_unit<T, true>(dst, e, acc, valid);
    else store_unit<T, false>(dst, e, acc, valid);
  }
}

}  // namespace pdcc

using namespace pdcc;

extern "C" {

uint64_t pdcc_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

void pdcc_reload_env(void) {
  PlanEnv e;
  read_plan_env(&e);
  (void)plan_env();
  g_plan_env = e;
}

static int set_device_checked(int device) {
  if (device < 0) return PDCC_OK;
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) {
    set_error("cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    return PDCC_ECUDA;
  }
  return PDCC_OK;
}

int pdcc_local_copy(void* dst, const void* src, uint64_t bytes, int device, void* stream) {
  if (bytes == 0 || dst == src) return PDCC_OK;
  if (!dst || !src) return PDCC_EINVAL;
  int rc = set_device_checked(device);
  if (rc) return rc;
  cudaError_t e;
  const bool tma_ok = ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src) | bytes) & 15u) == 0 &&
                      bytes >= (uint64_t(1) << 20) && plan_env().copy_tma != 0;
  if (tma_ok) {
    // big aligned copies go through the TMA engine (tma_copy.cuh), one CTA per SM
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      e = cudaFuncSetAttribute(tma_copy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               int(kTmaMaxSmemBytes));
      if (e != cudaSuccess) {
        set_error("tma_copy attribute: %s", cudaGetErrorString(e));
        return PDCC_ECUDA;
      }
      attr_set[dev] = true;
    }
    const PlanEnv& pe = plan_env();
    uint32_t chunk = uint32_t(pe.copy_chunk_kb > 0 ? pe.copy_chunk_kb : 16) << 10;
    int stages = pe.copy_stages > 1 ? pe.copy_stages : 2;
    if (stages > kTmaMaxStages) stages = kTmaMaxStages;
    while (uint64_t(chunk) * stages > kTmaMaxSmemBytes && stages > 2) --stages;
    if (uint64_t(chunk) * stages > kTmaMaxSmemBytes) chunk = (kTmaMaxSmemBytes / stages) & ~uint32_t(1023);
    const uint64_t per_sm = uint64_t(pe.copy_ctas_per_sm > 0 ? pe.copy_ctas_per_sm : 1);
    uint64_t chunks = (bytes + chunk - 1) / chunk;
    uint64_t blocks = chunks < per_sm * 148 ? chunks : per_sm * 148;
    tma_copy_kernel<<<dim3(unsigned(blocks)), dim3(32), size_t(chunk) * stages,
                      static_cast<cudaStream_t>(stream)>>>(static_cast<char*>(dst),
                                                           static_cast<const char*>(src), bytes,
                                                           chunk, stages);
  } else {
    uint64_t blocks = (bytes + (64 << 10) - 1) / (64 << 10);
    if (blocks > 4 * 148) blocks = 4 * 148;
    local_copy_kernel<<<dim3(unsigned(blocks)), dim3(kThreads), 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<char*>(dst), static_cast<const char*>(src), bytes);
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("local_copy launch: %s", cudaGetErrorString(e));
    return PDCC_ECUDA;
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return PDCC_OK;
}

int pdcc_local_scale_cast(void* dst, const void* src, uint64_t count, int dtype, int out_dtype,
                          float scale, int device, void* stream) {
  if (count == 0) return PDCC_OK;
  if (!dst || !src) return PDCC_EINVAL;
  const bool wide = out_dtype != dtype;
  if (wide && !(out_dtype == PDCC_F32 && (dtype == PDCC_BF16 || dtype == PDCC_F16))) {
    set_error("scale_cast supports bf16/f16 -> f32 only");
    return PDCC_EINVAL;
  }
  int rc = set_device_checked(device);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t esz = dtype_size(dtype);
  uint64_t units = (count * esz + 15) / 16;
  uint64_t blocks = (units + kThreads - 1) / kThreads;
  if (blocks > 8 * 148) blocks = 8 * 148;
  dim3 g{unsigned(blocks)}, b{unsigned(kThreads)};
  switch (dtype) {
    case PDCC_F32: scale_cast_kernel<float><<<g, b, 0, st>>>(dst, static_cast<const float*>(src), count, scale, false); break;
    case PDCC_BF16: scale_cast_kernel<__nv_bfloat16><<<g, b, 0, st>>>(dst, static_cast<const __nv_bfloat16*>(src), count, scale, wide); break;
    case PDCC_F16: scale_cast_kernel<__half><<<g, b, 0, st>>>(dst, static_cast<const __half*>(src), count, scale, wide); break;
    case PDCC_F64: scale_cast_kernel<double><<<g, b, 0, st>>>(dst, static_cast<const double*>(src), count, scale, false); break;
    default:
      set_error("scale_cast: unsupported dtype %d", dtype);
      return PDCC_EINVAL;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("scale_cast launch: %s", cudaGetErrorString(e));
    return PDCC_ECUDA;
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return PDCC_OK;
}

}  // extern "C"
