// Internal declarations shared by store.cpp, runtime.cu and kernels.cu.
#pragma once

#include <stdarg.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>

#include "../include/pdcc.h"

namespace pdcc {

// thread-local error string behind pdcc_last_error()
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));

// ---- layout of one rank's symmetric block -------------------------------------------
//
//   [0, kCtrlBytes)                 ctrl   local-only bookkeeping (epochs, error record)
//   [kCtrlBytes, kHeaderBytes)      pads   signal flags written by peers
//   [kHeaderBytes, +kLLBytes)       LL     kLLRings x PDCC_RING_DEPTH flag-in-data slots
//   [kDataOffset, +rings)           rings  max_blocks x ring_bytes; every call cuts its ring
//                                          into `depth` slots sized for its own tiles
//   [.., +heap_bytes)               heap   symmetric heap for zero-copy user tensors
//
// Every CTA index b owns ring b on every rank: slot(b, pos) sits at the same offset in
// every rank's block, so one offset addresses "the same tile" on all peers and on the
// multicast alias. A ring is only ever touched by CTAs with that index, which is what
// lets per-CTA flag counters order every reuse of a slot across calls.
// Slot geometry is per call (plan_collective): a call's first `depth` tiles wait until
// every rank has finished ALL earlier tiles of the ring, so two calls with different slot
// sizes never overlap inside a ring.
constexpr size_t kCtrlBytes = 64 * 1024;
constexpr int kNumChannels = 3;  // READY, MID, DONE
constexpr size_t kFlagWords = size_t(kNumChannels) * PDCC_MAX_BLOCKS * PDCC_MAX_RANKS;
// Arrival counters of the zero-copy kernels (multimem.red.add.u32 targets): they sit behind
// the flag channels inside the pad region, so that every rank can add to every rank's copy
// through the multicast alias. Word kSymArriveWord + 2*b + phase of the pad region.
constexpr size_t kSymArriveWord = kFlagWords;
// Progress counters of the two-phase zero-copy broadcast: word kSymProgWord + 2*b + half.
constexpr size_t kSymProgWord = kSymArriveWord + 2 * PDCC_MAX_BLOCKS;
constexpr size_t kPadBytes = (kFlagWords + 4 * PDCC_MAX_BLOCKS) * sizeof(uint32_t);
constexpr size_t kHeaderBytes = 2 * 1024 * 1024;  // keeps rings 2 MiB aligned
static_assert(kCtrlBytes + kPadBytes <= kHeaderBytes, "header too small");
// LL (low-latency) slots hold {data, flag, data, flag} packets only, never raw payload, so
// a stale word can never be mistaken for the current flag. They stay zero-initialised
// until first use and are only used by the first kLLRings CTA indices.
constexpr int kLLRings = 16;
constexpr size_t kLLSlotBytes = 256 * 1024;
constexpr size_t kLLBytes = size_t(kLLRings) * PDCC_RING_DEPTH * kLLSlotBytes;  // 16 MiB
constexpr size_t kDataOffset = kHeaderBytes + kLLBytes;
constexpr size_t kZeroBytes = kDataOffset;  // what comm creation must clear

// ctrl words (uint32 index)
constexpr int kCtrlErr = 0;        // first device-side timeout record: code
constexpr int kCtrlErrInfo = 1;    // .. chan | tile | peer packed
constexpr int kCtrlEpoch = 64;     // + b : tiles processed so far by CTA index b
constexpr int kCtrlSymEpoch = 256;  // + b : zero-copy (symmetric heap) calls made by CTA index b
// Optional timeline of CTA 0 (PDCC_TRACE=1): word kCtrlTrace counts the events, 8-word
// records {kind, tile, begin_lo, begin_hi, end_lo, end_hi, aux0, aux1} follow at +4.
// kind: 0..2 worker of that stage (| warp << 8): begin = started waiting for admission,
// end = tile finished, aux0 = ns spent waiting; 16 publisher: the fence, aux0 = tiles done
// per stage (10 bits each); 32..34 admitter: the moment tile `tile` of that stage was admitted.
constexpr int kCtrlTrace = 4096;
constexpr int kCtrlTraceCap = (int(kCtrlBytes / 4) - kCtrlTrace - 4) / 8;
constexpr int kCtrlSymProg = 512;   // + 2*b + half : tiles the two-phase zero-copy broadcast dealt so far

// What a kernel needs to know about the group, passed by value (88 bytes).
struct DevComm {
  int rank;
  int nranks;
  char* base[PDCC_MAX_RANKS];  // base[p] = local VA of rank p's symmetric block
  char* mc;                    // multicast alias of the block (nullptr: no NVLS)
  uint32_t ring_bytes;         // bytes of one CTA's ring (cut into slots per call)
  uint32_t timeout_ms;         // device-side wait budget
};

// One rank's view of one collective call.
struct RankArgs {
  DevComm comm;
  const void* in;
  void* out;
  void* list[PDCC_MAX_RANKS];
};

struct Plan {
  int algo;
  int nblocks;
  int threads;
  uint64_t tile_bytes;  // payload bytes per CTA per tile (see kernels.cuh)
  int ntiles;
  int depth;             // slots the ring is cut into for this call
  uint32_t slot_stride;  // bytes between consecutive slots of a ring
};

// launch.cu
// ring_limit: how many CTAs (= provisioned rings, capped by the SM count) a call may use.
int plan_collective(const DevComm& comm, int ring_limit, bool has_mc, const pdcc_op_t& op,
                    Plan* plan);
// Launches one kernel covering `nlocal` ranks (1 = real multi-GPU rank, >1 = emulation:
// cooperative launch with grid.y = nlocal, per-rank args read from the device array
// dev_ranks). All ranks must describe the same collective.
int launch_collective(const RankArgs* ranks, int nlocal, const RankArgs* dev_ranks,
                      const pdcc_op_t& op, const Plan& plan, void* stream);
size_t dtype_size(int dtype);
// Copies `n` per-rank argument records into device memory from KERNEL PARAMETERS (a tiny
// set-up kernel): nothing on the stream ever reads host memory, so the caller's array may
// live on the stack and the enqueued work stays valid after the caller returns.
int upload_rank_args(const RankArgs* host, int n, RankArgs* dev, void* stream);
// True when (dtype, redop) has an all_reduce kernel (P2P variants).
bool reduce_supported(int dtype, int redop);

// Zero-copy collectives on symmetric memory (sym.cuh). `in`/`out` are described by their
// local address plus the multicast alias / peer addresses of the same offset.
struct SymTensor {
  char* local = nullptr;
  char* mc = nullptr;
  char* peer[PDCC_MAX_RANKS] = {};
};
bool sym_supported(int coll, int dtype, int redop);
uint64_t sym_min_bytes(int nranks);  // at or below this size AUTO keeps the staged small-message kernels
int launch_sym(const DevComm& comm, int ring_limit, const pdcc_op_t& op, const SymTensor& in,
               const SymTensor& out, void* stream);

}  // namespace pdcc
