/*
 * pdcc.h — C ABI of libpdcc, the B200-native collective-communication runtime.
 *
 * Everything the Python package (pdcc_b200) does on a CUDA tensor goes through
 * the entry points declared here; they are plain C so that ctypes, cffi or a
 * C/C++ host program can call them without a torch build dependency.
 *
 * Conventions
 *   - Every function returns 0 on success and a negative PDCC_E* code on
 *     failure; pdcc_last_error() returns a thread-local human readable string.
 *   - `stream` arguments are cudaStream_t values passed as void* (0 = legacy
 *     default stream). All collectives are stream ordered: they enqueue one
 *     kernel on `stream` and return; nothing blocks the host.
 *   - Ranks inside a comm are *group* ranks (0..nranks-1). Translating the
 *     reference's global src/dst ranks to group ranks is done by the caller
 *     (pdcc_b200/distributed.py), as torch does in
 *     [torch] distributed/distributed_c10d.py:1183-1210.
 *
 * Reference parity: the reference (main.py) owns no native code; the pieces of
 * torch it drives transitively are listed in SURVEY.md §2.5. Each block below
 * names the transitive component it replaces.
 */
#ifndef PDCC_H_
#define PDCC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDCC_MAX_RANKS 8      /* one NVSwitch domain of an HGX B200 board      */
#define PDCC_MAX_BLOCKS 148   /* CTAs (= private tile rings) per collective: one per SM */
#define PDCC_RING_DEPTH 4     /* LL slots per CTA ring; a data ring is DEPTH x slot_bytes */

/* error codes */
#define PDCC_OK 0
#define PDCC_EINVAL (-1)
#define PDCC_ECUDA (-2)
#define PDCC_ESYS (-3)
#define PDCC_ETIMEOUT (-4)
#define PDCC_EUNSUPPORTED (-5)
#define PDCC_EPEER (-6) /* a device-side wait timed out (mismatched collective?) */

const char* pdcc_last_error(void);
const char* pdcc_version(void);

/* ------------------------------------------------------------------------ *
 * R1  Bootstrap key-value store (replaces torch TCPStore + PrefixStore,
 *     reached from main.py:92-94 through [torch] distributed/rendezvous.py).
 *     Rank 0 hosts a server thread; every rank (rank 0 included) is a client.
 * ------------------------------------------------------------------------ */
typedef struct pdcc_store_server pdcc_store_server_t;
typedef struct pdcc_store pdcc_store_t;

/* port 0 = pick a free port; the bound port is returned in *bound_port. */
int pdcc_store_server_start(const char* bind_addr, int port, pdcc_store_server_t** out,
                            int* bound_port);
int pdcc_store_server_stop(pdcc_store_server_t* srv);

int pdcc_store_connect(const char* host, int port, int timeout_ms, pdcc_store_t** out);
int pdcc_store_close(pdcc_store_t* st);
int pdcc_store_set(pdcc_store_t* st, const char* key, const void* val, int len);
/* Blocks until `key` exists or timeout_ms elapses. *len_out = value length
 * (may exceed cap: the value is truncated to cap bytes). */
int pdcc_store_get(pdcc_store_t* st, const char* key, void* buf, int cap, int* len_out,
                   int timeout_ms);
int pdcc_store_add(pdcc_store_t* st, const char* key, long long delta, long long* result);
int pdcc_store_check(pdcc_store_t* st, const char* key, int* exists);
int pdcc_store_delete(pdcc_store_t* st, const char* key);
/* Host barrier over the store: `name` must be unique per barrier instance. */
int pdcc_store_barrier(pdcc_store_t* st, const char* name, int nranks, int timeout_ms);

/* ------------------------------------------------------------------------ *
 * Capability probe (SURVEY.md §7.3 hard part 1): what handle kinds / NVLS
 * multicast the box offers. Safe to call with no GPU (returns PDCC_ECUDA).
 * ------------------------------------------------------------------------ */
typedef struct {
  int device;
  int sm_count;
  int cc_major, cc_minor;
  int vmm_supported;          /* cuMemCreate family                          */
  int posix_fd_supported;     /* CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR    */
  int fabric_supported;       /* CU_MEM_HANDLE_TYPE_FABRIC                   */
  int multicast_supported;    /* CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED     */
  int cooperative_launch;
  int device_count;
  size_t alloc_granularity;
  size_t mc_granularity;      /* recommended, for a 2-device object (0 = n/a) */
  size_t total_mem;
  char name[64];
} pdcc_caps_t;
int pdcc_probe(int device, pdcc_caps_t* caps);

/* ------------------------------------------------------------------------ *
 * R2/R3/R4  Communicator: symmetric arena + peer mapping + NVLS multicast +
 * signal pads for one group (replaces ProcessGroupNCCL / ncclCommInitRank,
 * reached from main.py:94 and every new_group call at main.py:11..75).
 * ------------------------------------------------------------------------ */
typedef struct pdcc_comm pdcc_comm_t;

#define PDCC_HANDLE_AUTO 0
#define PDCC_HANDLE_POSIX_FD 1  /* cuMemCreate + fd passed with SCM_RIGHTS   */
#define PDCC_HANDLE_LEGACY_IPC 2 /* cudaMalloc + cudaIpcGetMemHandle (no NVLS) */
#define PDCC_HANDLE_LOCAL 3     /* single-process comm (nranks == 1 or emu)  */

typedef struct {
  int rank;
  int nranks;
  int device;
  int handle_kind;        /* PDCC_HANDLE_*                                    */
  int multicast;          /* -1 auto, 0 never, 1 require                      */
  int timeout_ms;         /* host-side rendezvous timeout                     */
  int device_timeout_ms;  /* device-side flag wait timeout (0 = default 10 s) */
  size_t slot_bytes;      /* a CTA's ring is PDCC_RING_DEPTH x this; every call
                             cuts the ring into slots for its own tile size
                             (0 = default 512 KiB, i.e. 2 MiB rings)          */
  int max_blocks;         /* rings to provision (0 = PDCC_MAX_BLOCKS)         */
  size_t heap_bytes;      /* symmetric heap for zero-copy tensors (0 = none)  */
  char key[64];           /* unique per group: namespaces the store keys      */
} pdcc_comm_config_t;

typedef struct {
  int rank, nranks, device;
  int handle_kind;
  int has_multicast;
  int sm_count;
  int max_blocks;
  size_t slot_bytes;
  size_t block_bytes;     /* physical bytes of this rank's symmetric block    */
  size_t heap_bytes;
  void* heap_base;        /* local VA of the symmetric heap (NULL if none)    */
} pdcc_comm_info_t;

int pdcc_comm_create(pdcc_store_t* store, const pdcc_comm_config_t* cfg, pdcc_comm_t** out);
int pdcc_comm_destroy(pdcc_comm_t* comm);
int pdcc_comm_info(pdcc_comm_t* comm, pdcc_comm_info_t* info);
/* Tracing: with PDCC_TRACE=1 in the environment CTA 0 of every pipeline
 * kernel records a timeline of its stages (admission waits, tile work, publisher fences)
 * into the block; this copies it out (8 uint32 per event, layout in csrc/common.h) and
 * clears it. Synchronises the device. */
int pdcc_comm_read_trace(pdcc_comm_t* comm, uint32_t* out, int cap_events, int* count);
/* Synchronises the device and returns PDCC_EPEER if any kernel of this comm
 * hit a device-side wait timeout since the last check (details in last_error). */
int pdcc_comm_check(pdcc_comm_t* comm);

/* ------------------------------------------------------------------------ *
 * K1-K9  Collectives. One descriptor type for all of them; the named entry
 * points below fill it in. (Replaces Backend::{allreduce,reduce,broadcast,
 * scatter,gather,allgather}, [torch] include/.../c10d/Backend.hpp:140-382.)
 * ------------------------------------------------------------------------ */
enum pdcc_coll {
  PDCC_ALL_REDUCE = 0, /* main.py:23 */
  PDCC_REDUCE = 1,     /* main.py:14 */
  PDCC_BROADCAST = 2,  /* main.py:81 */
  PDCC_SCATTER = 3,    /* main.py:37,39 */
  PDCC_GATHER = 4,     /* main.py:52,54 */
  PDCC_ALL_GATHER = 5, /* main.py:68 */
  PDCC_BARRIER = 6,
  PDCC_NUM_COLLS = 7
};

enum pdcc_dtype {
  PDCC_F32 = 0, PDCC_BF16 = 1, PDCC_F16 = 2, PDCC_F64 = 3,
  PDCC_I32 = 4, PDCC_I64 = 5, PDCC_U8 = 6, PDCC_I8 = 7, PDCC_I16 = 8,
  PDCC_NUM_DTYPES = 9
};

/* numbering follows c10d ReduceOp ([torch] .../c10d/Types.hpp:37-48) */
enum pdcc_redop { PDCC_SUM = 0, PDCC_AVG = 1, PDCC_PRODUCT = 2, PDCC_MIN = 3, PDCC_MAX = 4 };

enum pdcc_algo {
  PDCC_ALGO_AUTO = 0,
  PDCC_ALGO_ONESHOT_P2P = 1,  /* every rank pulls every peer's tile and combines in registers */
  PDCC_ALGO_TWOSHOT_P2P = 2,  /* reduce-scatter + all-gather over peer pointers               */
  PDCC_ALGO_ONESHOT_NVLS = 3, /* multimem.ld_reduce of the whole tile on every rank           */
  PDCC_ALGO_TWOSHOT_NVLS = 4, /* multimem.ld_reduce own shard + multimem.st to all            */
  PDCC_ALGO_P2P = 5,          /* rooted / gather-type collectives: posted peer stores (push)  */
  PDCC_ALGO_NVLS = 6,         /* rooted / gather-type collectives: multicast                  */
  PDCC_ALGO_LL = 7,           /* any collective, small messages: flag-in-data push, no fences */
  PDCC_ALGO_PULL = 8,         /* all_gather: stage own shard, read every peer's (remote loads)  */
  PDCC_ALGO_SYM = 9           /* zero-copy on symmetric memory (pdcc_sym_alloc): multimem ops
                                 straight on the user's tensor, two barriers per call          */
};

typedef struct {
  int coll;       /* enum pdcc_coll                                           */
  int dtype;      /* enum pdcc_dtype of the tensors                           */
  int out_dtype;  /* all_reduce only: PDCC_F32 with a bf16/f16 input gives the
                     fused cast of the DP demo; otherwise == dtype            */
  int redop;      /* enum pdcc_redop (AVG = SUM with scale 1/nranks)          */
  int root;       /* group rank of src/dst for rooted collectives            */
  int algo;       /* enum pdcc_algo                                           */
  int nblocks;    /* CTAs, 0 = auto                                           */
  float scale;    /* all_reduce/reduce: result multiplied by this (1 = none)  */
  uint64_t count; /* elements per tensor (per list element for list ops)      */
  const void* in; /* input tensor (all_reduce/reduce/broadcast: may == out)   */
  void* out;      /* output tensor                                            */
  void* list[PDCC_MAX_RANKS]; /* scatter: inputs on root; gather/all_gather:
                                 outputs. One pointer per group rank.         */
} pdcc_op_t;

/* ------------------------------------------------------------------------ *
 * Symmetric memory. pdcc_sym_alloc is COLLECTIVE over
 * the communicator: every rank calls it with the same size in the same order
 * and gets memory that every rank has mapped and that is bound to a multicast
 * object of its own. A tensor placed there (same offset on every rank) takes
 * the zero-copy kernels: all_reduce / reduce / broadcast / all_gather work on
 * it directly with multimem instructions, no staging and no tile flags.
 * Contract: in one call either every rank passes symmetric tensors of the same
 * allocation and offset, or none does. AUTO picks the zero-copy kernels for
 * symmetric tensors above the small-message threshold; PDCC_ALGO_SYM forces
 * them (error if not applicable). Regions are released by pdcc_comm_destroy.
 * No counterpart in the reference; role of ncclMemAlloc + ncclCommRegister.
 * ------------------------------------------------------------------------ */
int pdcc_sym_alloc(pdcc_comm_t* comm, size_t bytes, void** ptr);
int pdcc_sym_query(pdcc_comm_t* comm, const void* ptr, size_t bytes, int* is_symmetric,
                   int* has_multicast);

/* Generic launcher: validates, picks algo/grid, enqueues one kernel. */
int pdcc_collective(pdcc_comm_t* comm, const pdcc_op_t* op, void* stream);

/* Named wrappers with the reference's call shapes. */
int pdcc_all_reduce(pdcc_comm_t* comm, void* buf, uint64_t count, int dtype, int redop,
                    void* stream); /* main.py:23 */
int pdcc_all_reduce_scaled(pdcc_comm_t* comm, const void* in, void* out, uint64_t count,
                           int dtype, int out_dtype, float scale, void* stream); /* README.md:5 */
int pdcc_reduce(pdcc_comm_t* comm, void* buf, uint64_t count, int dtype, int redop, int dst,
                void* stream); /* main.py:14 */
int pdcc_broadcast(pdcc_comm_t* comm, void* buf, uint64_t count, int dtype, int src,
                   void* stream); /* main.py:81 */
int pdcc_scatter(pdcc_comm_t* comm, void* out, void* const* scatter_list, uint64_t count,
                 int dtype, int src, void* stream); /* main.py:37,39 */
int pdcc_gather(pdcc_comm_t* comm, const void* in, void* const* gather_list, uint64_t count,
                int dtype, int dst, void* stream); /* main.py:52,54 */
int pdcc_all_gather(pdcc_comm_t* comm, void* const* out_list, const void* in, uint64_t count,
                    int dtype, void* stream); /* main.py:68 */
int pdcc_barrier(pdcc_comm_t* comm, void* stream);

/* What `algo`/`nblocks` AUTO resolves to for this op on this comm. */
int pdcc_plan(pdcc_comm_t* comm, const pdcc_op_t* op, int* algo, int* nblocks, int* threads,
              uint64_t* tile_bytes, int* ntiles, int* depth, uint32_t* slot_stride);

/* The same planner without a communicator (and without a GPU): what AUTO would resolve to
 * for `op` on a group of `nranks` with `slot_bytes` slots and `rings` usable CTAs, with or
 * without a multicast mapping. Used by the CPU tests of the size -> variant table. */
int pdcc_plan_dry(int nranks, int has_multicast, size_t slot_bytes, int rings, const pdcc_op_t* op,
                  int* algo, int* nblocks, int* threads, uint64_t* tile_bytes, int* ntiles,
                  int* depth, uint32_t* slot_stride);

/* The planner reads its PDCC_* knobs from the environment once and caches them (it runs on
 * every call). Call this after changing one in a live process (benchmarks do). */
void pdcc_reload_env(void);

/* Number of kernels this library has launched in this process (bench.py's
 * gpu_launches claim is read from here, not estimated). */
uint64_t pdcc_launch_count(void);

/* ------------------------------------------------------------------------ *
 * Single-GPU emulation: N virtual ranks on ONE device, each with its own
 * arena/pads, run as ONE cooperative kernel launch (grid.y = rank) so the
 * cross-rank flag protocol is exercised for real without N GPUs. Used by the
 * `-m gpu` tests on a 1-GPU box. P2P algorithms only (no multicast object).
 * ------------------------------------------------------------------------ */
typedef struct pdcc_emu pdcc_emu_t;
int pdcc_emu_create(int device, int nranks, size_t slot_bytes, int max_blocks,
                    int device_timeout_ms, pdcc_emu_t** out);
int pdcc_emu_destroy(pdcc_emu_t* emu);
/* ops[r] is rank r's view of the same collective (same coll/count/dtype/...). */
int pdcc_emu_collective(pdcc_emu_t* emu, const pdcc_op_t* ops, void* stream);
int pdcc_emu_check(pdcc_emu_t* emu);

/* ------------------------------------------------------------------------ *
 * Allocation-free variant of the two-rank emulated all_reduce. It has no
 * counterpart in the reference. For callers that cannot allocate device memory
 * when they need the collective (a plugin running after its host application
 * took all device memory into its own pool; a loop that cannot afford
 * cudaMalloc's device-wide synchronisation): the two symmetric blocks live in a
 * static __device__ array of about 44 MiB that every GPU loading the library
 * holds until exit. In-place SUM/PRODUCT/MIN/MAX over buf0 and buf1 (count
 * elements each, both on `device`); afterwards both hold the result.
 *   - count * sizeof(dtype) must not exceed PDCC_NOALLOC_MAX_BYTES; dtypes
 *     without a reduction kernel and tensors that are not on `device` are
 *     refused before anything is enqueued.
 *   - The first call on a device resolves the array's device address and zeroes
 *     its flags; it synchronises, so make it before any stream capture.
 *   - All calls on a device share the array: run them on one stream; calls
 *     that overlap on the GPU race.
 *   - Calls no cudaMalloc / cudaFree / allocating routine, reads no host memory
 *     from the stream, and leaves the caller's current device as it was.
 * pdcc_emu2_noalloc_check(device) waits for the device and returns PDCC_EPEER
 * if a device-side wait timed out in any no-alloc call since the last check
 * (the record is kept until it is read here); it allocates nothing either.
 * ------------------------------------------------------------------------ */
#define PDCC_NOALLOC_MAX_BYTES (64ull << 20)
int pdcc_emu2_all_reduce_noalloc(int device, void* buf0, void* buf1, uint64_t count, int dtype,
                                 int redop, void* stream);
int pdcc_emu2_noalloc_check(int device);

/* ------------------------------------------------------------------------ *
 * K9  Local helpers used by the world-size-1 degenerate paths and by the
 * DP demo: strided-free vector copy and fused scale+cast.
 * ------------------------------------------------------------------------ */
int pdcc_local_copy(void* dst, const void* src, uint64_t bytes, int device, void* stream);
int pdcc_local_scale_cast(void* dst, const void* src, uint64_t count, int dtype, int out_dtype,
                          float scale, int device, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PDCC_H_ */
