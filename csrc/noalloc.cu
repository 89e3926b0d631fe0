// Allocation-free entry point: pdcc_emu2_all_reduce_noalloc().
//
// Everything else in this library sets up its working memory with cudaMalloc / cuMemCreate
// when a communicator or an emulator is created. Some hosts cannot do that at the point
// where they need a collective: a plugin that runs inside an inference server after the
// server has handed every free byte of the device to its own pool (cudaMalloc then fails
// or fragments the pool), or a control loop that must not take cudaMalloc's implicit
// device-wide synchronisation. For them this file offers the two-rank emulated all_reduce
// (the same tile-pipeline kernels and the same flag protocol as pdcc_emu_collective with
// nranks = 2) with its two symmetric blocks living in a file-scope `static __device__`
// array instead of in allocated memory. The caller still owns the two tensors.
//
// Cost: the array is kNoallocScratchBytes (about 44 MiB) of device memory on every GPU
// that loads this module, from load until the process exits.
//
// Rules of use
//   * The first call on a device does the set-up for that device: it makes the device
//     current, asks the runtime where the array lives there (a host-side `&array` is only
//     a stand-in, not a device address), and zeroes the flag and counter regions. That
//     first call synchronises, so make it before any stream capture begins.
//   * All calls on a device share the one array. Its ring counters and flags deliberately
//     survive from call to call (they only ever grow, that is how slots are recycled), and
//     so does the device-side error record: a wait that timed out inside a kernel stays
//     recorded until pdcc_emu2_noalloc_check() reads and clears it.
//   * Because the array is shared, two calls that overlap on the GPU (different streams,
//     or a replayed launch racing a direct call) corrupt each other. Issue them on one
//     stream.
//   * Messages above PDCC_NOALLOC_MAX_BYTES are refused; so are dtypes without a reduction
//     kernel and tensors that do not live on `device` — all before anything is enqueued.
//   * The per-rank kernel arguments reach the device as kernel parameters of a small
//     set-up kernel, never through a copy from host memory, so nothing enqueued here keeps
//     pointing at the caller's (or this function's) stack.
//   * The caller's current device is left as it was found.
// Nothing here calls cudaMalloc, cudaFree or any routine that allocates.
#include <cuda_runtime.h>
#include <string.h>

#include <mutex>

#include "common.h"

using namespace pdcc;

namespace {

constexpr int kRanks = 2;
constexpr int kRings = 4;                         // CTAs per emulated rank
constexpr size_t kSlotBytes = 256 * 1024;         // a ring is PDCC_RING_DEPTH x this
constexpr size_t kBlockBytes = kDataOffset + size_t(kRings) * PDCC_RING_DEPTH * kSlotBytes;
constexpr size_t kArgsBytes = 4096;               // the per-rank kernel arguments
constexpr size_t kNoallocScratchBytes = kRanks * kBlockBytes + kArgsBytes;
static_assert(sizeof(RankArgs) * kRanks <= kArgsBytes, "argument area too small");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocScratchBytes];

constexpr int kMaxDevices = 64;
unsigned char* g_scratch_ptr[kMaxDevices] = {};  // device address of the array, per device
std::mutex g_setup_mu;

// Restores the caller's current device on every exit path.
struct DeviceGuard {
  int prev = -1;
  bool armed = false;
  int enter(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
      set_error("cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
      return PDCC_ECUDA;
    }
    armed = prev >= 0 && prev != device;
    return PDCC_OK;
  }
  ~DeviceGuard() {
    if (armed) cudaSetDevice(prev);
  }
};

// Once per device, with that device current: resolve the array's device address and put
// the regions that must start at zero (error record, ring counters, flags, LL packets)
// into that state.
int setup_device(int device, unsigned char** out) {
  std::lock_guard<std::mutex> g(g_setup_mu);
  if (g_scratch_ptr[device] != nullptr) {
    *out = g_scratch_ptr[device];
    return PDCC_OK;
  }
  unsigned char* ptr = nullptr;
  cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scratch);
  if (e != cudaSuccess || ptr == nullptr) {
    set_error("cudaGetSymbolAddress(no-alloc scratch) on device %d: %s", device,
              cudaGetErrorString(e));
    return PDCC_ECUDA;
  }
  for (int r = 0; r < kRanks; ++r) {
    e = cudaMemset(ptr + size_t(r) * kBlockBytes, 0, kZeroBytes);
    if (e != cudaSuccess) {
      set_error("zeroing the no-alloc scratch on device %d: %s", device, cudaGetErrorString(e));
      return PDCC_ECUDA;
    }
  }
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("no-alloc set-up on device %d: %s", device, cudaGetErrorString(e));
    return PDCC_ECUDA;
  }
  g_scratch_ptr[device] = ptr;
  *out = ptr;
  return PDCC_OK;
}

// The tensor must be device memory of `device` (or managed memory, which every device
// reaches): a host pointer or another GPU's memory would fault inside the kernel.
int check_on_device(const void* p, int device, const char* what) {
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    set_error("no-alloc all_reduce: cannot query %s: %s", what, cudaGetErrorString(e));
    return PDCC_EINVAL;
  }
  if (at.type == cudaMemoryTypeManaged) return PDCC_OK;
  if (at.type != cudaMemoryTypeDevice || at.device != device) {
    set_error("no-alloc all_reduce: %s is not device memory of device %d", what, device);
    return PDCC_EINVAL;
  }
  return PDCC_OK;
}

}  // namespace

extern "C" int pdcc_emu2_all_reduce_noalloc(int device, void* buf0, void* buf1, uint64_t count,
                                            int dtype, int redop, void* stream) {
  if (device < 0 || device >= kMaxDevices) {
    set_error("no-alloc all_reduce: device %d out of range", device);
    return PDCC_EINVAL;
  }
  if (buf0 == nullptr || buf1 == nullptr) {
    set_error("no-alloc all_reduce: both tensors must be given");
    return PDCC_EINVAL;
  }
  const size_t esz = dtype_size(dtype);
  if (esz == 0 || !reduce_supported(dtype, redop)) {
    set_error("no-alloc all_reduce: no reduction kernel for dtype %d with op %d", dtype, redop);
    return PDCC_EUNSUPPORTED;
  }
  if (count == 0) return PDCC_OK;
  if (count > PDCC_NOALLOC_MAX_BYTES / esz) {
    set_error("no-alloc all_reduce: %llu bytes requested, this entry point is limited to %llu; "
              "use pdcc_emu_collective (which allocates) for larger messages",
              static_cast<unsigned long long>(count * esz),
              static_cast<unsigned long long>(PDCC_NOALLOC_MAX_BYTES));
    return PDCC_EINVAL;
  }
  DeviceGuard guard;  // the address below is only valid with `device` current
  int rc = guard.enter(device);
  if (rc != PDCC_OK) return rc;
  if ((rc = check_on_device(buf0, device, "buf0")) != PDCC_OK) return rc;
  if ((rc = check_on_device(buf1, device, "buf1")) != PDCC_OK) return rc;
  unsigned char* scratch = nullptr;
  rc = setup_device(device, &scratch);
  if (rc != PDCC_OK) return rc;

  pdcc_op_t op;
  memset(&op, 0, sizeof(op));
  op.coll = PDCC_ALL_REDUCE;
  op.dtype = dtype;
  op.out_dtype = dtype;
  op.redop = redop;
  op.scale = 1.0f;
  op.count = count;

  void* bufs[kRanks] = {buf0, buf1};
  RankArgs host[kRanks];
  memset(host, 0, sizeof(host));
  for (int r = 0; r < kRanks; ++r) {
    DevComm& d = host[r].comm;
    d.rank = r;
    d.nranks = kRanks;
    for (int p = 0; p < kRanks; ++p)
      d.base[p] = reinterpret_cast<char*>(scratch) + size_t(p) * kBlockBytes;
    d.mc = nullptr;
    d.ring_bytes = uint32_t(kSlotBytes * PDCC_RING_DEPTH);
    d.timeout_ms = 5000;
    host[r].in = bufs[r];
    host[r].out = bufs[r];
  }
  Plan plan;
  rc = plan_collective(host[0].comm, kRings, false, op, &plan);
  if (rc != PDCC_OK) return rc;

  RankArgs* dev_args = reinterpret_cast<RankArgs*>(scratch + size_t(kRanks) * kBlockBytes);
  rc = upload_rank_args(host, kRanks, dev_args, stream);  // kernel parameters, not a memcpy
  if (rc != PDCC_OK) return rc;
  return launch_collective(host, kRanks, dev_args, op, plan, stream);
}

// Waits for the device and reports (then clears) a device-side wait that ran out of time in
// any no-alloc call since the last check. Allocates nothing either.
extern "C" int pdcc_emu2_noalloc_check(int device) {
  if (device < 0 || device >= kMaxDevices) {
    set_error("no-alloc check: device %d out of range", device);
    return PDCC_EINVAL;
  }
  DeviceGuard guard;
  int rc = guard.enter(device);
  if (rc != PDCC_OK) return rc;
  unsigned char* scratch = nullptr;
  {
    std::lock_guard<std::mutex> g(g_setup_mu);
    scratch = g_scratch_ptr[device];
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("no-alloc check on device %d: %s", device, cudaGetErrorString(e));
    return PDCC_ECUDA;
  }
  if (scratch == nullptr) return PDCC_OK;  // never used on this device
  for (int r = 0; r < kRanks; ++r) {
    uint32_t err[2] = {0, 0};
    unsigned char* rec = scratch + size_t(r) * kBlockBytes;
    e = cudaMemcpy(err, rec, sizeof(err), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
      set_error("no-alloc check: reading the error record: %s", cudaGetErrorString(e));
      return PDCC_ECUDA;
    }
    if (err[0] != 0) {
      const uint32_t info = err[1];
      uint32_t zero[2] = {0, 0};
      cudaMemcpy(rec, zero, sizeof(zero), cudaMemcpyHostToDevice);
      set_error("no-alloc all_reduce: a device-side wait timed out on emulated rank %d "
                "(channel %u, waiting on rank %u, flag value 0x%06x); the output of that call is "
                "not valid", r, info >> 28, (info >> 24) & 0xfu, info & 0xffffffu);
      return PDCC_EPEER;
    }
  }
  return PDCC_OK;
}
