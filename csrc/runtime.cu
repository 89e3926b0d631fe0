// R2/R3/R4 — communicator runtime.
//
// Per group and per rank this file owns ONE physical "symmetric block" (ctrl words,
// signal pads, tile rings, optional symmetric heap; layout in common.h), makes every
// peer's block a plain device pointer in this process, and — where the box allows —
// binds all blocks to one NVSwitch multicast object so that a single address reaches
// every GPU (multimem.*). Handle exchange, in order of preference:
//
//   POSIX fd   cuMemCreate + cuMemExportToShareableHandle, fd passed between the rank
//              processes with SCM_RIGHTS over abstract unix sockets whose names are
//              published in the bootstrap store. Required for multicast.
//   legacy IPC cudaMalloc + cudaIpcGetMemHandle through the store. P2P only.
//   local      single process (world size 1, or the N-virtual-rank emulation).
//
// The CUDA driver API is reached through cudaGetDriverEntryPoint so the library has no
// link-time dependency on libcuda.so.1 and still loads on a CPU-only box (where the
// store and the argument checks remain usable and everything CUDA returns PDCC_ECUDA).
//
// Role parity: ProcessGroupNCCL construction / ncclCommInitRank / NVLS setup, reached
// from the reference at main.py:94 and at every new_group (main.py:11,21,31,46,63,75).
#include <cuda.h>
#include <cuda_runtime.h>
#include <errno.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <chrono>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "common.h"

#define PDCC_MIN_CTAS_PER_SM_HOST 1  // keep in sync with kMinCtasPerSm (kernels.cuh)

namespace pdcc {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

#define CUDA_TRY(expr)                                                              \
  do {                                                                              \
    cudaError_t e__ = (expr);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__,  \
                __LINE__);                                                          \
      return PDCC_ECUDA;                                                            \
    }                                                                               \
  } while (0)

// ---------------------------------------------------------------- driver API table ---
struct Driver {
  bool ok = false;
#define PDCC_DRV(name) decltype(&name) p_##name = nullptr;
  PDCC_DRV(cuGetErrorString)
  PDCC_DRV(cuDeviceGet)
  PDCC_DRV(cuDeviceGetAttribute)
  PDCC_DRV(cuMemGetAllocationGranularity)
  PDCC_DRV(cuMemCreate)
  PDCC_DRV(cuMemRelease)
  PDCC_DRV(cuMemExportToShareableHandle)
  PDCC_DRV(cuMemImportFromShareableHandle)
  PDCC_DRV(cuMemAddressReserve)
  PDCC_DRV(cuMemAddressFree)
  PDCC_DRV(cuMemMap)
  PDCC_DRV(cuMemUnmap)
  PDCC_DRV(cuMemSetAccess)
  PDCC_DRV(cuMulticastCreate)
  PDCC_DRV(cuMulticastAddDevice)
  PDCC_DRV(cuMulticastBindMem)
  PDCC_DRV(cuMulticastGetGranularity)
#undef PDCC_DRV
};

static Driver g_drv;
static std::once_flag g_drv_once;

static void load_driver() {
  Driver d;
  bool all = true;
#define PDCC_LOAD(name)                                                                     \
  do {                                                                                      \
    void* fn = nullptr;                                                                     \
    cudaDriverEntryPointQueryResult q;                                                      \
    if (cudaGetDriverEntryPoint(#name, &fn, cudaEnableDefault, &q) != cudaSuccess || !fn)   \
      all = false;                                                                          \
    d.p_##name = reinterpret_cast<decltype(&name)>(fn);                                     \
  } while (0)
  PDCC_LOAD(cuGetErrorString);
  PDCC_LOAD(cuDeviceGet);
  PDCC_LOAD(cuDeviceGetAttribute);
  PDCC_LOAD(cuMemGetAllocationGranularity);
  PDCC_LOAD(cuMemCreate);
  PDCC_LOAD(cuMemRelease);
  PDCC_LOAD(cuMemExportToShareableHandle);
  PDCC_LOAD(cuMemImportFromShareableHandle);
  PDCC_LOAD(cuMemAddressReserve);
  PDCC_LOAD(cuMemAddressFree);
  PDCC_LOAD(cuMemMap);
  PDCC_LOAD(cuMemUnmap);
  PDCC_LOAD(cuMemSetAccess);
  PDCC_LOAD(cuMulticastCreate);
  PDCC_LOAD(cuMulticastAddDevice);
  PDCC_LOAD(cuMulticastBindMem);
  PDCC_LOAD(cuMulticastGetGranularity);
#undef PDCC_LOAD
  (void)cudaGetLastError();
  d.ok = all;
  g_drv = d;
}

static const Driver& drv() {
  std::call_once(g_drv_once, load_driver);
  return g_drv;
}

static const char* cu_err(CUresult r) {
  const char* s = nullptr;
  if (drv().p_cuGetErrorString && drv().p_cuGetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown CUresult";
}

#define CU_TRY(call)                                                                 \
  do {                                                                               \
    CUresult r__ = drv().p_##call;                                                   \
    if (r__ != CUDA_SUCCESS) {                                                       \
      set_error("%s failed: %s (%s:%d)", #call, cu_err(r__), __FILE__, __LINE__);    \
      return PDCC_ECUDA;                                                             \
    }                                                                                \
  } while (0)

static size_t round_up_sz(size_t v, size_t m) { return (v + m - 1) / m * m; }

// ---------------------------------------------------------------- fd passing ---------
// One message = {kind, src_rank} + one fd in the ancillary data.
struct FdMsg {
  int kind;  // 0: memory handle, 1: multicast handle
  int src;
};

static int make_listener(const std::string& name) {
  int fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
  if (fd < 0) return -1;
  sockaddr_un addr{};
  addr.sun_family = AF_UNIX;
  // abstract namespace: leading NUL, no filesystem entry to clean up
  memcpy(addr.sun_path + 1, name.data(), std::min(name.size(), sizeof(addr.sun_path) - 2));
  socklen_t len = socklen_t(offsetof(sockaddr_un, sun_path) + 1 + name.size());
  if (::bind(fd, reinterpret_cast<sockaddr*>(&addr), len) != 0 || ::listen(fd, 64) != 0) {
    ::close(fd);
    return -1;
  }
  return fd;
}

static int connect_to(const std::string& name, int timeout_ms) {
  auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  for (;;) {
    int fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
    if (fd < 0) return -1;
    sockaddr_un addr{};
    addr.sun_family = AF_UNIX;
    memcpy(addr.sun_path + 1, name.data(), std::min(name.size(), sizeof(addr.sun_path) - 2));
    socklen_t len = socklen_t(offsetof(sockaddr_un, sun_path) + 1 + name.size());
    if (::connect(fd, reinterpret_cast<sockaddr*>(&addr), len) == 0) return fd;
    ::close(fd);
    if (std::chrono::steady_clock::now() > deadline) return -1;
    std::this_thread::sleep_for(std::chrono::milliseconds(5));
  }
}

static bool send_fd(int sock, const FdMsg& m, int fd) {
  msghdr msg{};
  iovec io{const_cast<FdMsg*>(&m), sizeof(m)};
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  char ctl[CMSG_SPACE(sizeof(int))];
  memset(ctl, 0, sizeof(ctl));
  msg.msg_control = ctl;
  msg.msg_controllen = sizeof(ctl);
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  return ::sendmsg(sock, &msg, MSG_NOSIGNAL) == ssize_t(sizeof(m));
}

static bool recv_fd(int sock, FdMsg* m, int* fd) {
  msghdr msg{};
  iovec io{m, sizeof(*m)};
  msg.msg_iov = &io;
  msg.msg_iovlen = 1;
  char ctl[CMSG_SPACE(sizeof(int))];
  memset(ctl, 0, sizeof(ctl));
  msg.msg_control = ctl;
  msg.msg_controllen = sizeof(ctl);
  ssize_t k;
  do {
    k = ::recvmsg(sock, &msg, 0);
  } while (k < 0 && errno == EINTR);
  if (k != ssize_t(sizeof(*m))) return false;
  cmsghdr* c = CMSG_FIRSTHDR(&msg);
  if (!c || c->cmsg_level != SOL_SOCKET || c->cmsg_type != SCM_RIGHTS) return false;
  memcpy(fd, CMSG_DATA(c), sizeof(int));
  return true;
}

// Accept `count` messages on `lsock`; out[src] = fd for messages of `kind`.
static bool accept_fds(int lsock, int count, int kind, std::vector<int>* out, int timeout_ms) {
  auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  int got = 0;
  while (got < count) {
    timeval tv{0, 200000};
    fd_set rs;
    FD_ZERO(&rs);
    FD_SET(lsock, &rs);
    int r = ::select(lsock + 1, &rs, nullptr, nullptr, &tv);
    if (r <= 0) {
      if (std::chrono::steady_clock::now() > deadline) return false;
      continue;
    }
    int s = ::accept(lsock, nullptr, nullptr);
    if (s < 0) continue;
    FdMsg m;
    int fd = -1;
    bool ok = recv_fd(s, &m, &fd);
    ::close(s);
    if (!ok) return false;
    if (m.kind != kind || m.src < 0 || m.src >= int(out->size())) {
      if (fd >= 0) ::close(fd);
      return false;
    }
    (*out)[m.src] = fd;
    ++got;
  }
  return true;
}

}  // namespace pdcc

using namespace pdcc;

// ---------------------------------------------------------------- comm object --------
struct Mapping {
  CUmemGenericAllocationHandle handle = 0;
  CUdeviceptr va = 0;
  size_t size = 0;
  bool mapped = false;
  bool have_handle = false;
};

// One symmetric allocation: the same number of bytes on every rank, every rank's copy
// mapped into this process, and (where the box allows) one multicast alias over all of
// them. The communicator's own block is region 0; pdcc_sym_alloc adds more.
struct Region {
  size_t bytes = 0;
  Mapping peers[PDCC_MAX_RANKS];  // POSIX-fd path
  Mapping mc;
  bool has_mc = false;
  void* legacy_base = nullptr;    // cudaMalloc'd (legacy IPC / local)
  void* legacy_peer[PDCC_MAX_RANKS] = {};
  char* base[PDCC_MAX_RANKS] = {};  // base[p] = local VA of rank p's copy
  char* mcp = nullptr;              // multicast alias (nullptr: none)
};

constexpr int kMaxRegions = 64;

struct pdcc_comm {
  pdcc_comm_config_t cfg{};
  pdcc_store_t* store = nullptr;   // kept for later symmetric allocations (caller owns it)
  bool want_mc = false;            // every rank reported multicast support
  Region* regions[kMaxRegions] = {};
  int nregions = 0;                // region counter doubles as the store-key suffix
  int kind = PDCC_HANDLE_LOCAL;
  int sm_count = 0;
  int max_blocks = PDCC_MAX_BLOCKS;
  size_t slot_bytes = 0;   // configured quarter-ring (ring_bytes = PDCC_RING_DEPTH x this)
  size_t ring_bytes = 0;
  size_t block_bytes = 0;
  size_t heap_off = 0;
  bool has_mc = false;
  Region blk;  // the communicator's own block (ctrl, pads, LL slots, rings, heap)
  DevComm dev{};
};

static void unmap(Mapping& m) {
  const Driver& d = drv();
  if (m.mapped) {
    d.p_cuMemUnmap(m.va, m.size);
    m.mapped = false;
  }
  if (m.va) {
    d.p_cuMemAddressFree(m.va, m.size);
    m.va = 0;
  }
  if (m.have_handle) {
    d.p_cuMemRelease(m.handle);
    m.have_handle = false;
  }
}

static int map_handle(Mapping& m, size_t size, int device) {
  m.size = size;
  CU_TRY(cuMemAddressReserve(&m.va, size, 0, 0, 0));
  CU_TRY(cuMemMap(m.va, size, 0, m.handle, 0));
  m.mapped = true;
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CU_TRY(cuMemSetAccess(m.va, size, &acc, 1));
  return PDCC_OK;
}

static std::string skey(const pdcc_comm_config_t& cfg, const char* what, int idx = -1) {
  char buf[192];
  if (idx >= 0) snprintf(buf, sizeof(buf), "%s/%s/%d", cfg.key, what, idx);
  else snprintf(buf, sizeof(buf), "%s/%s", cfg.key, what);
  return buf;
}

static int store_put_int(pdcc_store_t* st, const std::string& key, int v) {
  return pdcc_store_set(st, key.c_str(), &v, sizeof(v));
}
static int store_get_int(pdcc_store_t* st, const std::string& key, int* v, int timeout_ms) {
  int len = 0;
  int rc = pdcc_store_get(st, key.c_str(), v, sizeof(*v), &len, timeout_ms);
  if (rc == PDCC_OK && len != int(sizeof(*v))) {
    set_error("store key %s has unexpected length %d", key.c_str(), len);
    return PDCC_ESYS;
  }
  return rc;
}

extern "C" {

const char* pdcc_last_error(void) { return g_err; }
const char* pdcc_version(void) { return "pdcc 0.1 (sm_100a)"; }

int pdcc_probe(int device, pdcc_caps_t* caps) {
  if (!caps) return PDCC_EINVAL;
  memset(caps, 0, sizeof(*caps));
  caps->device = device;
  int count = 0;
  CUDA_TRY(cudaGetDeviceCount(&count));
  caps->device_count = count;
  if (device < 0 || device >= count) {
    set_error("device %d out of range (%d devices)", device, count);
    return PDCC_EINVAL;
  }
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  snprintf(caps->name, sizeof(caps->name), "%s", prop.name);
  caps->sm_count = prop.multiProcessorCount;
  caps->cc_major = prop.major;
  caps->cc_minor = prop.minor;
  caps->total_mem = prop.totalGlobalMem;
  caps->cooperative_launch = prop.cooperativeLaunch;
  CUDA_TRY(cudaSetDevice(device));
  CUDA_TRY(cudaFree(nullptr));
  const Driver& d = drv();
  if (!d.ok) return PDCC_OK;  // runtime-only box: everything else stays 0
  CUdevice cudev;
  CU_TRY(cuDeviceGet(&cudev, device));
  int v = 0;
  if (d.p_cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED,
                               cudev) == CUDA_SUCCESS)
    caps->vmm_supported = v;
  if (d.p_cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED,
                               cudev) == CUDA_SUCCESS)
    caps->posix_fd_supported = v;
  if (d.p_cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, cudev) ==
      CUDA_SUCCESS)
    caps->fabric_supported = v;
  if (d.p_cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev) == CUDA_SUCCESS)
    caps->multicast_supported = v;
  CUmemAllocationProp ap{};
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t g = 0;
  if (caps->vmm_supported &&
      d.p_cuMemGetAllocationGranularity(&g, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) ==
          CUDA_SUCCESS)
    caps->alloc_granularity = g;
  if (caps->multicast_supported && count >= 2) {
    CUmulticastObjectProp mp{};
    mp.numDevices = 2;
    mp.size = 2 << 20;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    if (d.p_cuMulticastGetGranularity(&g, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) ==
        CUDA_SUCCESS)
      caps->mc_granularity = g;
  }
  return PDCC_OK;
}

static int fill_dev(pdcc_comm* c) {
  c->dev.rank = c->cfg.rank;
  c->dev.nranks = c->cfg.nranks;
  c->dev.ring_bytes = uint32_t(c->ring_bytes);
  c->dev.timeout_ms = uint32_t(c->cfg.device_timeout_ms > 0 ? c->cfg.device_timeout_ms : 10000);
  for (int p = 0; p < PDCC_MAX_RANKS; ++p) c->dev.base[p] = nullptr;
  for (int p = 0; p < c->cfg.nranks; ++p) c->dev.base[p] = c->blk.base[p];
  c->dev.mc = c->has_mc ? c->blk.mcp : nullptr;
  return PDCC_OK;
}

static void sizes_from_cfg(pdcc_comm* c) {
  // One CTA's ring is PDCC_RING_DEPTH x slot_bytes; every call cuts it into slots of its own
  // size (plan_collective). 2 MiB per ring by default: 148 rings + header = 314 MiB per
  // communicator (1.2 GB in round 1). Measured at 4 and 8 GPUs (profiles/r2): 2 MiB rings are
  // 3-10 % faster than 1 MiB ones on reduce / broadcast (eight 64 KiB tiles in flight also
  // for the layouts that double or n-fold a slot), 4 MiB rings add nothing.
  c->slot_bytes = c->cfg.slot_bytes ? round_up_sz(c->cfg.slot_bytes, 4096) : (size_t(512) << 10);
  c->ring_bytes = c->slot_bytes * PDCC_RING_DEPTH;
  c->max_blocks = c->cfg.max_blocks > 0 ? c->cfg.max_blocks : PDCC_MAX_BLOCKS;
  if (c->max_blocks > PDCC_MAX_BLOCKS) c->max_blocks = PDCC_MAX_BLOCKS;
  c->heap_off = round_up_sz(kDataOffset + size_t(c->max_blocks) * c->ring_bytes, size_t(2) << 20);
  c->block_bytes = c->heap_off + c->cfg.heap_bytes;
}

static void region_release(pdcc_comm* c, Region* rg) {
  if (c->kind == PDCC_HANDLE_POSIX_FD) {
    unmap(rg->mc);
    for (int p = 0; p < c->cfg.nranks; ++p) unmap(rg->peers[p]);
  } else {
    for (int p = 0; p < c->cfg.nranks; ++p)
      if (p != c->cfg.rank && rg->legacy_peer[p]) cudaIpcCloseMemHandle(rg->legacy_peer[p]);
    if (rg->legacy_base) cudaFree(rg->legacy_base);
    rg->legacy_base = nullptr;
  }
}

int pdcc_comm_destroy(pdcc_comm_t* c) {
  if (!c) return PDCC_EINVAL;
  cudaSetDevice(c->cfg.device);
  cudaDeviceSynchronize();
  for (int i = 0; i < c->nregions; ++i) {
    if (c->regions[i]) {
      region_release(c, c->regions[i]);
      delete c->regions[i];
    }
  }
  region_release(c, &c->blk);
  delete c;
  return PDCC_OK;
}

// Closes a set of file descriptors on every exit path.
struct FdBag {
  std::vector<int> fds;
  int add(int fd) {
    if (fd >= 0) fds.push_back(fd);
    return fd;
  }
  void close_one(int fd) {
    for (auto& f : fds)
      if (f == fd && fd >= 0) {
        ::close(f);
        f = -1;
      }
  }
  ~FdBag() {
    for (int f : fds)
      if (f >= 0) ::close(f);
  }
};

// Builds region `rg` (rg->bytes requested, rounded up to the allocation / multicast
// granularity): my copy, every peer's copy mapped here, and the multicast alias when
// want_mc and every rank manages to bind. `tag` namespaces the store keys and the socket
// name, so any number of regions can be built one after another on one communicator.
static int create_posix(pdcc_comm* c, pdcc_store_t* st, bool want_mc, Region* rg, const char* tag) {
  const Driver& d = drv();
  const pdcc_comm_config_t& cfg = c->cfg;
  const int n = cfg.nranks, me = cfg.rank, tmo = cfg.timeout_ms;

  CUmemAllocationProp ap{};
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = cfg.device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  CU_TRY(cuMemGetAllocationGranularity(&gran, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  CUmulticastObjectProp mp{};
  if (want_mc) {
    mp.numDevices = unsigned(n);
    mp.size = rg->bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d.p_cuMulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) ==
            CUDA_SUCCESS && mg > gran)
      gran = mg;
  }
  rg->bytes = round_up_sz(rg->bytes, gran);
  mp.size = rg->bytes;

  // my block
  Mapping& mine = rg->peers[me];
  CU_TRY(cuMemCreate(&mine.handle, rg->bytes, &ap, 0));
  mine.have_handle = true;
  FdBag bag;  // every early return below closes what is still open
  int my_fd = -1;
  CU_TRY(cuMemExportToShareableHandle(&my_fd, mine.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR,
                                      0));
  bag.add(my_fd);

  // listener, published in the store
  char sname[160];
  snprintf(sname, sizeof(sname), "pdcc-%s-%s%d-%d", cfg.key, tag, me, int(getpid()));
  int lsock = make_listener(sname);
  if (lsock < 0) {
    set_error("cannot bind unix socket %s: %s", sname, strerror(errno));
    return PDCC_ESYS;
  }
  bag.add(lsock);
  int rc = pdcc_store_set(st, skey(cfg, (std::string(tag) + "sock").c_str(), me).c_str(), sname, int(strlen(sname)));
  if (rc) return rc;

  // receive in a helper thread while this thread sends
  std::vector<int> fds(n, -1);
  bool recv_ok = false;
  std::thread rx([&] { recv_ok = accept_fds(lsock, n - 1, 0, &fds, tmo); });
  bool send_ok = true;
  std::string send_err;
  for (int i = 1; i < n && send_ok; ++i) {
    const int p = (me + i) % n;
    char pname[160] = {0};
    int len = 0;
    rc = pdcc_store_get(st, skey(cfg, (std::string(tag) + "sock").c_str(), p).c_str(), pname, sizeof(pname) - 1, &len, tmo);
    if (rc) {
      send_ok = false;
      break;
    }
    int s = connect_to(std::string(pname, size_t(len)), tmo);
    if (s < 0 || !send_fd(s, FdMsg{0, me}, my_fd)) {
      send_err = std::string("sending fd to rank ") + std::to_string(p) + ": " + strerror(errno);
      send_ok = false;
    }
    if (s >= 0) ::close(s);
  }
  rx.join();
  bag.close_one(my_fd);
  for (int p = 0; p < n; ++p) bag.add(fds[p]);
  if (!send_ok || !recv_ok) {
    set_error("fd exchange failed (%s%s)", send_ok ? "" : send_err.c_str(),
              recv_ok ? "" : " receive timed out");
    return PDCC_ESYS;
  }

  // map every block (mine included) into this process
  rc = map_handle(mine, rg->bytes, cfg.device);
  if (rc) return rc;
  for (int p = 0; p < n; ++p) {
    if (p == me) continue;
    Mapping& m = rg->peers[p];
    CUresult r = d.p_cuMemImportFromShareableHandle(
        &m.handle, reinterpret_cast<void*>(static_cast<uintptr_t>(fds[p])),
        CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
    bag.close_one(fds[p]);
    if (r != CUDA_SUCCESS) {
      set_error("cuMemImportFromShareableHandle(rank %d): %s", p, cu_err(r));
      return PDCC_ECUDA;
    }
    m.have_handle = true;
    rc = map_handle(m, rg->bytes, cfg.device);
    if (rc) return rc;
  }

  // Nobody may still be collecting memory handles when rank 0 starts handing out the
  // multicast handle on the same sockets.
  rc = pdcc_store_barrier(st, skey(cfg, (std::string(tag) + "memx").c_str()).c_str(), n, tmo);
  if (rc) return rc;

  // ---- multicast: every step is voted on so that all ranks agree on has_mc ----------
  rg->has_mc = false;
  if (want_mc) {
    int ok = 1;
    CUdevice cudev = 0;
    if (d.p_cuDeviceGet(&cudev, cfg.device) != CUDA_SUCCESS) ok = 0;
    if (me == 0) {
      int mfd = -1;
      if (ok && d.p_cuMulticastCreate(&rg->mc.handle, &mp) != CUDA_SUCCESS) ok = 0;
      if (ok) rg->mc.have_handle = true;
      if (ok && d.p_cuMemExportToShareableHandle(&mfd, rg->mc.handle,
                                                 CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR,
                                                 0) != CUDA_SUCCESS)
        ok = 0;
      store_put_int(st, skey(cfg, (std::string(tag) + "mc_created").c_str()), ok);
      if (ok) {
        for (int p = 1; p < n; ++p) {
          char pname[160] = {0};
          int len = 0;
          if (pdcc_store_get(st, skey(cfg, (std::string(tag) + "sock").c_str(), p).c_str(), pname, sizeof(pname) - 1, &len,
                             tmo))
            continue;
          int s = connect_to(std::string(pname, size_t(len)), tmo);
          if (s >= 0) {
            send_fd(s, FdMsg{1, 0}, mfd);
            ::close(s);
          }
        }
        ::close(mfd);
      }
    } else {
      int created = 0;
      if (store_get_int(st, skey(cfg, (std::string(tag) + "mc_created").c_str()), &created, tmo) != PDCC_OK) created = 0;
      if (!created) {
        ok = 0;
      } else {
        std::vector<int> mfds(n, -1);
        if (!accept_fds(lsock, 1, 1, &mfds, tmo)) {
          ok = 0;
        } else {
          if (d.p_cuMemImportFromShareableHandle(
                  &rg->mc.handle, reinterpret_cast<void*>(static_cast<uintptr_t>(mfds[0])),
                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) != CUDA_SUCCESS)
            ok = 0;
          else
            rg->mc.have_handle = true;
          ::close(mfds[0]);
        }
      }
    }
    if (ok && d.p_cuMulticastAddDevice(rg->mc.handle, cudev) != CUDA_SUCCESS) ok = 0;
    // vote 1: everyone added its device (binding before that is an error)
    store_put_int(st, skey(cfg, (std::string(tag) + "mc_add").c_str(), me), ok);
    for (int p = 0; p < n; ++p) {
      int v = 0;
      if (store_get_int(st, skey(cfg, (std::string(tag) + "mc_add").c_str(), p), &v, tmo) != PDCC_OK || !v) ok = 0;
    }
    if (ok && d.p_cuMulticastBindMem(rg->mc.handle, 0, mine.handle, 0, rg->bytes, 0) !=
                  CUDA_SUCCESS)
      ok = 0;
    if (ok) {
      rg->mc.size = rg->bytes;
      if (d.p_cuMemAddressReserve(&rg->mc.va, rg->bytes, gran, 0, 0) != CUDA_SUCCESS) {
        ok = 0;
        rg->mc.va = 0;
      }
    }
    if (ok) {
      if (d.p_cuMemMap(rg->mc.va, rg->bytes, 0, rg->mc.handle, 0) != CUDA_SUCCESS) ok = 0;
      else rg->mc.mapped = true;
    }
    if (ok) {
      CUmemAccessDesc acc{};
      acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
      acc.location.id = cfg.device;
      acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
      if (d.p_cuMemSetAccess(rg->mc.va, rg->bytes, &acc, 1) != CUDA_SUCCESS) ok = 0;
    }
    // vote 2: everyone bound and mapped
    store_put_int(st, skey(cfg, (std::string(tag) + "mc_bound").c_str(), me), ok);
    for (int p = 0; p < n; ++p) {
      int v = 0;
      if (store_get_int(st, skey(cfg, (std::string(tag) + "mc_bound").c_str(), p), &v, tmo) != PDCC_OK || !v) ok = 0;
    }
    rg->has_mc = ok != 0;
    if (!rg->has_mc) unmap(rg->mc);
  }
  for (int p = 0; p < n; ++p) rg->base[p] = reinterpret_cast<char*>(rg->peers[p].va);
  rg->mcp = rg->has_mc ? reinterpret_cast<char*>(rg->mc.va) : nullptr;
  return PDCC_OK;
}

static int create_legacy(pdcc_comm* c, pdcc_store_t* st, Region* rg, const char* tag) {
  const pdcc_comm_config_t& cfg = c->cfg;
  const int n = cfg.nranks, me = cfg.rank, tmo = cfg.timeout_ms;
  CUDA_TRY(cudaMalloc(&rg->legacy_base, rg->bytes));
  rg->legacy_peer[me] = rg->legacy_base;
  rg->base[me] = static_cast<char*>(rg->legacy_base);
  if (n == 1) return PDCC_OK;
  cudaIpcMemHandle_t h;
  CUDA_TRY(cudaIpcGetMemHandle(&h, rg->legacy_base));
  int rc = pdcc_store_set(st, skey(cfg, (std::string(tag) + "ipc").c_str(), me).c_str(), &h, sizeof(h));
  if (rc) return rc;
  for (int p = 0; p < n; ++p) {
    if (p == me) continue;
    cudaIpcMemHandle_t ph;
    int len = 0;
    rc = pdcc_store_get(st, skey(cfg, (std::string(tag) + "ipc").c_str(), p).c_str(), &ph, sizeof(ph), &len, tmo);
    if (rc) return rc;
    CUDA_TRY(cudaIpcOpenMemHandle(&rg->legacy_peer[p], ph, cudaIpcMemLazyEnablePeerAccess));
    rg->base[p] = static_cast<char*>(rg->legacy_peer[p]);
  }
  return PDCC_OK;
}

int pdcc_comm_create(pdcc_store_t* st, const pdcc_comm_config_t* cfg_in, pdcc_comm_t** out) {
  if (!cfg_in || !out) return PDCC_EINVAL;
  pdcc_comm_config_t cfg = *cfg_in;
  if (cfg.nranks < 1 || cfg.nranks > PDCC_MAX_RANKS || cfg.rank < 0 || cfg.rank >= cfg.nranks) {
    set_error("bad rank/nranks %d/%d (max %d ranks)", cfg.rank, cfg.nranks, PDCC_MAX_RANKS);
    return PDCC_EINVAL;
  }
  if (cfg.nranks > 1 && !st) {
    set_error("a store is required for nranks > 1");
    return PDCC_EINVAL;
  }
  if (cfg.timeout_ms <= 0) cfg.timeout_ms = 120000;
  cfg.key[sizeof(cfg.key) - 1] = 0;

  pdcc_caps_t caps;
  int rc = pdcc_probe(cfg.device, &caps);
  if (rc) return rc;

  auto* c = new pdcc_comm();
  c->cfg = cfg;
  c->store = st;
  c->sm_count = caps.sm_count;
  sizes_from_cfg(c);
  c->blk.bytes = c->block_bytes;

  const int n = cfg.nranks, me = cfg.rank;
  if (n == 1) {
    c->kind = PDCC_HANDLE_LOCAL;
    rc = create_legacy(c, st, &c->blk, "");
  } else {
    // agree on the handle kind: everyone publishes what it can do
    int mine[2] = {caps.vmm_supported && caps.posix_fd_supported && drv().ok,
                   caps.multicast_supported};
    if (cfg.handle_kind == PDCC_HANDLE_LEGACY_IPC) mine[0] = 0;
    if (cfg.multicast == 0) mine[1] = 0;
    rc = pdcc_store_set(st, skey(cfg, "caps", me).c_str(), mine, sizeof(mine));
    bool all_posix = true, all_mc = true;
    for (int p = 0; p < n && rc == PDCC_OK; ++p) {
      int theirs[2] = {0, 0};
      int len = 0;
      rc = pdcc_store_get(st, skey(cfg, "caps", p).c_str(), theirs, sizeof(theirs), &len,
                          cfg.timeout_ms);
      all_posix = all_posix && theirs[0];
      all_mc = all_mc && theirs[1];
    }
    if (rc == PDCC_OK) {
      if (all_posix) {
        c->kind = PDCC_HANDLE_POSIX_FD;
        c->want_mc = all_mc;
        rc = create_posix(c, st, all_mc, &c->blk, "");
        c->has_mc = c->blk.has_mc;
        c->block_bytes = c->blk.bytes;
      } else if (cfg.handle_kind == PDCC_HANDLE_POSIX_FD) {
        set_error("POSIX fd handles requested but not supported on every rank");
        rc = PDCC_EUNSUPPORTED;
      } else {
        c->kind = PDCC_HANDLE_LEGACY_IPC;
        rc = create_legacy(c, st, &c->blk, "");
      }
    }
    if (rc == PDCC_OK && cfg.multicast == 1 && !c->has_mc) {
      set_error("multicast required but could not be set up on every rank");
      rc = PDCC_EUNSUPPORTED;
    }
  }
  if (rc != PDCC_OK) {
    std::string keep = g_err;
    pdcc_comm_destroy(c);
    set_error("%s", keep.c_str());
    return rc;
  }
  fill_dev(c);

  // zero ctrl + pads + LL slots of my block, then make sure nobody signals into a block that has
  // not been zeroed yet
  cudaError_t e = cudaMemset(c->dev.base[me], 0, kZeroBytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("zeroing the block header: %s", cudaGetErrorString(e));
    pdcc_comm_destroy(c);
    return PDCC_ECUDA;
  }
  if (n > 1) {
    rc = pdcc_store_barrier(st, skey(cfg, "ready").c_str(), n, cfg.timeout_ms);
    if (rc) {
      pdcc_comm_destroy(c);
      return rc;
    }
  }
  *out = c;
  return PDCC_OK;
}

// CTAs of one call must all be resident (they wait on their peers): one per SM.
static int ring_limit(const pdcc_comm* c) {
  const int resident = c->sm_count > 0 ? c->sm_count * PDCC_MIN_CTAS_PER_SM_HOST : c->max_blocks;
  return c->max_blocks < resident ? c->max_blocks : resident;
}

int pdcc_comm_info(pdcc_comm_t* c, pdcc_comm_info_t* info) {
  if (!c || !info) return PDCC_EINVAL;
  memset(info, 0, sizeof(*info));
  info->rank = c->cfg.rank;
  info->nranks = c->cfg.nranks;
  info->device = c->cfg.device;
  info->handle_kind = c->kind;
  info->has_multicast = c->has_mc;
  info->sm_count = c->sm_count;
  info->max_blocks = c->max_blocks;
  info->slot_bytes = c->slot_bytes;
  info->block_bytes = c->block_bytes;
  info->heap_bytes = c->cfg.heap_bytes;
  info->heap_base = c->cfg.heap_bytes ? c->dev.base[c->cfg.rank] + c->heap_off : nullptr;
  return PDCC_OK;
}

static int check_block(char* base, int rank) {
  uint32_t err[2] = {0, 0};
  CUDA_TRY(cudaMemcpy(err, base, sizeof(err), cudaMemcpyDeviceToHost));
  if (err[0] != 0) {
    const uint32_t info = err[1];
    set_error("rank %d: device-side wait timed out (channel %u, waiting on rank %u, flag value "
              "0x%06x): peers are not running the same collective sequence",
              rank, info >> 28, (info >> 24) & 0xfu, info & 0xffffffu);
    uint32_t zero[2] = {0, 0};
    cudaMemcpy(base, zero, sizeof(zero), cudaMemcpyHostToDevice);
    return PDCC_EPEER;
  }
  return PDCC_OK;
}

// Copies the timeline CTA 0 recorded (PDCC_TRACE=1) to `out` (8 uint32 per event) and
// clears it. Synchronises the device. *count = events recorded (may exceed cap: truncated).
int pdcc_comm_read_trace(pdcc_comm_t* c, uint32_t* out, int cap_events, int* count) {
  if (!c || !out || !count) return PDCC_EINVAL;
  CUDA_TRY(cudaSetDevice(c->cfg.device));
  CUDA_TRY(cudaDeviceSynchronize());
  char* base = c->dev.base[c->cfg.rank];
  uint32_t n = 0;
  CUDA_TRY(cudaMemcpy(&n, base + size_t(kCtrlTrace) * 4, 4, cudaMemcpyDeviceToHost));
  *count = int(n);
  int take = int(n) < kCtrlTraceCap ? int(n) : kCtrlTraceCap;
  if (take > cap_events) take = cap_events;
  if (take > 0)
    CUDA_TRY(cudaMemcpy(out, base + size_t(kCtrlTrace + 4) * 4, size_t(take) * 32, cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemset(base + size_t(kCtrlTrace) * 4, 0, 4));
  return PDCC_OK;
}

int pdcc_comm_check(pdcc_comm_t* c) {
  if (!c) return PDCC_EINVAL;
  CUDA_TRY(cudaSetDevice(c->cfg.device));
  CUDA_TRY(cudaDeviceSynchronize());
  return check_block(c->dev.base[c->cfg.rank], c->cfg.rank);
}

int pdcc_plan(pdcc_comm_t* c, const pdcc_op_t* op, int* algo, int* nblocks, int* threads,
              uint64_t* tile_bytes, int* ntiles, int* depth, uint32_t* slot_stride) {
  if (!c || !op) return PDCC_EINVAL;
  Plan p;
  int rc = plan_collective(c->dev, ring_limit(c), c->has_mc, *op, &p);
  if (rc) return rc;
  if (algo) *algo = p.algo;
  if (nblocks) *nblocks = p.nblocks;
  if (threads) *threads = p.threads;
  if (tile_bytes) *tile_bytes = p.tile_bytes;
  if (ntiles) *ntiles = p.ntiles;
  if (depth) *depth = p.depth;
  if (slot_stride) *slot_stride = p.slot_stride;
  return PDCC_OK;
}

int pdcc_plan_dry(int nranks, int has_multicast, size_t slot_bytes, int rings, const pdcc_op_t* op,
                  int* algo, int* nblocks, int* threads, uint64_t* tile_bytes, int* ntiles,
                  int* depth, uint32_t* slot_stride) {
  if (!op || nranks < 1 || nranks > PDCC_MAX_RANKS) return PDCC_EINVAL;
  DevComm dev;
  memset(&dev, 0, sizeof(dev));
  dev.rank = 0;
  dev.nranks = nranks;
  dev.ring_bytes = uint32_t((slot_bytes ? slot_bytes : (size_t(512) << 10)) * PDCC_RING_DEPTH);
  dev.mc = has_multicast ? reinterpret_cast<char*>(uintptr_t(16)) : nullptr;  // never dereferenced
  Plan p;
  int rc = plan_collective(dev, rings > 0 ? rings : PDCC_MAX_BLOCKS, has_multicast != 0, *op, &p);
  if (rc) return rc;
  if (algo) *algo = p.algo;
  if (nblocks) *nblocks = p.nblocks;
  if (threads) *threads = p.threads;
  if (tile_bytes) *tile_bytes = p.tile_bytes;
  if (ntiles) *ntiles = p.ntiles;
  if (depth) *depth = p.depth;
  if (slot_stride) *slot_stride = p.slot_stride;
  return PDCC_OK;
}

static int validate_ptrs(const pdcc_op_t& op, int rank, int n) {
  if (op.count == 0 || op.coll == PDCC_BARRIER) return PDCC_OK;
  auto need = [&](const void* p, const char* what) {
    if (!p) {
      set_error("collective %d: %s pointer is null on rank %d", op.coll, what, rank);
      return false;
    }
    return true;
  };
  switch (op.coll) {
    case PDCC_ALL_REDUCE:
      if (!need(op.in, "in") || !need(op.out, "out")) return PDCC_EINVAL;
      break;
    case PDCC_REDUCE:
      if (!need(op.in, "in")) return PDCC_EINVAL;
      if (rank == op.root && !need(op.out, "out")) return PDCC_EINVAL;
      break;
    case PDCC_BROADCAST:
      if (rank == op.root ? !need(op.in, "in") : !need(op.out, "out")) return PDCC_EINVAL;
      break;
    case PDCC_SCATTER:
      if (!need(op.out, "out")) return PDCC_EINVAL;
      if (rank == op.root)
        for (int j = 0; j < n; ++j)
          if (!need(op.list[j], "scatter_list[j]")) return PDCC_EINVAL;
      break;
    case PDCC_GATHER:
      if (!need(op.in, "in")) return PDCC_EINVAL;
      if (rank == op.root)
        for (int j = 0; j < n; ++j)
          if (!need(op.list[j], "gather_list[j]")) return PDCC_EINVAL;
      break;
    case PDCC_ALL_GATHER:
      if (!need(op.in, "in")) return PDCC_EINVAL;
      for (int j = 0; j < n; ++j)
        if (!need(op.list[j], "tensor_list[j]")) return PDCC_EINVAL;
      break;
    default:
      break;
  }
  return PDCC_OK;
}

static void fill_rank_args(RankArgs* ra, const DevComm& dev, const pdcc_op_t& op) {
  ra->comm = dev;
  ra->in = op.in;
  ra->out = op.out;
  for (int j = 0; j < PDCC_MAX_RANKS; ++j) ra->list[j] = op.list[j];
}

// ---------------------------------------------------------------- symmetric memory ----
// Collective: every rank of the communicator calls it with the same size, in the same
// order. The new region is mapped on every rank and (when the communicator has multicast)
// bound to its own multicast object, so any range of it has a peer address on every rank
// and one multicast alias. Regions live until the communicator is destroyed.
int pdcc_sym_alloc(pdcc_comm_t* c, size_t bytes, void** ptr) {
  if (!c || !ptr || bytes == 0) return PDCC_EINVAL;
  if (c->nregions >= kMaxRegions) {
    set_error("at most %d symmetric regions per communicator", kMaxRegions);
    return PDCC_EINVAL;
  }
  CUDA_TRY(cudaSetDevice(c->cfg.device));
  auto* rg = new Region();
  rg->bytes = round_up_sz(bytes, size_t(2) << 20);
  char tag[32];
  snprintf(tag, sizeof(tag), "sym%d/", c->nregions);
  int rc;
  if (c->kind == PDCC_HANDLE_POSIX_FD) rc = create_posix(c, c->store, c->want_mc && c->has_mc, rg, tag);
  else rc = create_legacy(c, c->store, rg, tag);
  if (rc == PDCC_OK && c->cfg.nranks > 1) {
    // nobody may touch the region before every rank has mapped it
    rc = pdcc_store_barrier(c->store, skey(c->cfg, (std::string(tag) + "ready").c_str()).c_str(),
                            c->cfg.nranks, c->cfg.timeout_ms);
  }
  if (rc != PDCC_OK) {
    std::string keep = g_err;
    region_release(c, rg);
    delete rg;
    set_error("%s", keep.c_str());
    return rc;
  }
  c->regions[c->nregions++] = rg;
  *ptr = rg->base[c->cfg.rank];
  return PDCC_OK;
}

// The region that holds [p, p + len) entirely, or nullptr.
static Region* find_region(pdcc_comm* c, const void* p, size_t len, size_t* off) {
  const char* q = static_cast<const char*>(p);
  for (int i = 0; i < c->nregions; ++i) {
    Region* rg = c->regions[i];
    const char* lo = rg->base[c->cfg.rank];
    if (q >= lo && q + len <= lo + rg->bytes) {
      *off = size_t(q - lo);
      return rg;
    }
  }
  // the heap inside the communicator's own block counts as symmetric memory too
  if (c->cfg.heap_bytes) {
    const char* lo = c->blk.base[c->cfg.rank] + c->heap_off;
    if (q >= lo && q + len <= lo + c->cfg.heap_bytes) {
      *off = size_t(q - c->blk.base[c->cfg.rank]);
      return &c->blk;
    }
  }
  return nullptr;
}

int pdcc_sym_query(pdcc_comm_t* c, const void* ptr, size_t bytes, int* is_symmetric,
                   int* has_multicast) {
  if (!c || !ptr) return PDCC_EINVAL;
  size_t off = 0;
  Region* rg = find_region(c, ptr, bytes, &off);
  if (is_symmetric) *is_symmetric = rg != nullptr;
  if (has_multicast) *has_multicast = rg != nullptr && rg->has_mc;
  return PDCC_OK;
}

static void sym_tensor(const pdcc_comm* c, const Region* rg, size_t off, SymTensor* t) {
  t->local = rg->base[c->cfg.rank] + off;
  t->mc = rg->mcp ? rg->mcp + off : nullptr;
  for (int p = 0; p < c->cfg.nranks; ++p) t->peer[p] = rg->base[p] + off;
}

// Decides whether `op` runs on the zero-copy path and, if so, launches it.
//   *taken = 1: launched (or failed with the returned code); 0: use the staged path.
// The decision only depends on things that are the same on every rank under the contract
// "symmetric on every rank or on none" (sym.cuh): size, dtype, op, and whether THIS rank's
// tensor is symmetric.
static int try_symmetric(pdcc_comm* c, const pdcc_op_t& op, void* stream, int* taken) {
  *taken = 0;
  const int n = c->cfg.nranks, me = c->cfg.rank;
  if (n < 2 || !c->has_mc || c->nregions + (c->cfg.heap_bytes ? 1 : 0) == 0) return PDCC_OK;
  if (op.algo != PDCC_ALGO_AUTO && op.algo != PDCC_ALGO_SYM &&
      !(op.coll == PDCC_BROADCAST && op.algo == PDCC_ALGO_NVLS))
    return PDCC_OK;
  const bool forced = op.algo == PDCC_ALGO_SYM;
  const uint64_t bytes = op.count * dtype_size(op.dtype);
  auto refuse = [&](const char* why) {
    if (!forced) return PDCC_OK;
    set_error("zero-copy path not applicable: %s", why);
    return PDCC_EINVAL;
  };
  if (op.coll != PDCC_ALL_REDUCE && op.coll != PDCC_REDUCE && op.coll != PDCC_BROADCAST &&
      op.coll != PDCC_ALL_GATHER)
    return refuse("collective has no zero-copy kernel");
  if (!sym_supported(op.coll, op.dtype, op.redop)) return refuse("the switch cannot reduce this dtype/op");
  if (bytes == 0 || (bytes & 15u)) return refuse("size is not a multiple of 16 bytes");
  if (op.out_dtype != op.dtype) return refuse("fused cast");
  auto aligned = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };

  SymTensor in, out;
  size_t off = 0;
  Region* rg = nullptr;
  switch (op.coll) {
    case PDCC_ALL_REDUCE: {
      if (!aligned(op.in) || !aligned(op.out)) return refuse("tensor not 16-byte aligned");
      rg = find_region(c, op.in, bytes, &off);
      if (!rg || !rg->has_mc) return refuse("input is not in multicast-bound symmetric memory");
      sym_tensor(c, rg, off, &in);
      rg = find_region(c, op.out, bytes, &off);
      if (!rg || !rg->has_mc) return refuse("output is not in multicast-bound symmetric memory");
      sym_tensor(c, rg, off, &out);
      break;
    }
    case PDCC_REDUCE: {
      if (!aligned(op.in)) return refuse("tensor not 16-byte aligned");
      rg = find_region(c, op.in, bytes, &off);
      if (!rg || !rg->has_mc) return refuse("input is not in multicast-bound symmetric memory");
      if (me == op.root && op.out != op.in) {
        set_error("reduce on a symmetric tensor is in place on the destination rank");
        return PDCC_EINVAL;
      }
      sym_tensor(c, rg, off, &in);
      out = in;  // the destination is the root's copy of the same tensor
      break;
    }
    case PDCC_BROADCAST: {
      const void* t = me == op.root ? op.in : op.out;
      if (!aligned(t)) return refuse("tensor not 16-byte aligned");
      rg = find_region(c, t, bytes, &off);
      if (!rg || !rg->has_mc) return refuse("tensor is not in multicast-bound symmetric memory");
      sym_tensor(c, rg, off, &in);
      out = in;
      break;
    }
    default: {  // all_gather: every list[j] symmetric; list[me] is where my shard lands
      if (!aligned(op.in)) return refuse("input not 16-byte aligned");
      for (int j = 0; j < n; ++j) {
        size_t o = 0;
        Region* r = find_region(c, op.list[j], bytes, &o);
        if (!r || !r->has_mc || !aligned(op.list[j])) return refuse("an output tensor is not symmetric");
        if (j == me) {
          rg = r;
          off = o;
        }
      }
      sym_tensor(c, rg, off, &out);
      in.local = const_cast<char*>(static_cast<const char*>(op.in));
      break;
    }
  }
  // tiny messages stay on the flag-in-data kernels (two barriers cost more than they do)
  if (!forced && bytes <= sym_min_bytes(n)) return PDCC_OK;
  // Measured (bench.py "symmetric_memory", profiles/r2): below four ranks the staged push
  // kernels beat the switch for everything but all_gather (2 GPUs, 256 MiB: all_reduce 0.53 ms
  // staged vs 0.71 zero-copy); a large broadcast is better off with the staged two-phase
  // kernel at every group size (8 GPUs, 64 MiB: 161 vs 193 us).
  if (!forced && n < 4 && op.coll != PDCC_ALL_GATHER) return PDCC_OK;
  if (!forced && op.coll == PDCC_BROADCAST && bytes >= (uint64_t(16) << 20)) return PDCC_OK;
  *taken = 1;
  return launch_sym(c->dev, ring_limit(c), op, in, out, stream);
}

int pdcc_collective(pdcc_comm_t* c, const pdcc_op_t* op_in, void* stream) {
  if (!c || !op_in) return PDCC_EINVAL;
  pdcc_op_t op = *op_in;
  if (op.coll != PDCC_ALL_REDUCE) op.out_dtype = op.dtype;
  if (op.count == 0 && op.coll != PDCC_BARRIER) return PDCC_OK;
  int rc = validate_ptrs(op, c->cfg.rank, c->cfg.nranks);
  if (rc) return rc;
  int taken = 0;
  rc = try_symmetric(c, op, stream, &taken);
  if (rc || taken) return rc;
  if (op.algo == PDCC_ALGO_SYM) {
    set_error("zero-copy path requested but the tensors are not symmetric memory of this communicator");
    return PDCC_EINVAL;
  }
  Plan p;
  rc = plan_collective(c->dev, ring_limit(c), c->has_mc, op, &p);
  if (rc) return rc;
  RankArgs ra;
  fill_rank_args(&ra, c->dev, op);
  return launch_collective(&ra, 1, nullptr, op, p, stream);
}

static pdcc_op_t base_op(int coll, uint64_t count, int dtype) {
  pdcc_op_t op;
  memset(&op, 0, sizeof(op));
  op.coll = coll;
  op.count = count;
  op.dtype = dtype;
  op.out_dtype = dtype;
  op.redop = PDCC_SUM;
  op.scale = 1.0f;
  return op;
}

int pdcc_all_reduce(pdcc_comm_t* c, void* buf, uint64_t count, int dtype, int redop,
                    void* stream) {
  pdcc_op_t op = base_op(PDCC_ALL_REDUCE, count, dtype);
  op.in = buf;
  op.out = buf;
  op.redop = redop;
  return pdcc_collective(c, &op, stream);
}

int pdcc_all_reduce_scaled(pdcc_comm_t* c, const void* in, void* out, uint64_t count, int dtype,
                           int out_dtype, float scale, void* stream) {
  pdcc_op_t op = base_op(PDCC_ALL_REDUCE, count, dtype);
  op.in = in;
  op.out = out;
  op.out_dtype = out_dtype;
  op.scale = scale;
  return pdcc_collective(c, &op, stream);
}

int pdcc_reduce(pdcc_comm_t* c, void* buf, uint64_t count, int dtype, int redop, int dst,
                void* stream) {
  pdcc_op_t op = base_op(PDCC_REDUCE, count, dtype);
  op.in = buf;
  op.out = buf;
  op.redop = redop;
  op.root = dst;
  return pdcc_collective(c, &op, stream);
}

int pdcc_broadcast(pdcc_comm_t* c, void* buf, uint64_t count, int dtype, int src, void* stream) {
  pdcc_op_t op = base_op(PDCC_BROADCAST, count, dtype);
  op.in = buf;
  op.out = buf;
  op.root = src;
  return pdcc_collective(c, &op, stream);
}

int pdcc_scatter(pdcc_comm_t* c, void* out, void* const* scatter_list, uint64_t count, int dtype,
                 int src, void* stream) {
  if (!c) return PDCC_EINVAL;
  pdcc_op_t op = base_op(PDCC_SCATTER, count, dtype);
  op.out = out;
  op.root = src;
  if (scatter_list)
    for (int j = 0; j < c->cfg.nranks; ++j) op.list[j] = scatter_list[j];
  return pdcc_collective(c, &op, stream);
}

int pdcc_gather(pdcc_comm_t* c, const void* in, void* const* gather_list, uint64_t count,
                int dtype, int dst, void* stream) {
  if (!c) return PDCC_EINVAL;
  pdcc_op_t op = base_op(PDCC_GATHER, count, dtype);
  op.in = in;
  op.root = dst;
  if (gather_list)
    for (int j = 0; j < c->cfg.nranks; ++j) op.list[j] = gather_list[j];
  return pdcc_collective(c, &op, stream);
}

int pdcc_all_gather(pdcc_comm_t* c, void* const* out_list, const void* in, uint64_t count,
                    int dtype, void* stream) {
  if (!c || !out_list) return PDCC_EINVAL;
  pdcc_op_t op = base_op(PDCC_ALL_GATHER, count, dtype);
  op.in = in;
  for (int j = 0; j < c->cfg.nranks; ++j) op.list[j] = out_list[j];
  return pdcc_collective(c, &op, stream);
}

int pdcc_barrier(pdcc_comm_t* c, void* stream) {
  pdcc_op_t op = base_op(PDCC_BARRIER, 0, PDCC_U8);
  return pdcc_collective(c, &op, stream);
}

}  // extern "C"

// ---------------------------------------------------------------- emulation ----------
struct pdcc_emu {
  int device = 0;
  int nranks = 0;
  int sm_count = 0;
  int max_blocks = 0;
  size_t slot_bytes = 0;
  size_t block_bytes = 0;
  char* blocks[PDCC_MAX_RANKS] = {};
  RankArgs* dev_args = nullptr;
  DevComm dev[PDCC_MAX_RANKS];
};

extern "C" {

int pdcc_emu_destroy(pdcc_emu_t* e) {
  if (!e) return PDCC_EINVAL;
  cudaSetDevice(e->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < e->nranks; ++r)
    if (e->blocks[r]) cudaFree(e->blocks[r]);
  if (e->dev_args) cudaFree(e->dev_args);
  delete e;
  return PDCC_OK;
}

int pdcc_emu_create(int device, int nranks, size_t slot_bytes, int max_blocks,
                    int device_timeout_ms, pdcc_emu_t** out) {
  if (!out || nranks < 1 || nranks > PDCC_MAX_RANKS) {
    set_error("emu: nranks must be 1..%d", PDCC_MAX_RANKS);
    return PDCC_EINVAL;
  }
  pdcc_caps_t caps;
  int rc = pdcc_probe(device, &caps);
  if (rc) return rc;
  if (!caps.cooperative_launch) {
    set_error("emu: device %d has no cooperative launch", device);
    return PDCC_EUNSUPPORTED;
  }
  auto* e = new pdcc_emu();
  e->device = device;
  e->nranks = nranks;
  e->sm_count = caps.sm_count;
  e->slot_bytes = slot_bytes ? round_up_sz(slot_bytes, 4096) : (size_t(256) << 10);
  e->max_blocks = max_blocks > 0 ? max_blocks : 16;
  if (e->max_blocks > PDCC_MAX_BLOCKS) e->max_blocks = PDCC_MAX_BLOCKS;
  // every emulated rank's CTAs must be resident at once
  if (e->max_blocks * nranks > caps.sm_count) e->max_blocks = caps.sm_count / nranks;
  e->block_bytes = kDataOffset + size_t(e->max_blocks) * PDCC_RING_DEPTH * e->slot_bytes;  // rings of DEPTH x slot
  for (int r = 0; r < nranks; ++r) {
    cudaError_t ce = cudaMalloc(reinterpret_cast<void**>(&e->blocks[r]), e->block_bytes);
    if (ce == cudaSuccess) ce = cudaMemset(e->blocks[r], 0, kZeroBytes);
    if (ce != cudaSuccess) {
      set_error("emu: allocating block %d: %s", r, cudaGetErrorString(ce));
      pdcc_emu_destroy(e);
      return PDCC_ECUDA;
    }
  }
  if (cudaMalloc(reinterpret_cast<void**>(&e->dev_args), sizeof(RankArgs) * PDCC_MAX_RANKS) !=
      cudaSuccess) {
    set_error("emu: allocating args");
    pdcc_emu_destroy(e);
    return PDCC_ECUDA;
  }
  for (int r = 0; r < nranks; ++r) {
    DevComm& d = e->dev[r];
    memset(&d, 0, sizeof(d));
    d.rank = r;
    d.nranks = nranks;
    for (int p = 0; p < nranks; ++p) d.base[p] = e->blocks[p];
    d.mc = nullptr;
    d.ring_bytes = uint32_t(e->slot_bytes * PDCC_RING_DEPTH);
    d.timeout_ms = uint32_t(device_timeout_ms > 0 ? device_timeout_ms : 5000);
  }
  cudaDeviceSynchronize();
  *out = e;
  return PDCC_OK;
}

int pdcc_emu_collective(pdcc_emu_t* e, const pdcc_op_t* ops, void* stream) {
  if (!e || !ops) return PDCC_EINVAL;
  CUDA_TRY(cudaSetDevice(e->device));
  pdcc_op_t op0 = ops[0];
  if (op0.coll != PDCC_ALL_REDUCE) op0.out_dtype = op0.dtype;
  if (op0.count == 0 && op0.coll != PDCC_BARRIER) return PDCC_OK;
  RankArgs host[PDCC_MAX_RANKS];
  for (int r = 0; r < e->nranks; ++r) {
    const pdcc_op_t& o = ops[r];
    if (o.coll != op0.coll || o.count != op0.count || o.dtype != op0.dtype ||
        o.redop != op0.redop || o.root != op0.root || o.algo != op0.algo ||
        o.nblocks != op0.nblocks) {
      set_error("emu: rank %d describes a different collective than rank 0", r);
      return PDCC_EINVAL;
    }
    int rc = validate_ptrs(o, r, e->nranks);
    if (rc) return rc;
    fill_rank_args(&host[r], e->dev[r], o);
  }
  Plan p;
  const int limit = e->max_blocks;
  int rc = plan_collective(e->dev[0], limit, false, op0, &p);
  if (rc) return rc;
  if (e->nranks == 1) return launch_collective(host, 1, nullptr, op0, p, stream);
  // by value through a set-up kernel: the stream never reads `host` (common.h)
  rc = upload_rank_args(host, e->nranks, e->dev_args, stream);
  if (rc) return rc;
  return launch_collective(host, e->nranks, e->dev_args, op0, p, stream);
}

int pdcc_emu_check(pdcc_emu_t* e) {
  if (!e) return PDCC_EINVAL;
  CUDA_TRY(cudaSetDevice(e->device));
  CUDA_TRY(cudaDeviceSynchronize());
  for (int r = 0; r < e->nranks; ++r) {
    int rc = check_block(e->blocks[r], r);
    if (rc) return rc;
  }
  return PDCC_OK;
}

}  // extern "C"
