// K9 on the TMA engine: a local device-to-device copy as 1-D bulk async copies
// (cp.async.bulk, SASS UBLKCP) global -> shared -> global through a ring of stages. One
// elected thread per CTA drives the whole pipeline; the SM's load/store units and register
// file are not involved at all, which is what lets a copy sit within a few percent of the
// HBM roofline (the LDG/STG version measured 0.87 of it: profiles/ncu_1gpu_kernels_raw.csv).
//
// Used by the world-size-1 forms of scatter / gather / all_gather (pdcc_local_copy), i.e.
// the 1-GPU case of the reference's list collectives (main.py:37,52,68).
#pragma once

#include <stdint.h>

namespace pdcc {

constexpr int kTmaMaxStages = 16;
// Defaults from benchmarks/copy_sweep.py (profiles/r2/copy_sweep_1gpu.jsonl, 256 MiB fp32):
// 4 x 16 KiB stages, ONE CTA per SM: 0.0848 ms = 6.33 TB/s (0.97 of the measured copy peak;
// torch's D2D copy 0.0843 ms). Deeper or wider is slower: 6 stages x 2 CTAs/SM 0.0927 ms.
constexpr int kTmaStages = 4;
constexpr uint32_t kTmaChunk = 16 * 1024;
constexpr uint32_t kTmaMaxSmemBytes = 200 * 1024;           // what the launcher opts in to

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// global -> shared, completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes,
                                          uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// shared -> global, tracked by bulk groups
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// dst, src 16-byte aligned; bytes a multiple of 16. Chunk k of CTA b is chunk k*grid + b of
// the buffer, so all CTAs stream through the same region of HBM at the same time and finish
// together.
// chunk_bytes (a multiple of 16) and nstages (<= kTmaMaxStages) are launch parameters: the
// dynamic shared memory of the launch is nstages * chunk_bytes.
__global__ void __launch_bounds__(32) tma_copy_kernel(char* __restrict__ dst,
                                                      const char* __restrict__ src,
                                                      uint64_t bytes, uint32_t chunk_bytes,
                                                      int nstages) {
  extern __shared__ __align__(128) unsigned char stage_mem[];
  __shared__ __align__(8) uint64_t bars[kTmaMaxStages];
  if (threadIdx.x != 0) return;
  const uint32_t kTmaChunk = chunk_bytes;  // shadows the default: same loop as before
  const int kTmaStages = nstages;

  const uint64_t nchunks = (bytes + kTmaChunk - 1) / kTmaChunk;
  if (blockIdx.x >= nchunks) return;
  const uint64_t mine = (nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x;
  auto offset = [&](uint64_t k) { return (k * gridDim.x + blockIdx.x) * uint64_t(kTmaChunk); };
  auto length = [&](uint64_t k) {
    const uint64_t off = offset(k);
    return uint32_t(bytes - off < kTmaChunk ? bytes - off : kTmaChunk);
  };

  for (int s = 0; s < kTmaStages; ++s) mbar_init(&bars[s], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");

  const uint64_t ahead = mine < uint64_t(kTmaStages) ? mine : uint64_t(kTmaStages);
  for (uint64_t k = 0; k < ahead; ++k) {
    mbar_arrive_expect_tx(&bars[k], length(k));
    bulk_load(stage_mem + k * kTmaChunk, src + offset(k), length(k), &bars[k]);
  }
  for (uint64_t k = 0; k < mine; ++k) {
    const int s = int(k % kTmaStages);
    const uint32_t parity = uint32_t(k / kTmaStages) & 1u;
    while (!mbar_try_wait(&bars[s], parity)) {
    }
    bulk_store(dst + offset(k), stage_mem + size_t(s) * kTmaChunk, length(k));
    // The stage of chunk k-1 is free once its store has finished READING shared memory:
    // allow the store just issued to be pending, nothing older.
    if (k >= 1 && k - 1 + kTmaStages < mine) {
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      const uint64_t kn = k - 1 + kTmaStages;
      const int sp = int((k - 1) % kTmaStages);
      mbar_arrive_expect_tx(&bars[sp], length(kn));
      bulk_load(stage_mem + size_t(sp) * kTmaChunk, src + offset(kn), length(kn), &bars[sp]);
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // stores done before exit
}

}  // namespace pdcc
