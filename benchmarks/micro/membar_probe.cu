// What does MEMBAR.SYS cost while NVLink stores are in flight, and does it stall the other
// warps of the SM?  Single process, two GPUs with peer access. Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o membar_probe membar_probe.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

__device__ __forceinline__ void st16(char* p, uint4 v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld16(const char* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t gtime() { uint64_t t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

// mode 0: every warp copies src(local) -> dst(peer); no fences.
// mode 1: same, plus after every `chunk` bytes per CTA: __syncthreads, thread 0 fence.sys (lockstep).
// mode 2: warps 1.. copy continuously; warp 0 lane 0 issues fence.sys back to back until the copy is done.
// mode 3: like 1 but the fence is issued by warp 0 (which does not copy) after the copy warps
//         signalled through shared memory; copy warps never wait (decoupled control warp).
__global__ void __launch_bounds__(512, 1) probe(char* dst, const char* src, size_t bytes_per_cta, size_t chunk,
                                                int mode, unsigned long long* stats) {
  __shared__ volatile unsigned done_chunks;
  __shared__ volatile unsigned stop;
  if (threadIdx.x == 0) { done_chunks = 0; stop = 0; }
  __syncthreads();
  const size_t base = size_t(blockIdx.x) * bytes_per_cta;
  const int warp = threadIdx.x >> 5;
  const bool control = (mode == 2 || mode == 3) && warp == 0;
  const int nt = (mode == 2 || mode == 3) ? int(blockDim.x) - 32 : int(blockDim.x);
  const int tid = (mode == 2 || mode == 3) ? int(threadIdx.x) - 32 : int(threadIdx.x);
  const size_t nchunks = bytes_per_cta / chunk;
  if (control) {
    unsigned long long nf = 0, tf = 0;
    if (threadIdx.x == 0) {
      if (mode == 2) {
        while (!stop) { uint64_t t0 = gtime(); asm volatile("fence.acq_rel.sys;" ::: "memory"); tf += gtime() - t0; ++nf; }
      } else {
        unsigned seen = 0;
        const unsigned per = (blockDim.x - 32) / 32;
        while (seen < nchunks) {
          while (done_chunks < (seen + 1) * per) {}
          uint64_t t0 = gtime(); asm volatile("fence.acq_rel.sys;" ::: "memory"); tf += gtime() - t0; ++nf; ++seen;
        }
      }
      if (blockIdx.x == 0) { stats[0] = nf; stats[1] = tf; }
    }
    return;
  }
  unsigned long long tf = 0, nf = 0;
  for (size_t c = 0; c < nchunks; ++c) {
    const size_t off = base + c * chunk;
    const size_t n16 = chunk >> 4;
    for (size_t i = tid; i < n16; i += size_t(nt) * 8) {
      uint4 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) { size_t j = i + size_t(k) * nt; v[k] = ld16(src + off + ((j < n16 ? j : i) << 4)); }
#pragma unroll
      for (int k = 0; k < 8; ++k) { size_t j = i + size_t(k) * nt; if (j < n16) st16(dst + off + (j << 4), v[k]); }
    }
    if (mode == 1) {
      __syncthreads();
      if (threadIdx.x == 0) { uint64_t t0 = gtime(); asm volatile("fence.acq_rel.sys;" ::: "memory"); tf += gtime() - t0; ++nf; }
      __syncthreads();
    } else if (mode == 3) {
      __syncwarp();
      if ((threadIdx.x & 31) == 0) { __threadfence_block(); atomicAdd((unsigned*)&done_chunks, 1u); }
    }
  }
  if (mode == 2) {
    asm volatile("bar.sync 1, %0;" ::"r"(nt) : "memory");
    if (tid == 0) stop = 1;
  }
  if (mode == 1 && blockIdx.x == 0 && threadIdx.x == 0) { stats[0] = nf; stats[1] = tf; }
}

int main(int argc, char** argv) {
  int ndev = 0; CK(cudaGetDeviceCount(&ndev));
  const int peer = ndev > 1 ? 1 : 0;
  CK(cudaSetDevice(0));
  if (peer) { cudaDeviceEnablePeerAccess(1, 0); }
  const size_t total = size_t(1) << 30;
  char *src, *dst; unsigned long long* stats;
  CK(cudaMalloc(&src, total)); CK(cudaMemset(src, 1, total));
  CK(cudaMalloc(&stats, 64));
  CK(cudaSetDevice(peer)); CK(cudaMalloc(&dst, total)); CK(cudaSetDevice(0));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int grids[] = {148, 64};
  for (int gi = 0; gi < 2; ++gi) {
    const int grid = grids[gi];
    const size_t per_cta = (total / grid) & ~size_t((1 << 20) - 1);
    for (int mode = 0; mode < 4; ++mode) {
      for (size_t chunk = 32 << 10; chunk <= (1 << 20); chunk <<= 1) {
        if ((mode == 0 || mode == 2) && chunk != (256 << 10)) continue;
        CK(cudaMemset(stats, 0, 64));
        for (int rep = 0; rep < 2; ++rep) {
          CK(cudaEventRecord(e0));
          probe<<<grid, 512>>>(dst, src, per_cta, chunk, mode, stats);
          CK(cudaEventRecord(e1));
          CK(cudaDeviceSynchronize());
        }
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        unsigned long long h[2]; CK(cudaMemcpy(h, stats, 16, cudaMemcpyDeviceToHost));
        printf("peer=%d grid=%3d mode=%d chunk=%7zu  %8.3f ms  %7.1f GB/s  fences(cta0)=%llu avg_fence_us=%.2f\n", peer, grid, mode,
               chunk, ms, double(per_cta) * grid / ms / 1e6, h[0], h[0] ? double(h[1]) / h[0] / 1e3 : 0.0);
      }
    }
  }
  return 0;
}
