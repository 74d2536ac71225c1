// L2 -> shared-memory delivery probe (bulk async copies, no consumer): which limit do the operand streams of the
// tcgen05 kernels sit on -- a per-SM ingest rate or a chip-wide L2 rate -- and does cluster multicast raise the
// delivered bytes?   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe_l2 tools/probe_l2.cu
//   modes: 0 = every CTA streams its own region, 1 = every CTA streams the SAME region (weights-like),
//          2 = clusters of 2, each CTA fetches half of every chunk and multicasts it to both (same region for all),
//          3 = clusters of 2, multicast, each cluster its own region
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok)
               : "r"(smem_u32(bar)), "r"(parity)
               : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity))
    if (clock64() - t0 > 400000000LL) return false;  // bounded: a protocol bug ends the kernel, never hangs the GPU
  return true;
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_load_mc(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
      : "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");
}

static constexpr int SLOTS = 6;

// chunk: bytes per ring slot; region: bytes of the source window one CTA (cluster) cycles through
__global__ void __launch_bounds__(32, 1) probe_kernel(const uint8_t* src, size_t buf_bytes, int chunk, size_t region, int iters,
                                                      int mode, int* err) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + (size_t)SLOTS * chunk);
  const bool mc = mode >= 2;
  const uint32_t rank = mc ? cluster_rank() : 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < SLOTS; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  if (mc) cluster_sync_all();
  if (threadIdx.x == 0) {
    const size_t unit = mc ? blockIdx.x / 2 : blockIdx.x;
    const size_t base = (mode == 1 || mode == 2) ? 0 : (unit * region) % (buf_bytes - region + 1);
    const int per = (int)(region / chunk);
    for (int it = 0; it < iters; ++it) {
      const int st = it % SLOTS;
      if (it >= SLOTS && !mbar_wait(&full[st], ((it / SLOTS) - 1) & 1)) { atomicAdd(err, 1); break; }
      mbar_expect_tx(&full[st], (uint32_t)chunk);
      const uint8_t* s = src + base + (size_t)(it % per) * chunk;
      uint8_t* d = smem + (size_t)st * chunk;
      if (!mc) bulk_load(d, s, (uint32_t)chunk, &full[st]);
      else bulk_load_mc(d + rank * (chunk / 2), s + rank * (chunk / 2), (uint32_t)(chunk / 2), &full[st], (uint16_t)3);
    }
    for (int it = (iters > SLOTS ? iters - SLOTS : 0); it < iters; ++it)
      if (!mbar_wait(&full[it % SLOTS], (it / SLOTS) & 1)) { atomicAdd(err, 1); break; }
  }
  __syncwarp();
  if (mc) cluster_sync_all();  // no CTA may exit while its peer still multicasts into it
}

int main() {
  const size_t buf_bytes = 48u << 20;  // L2-resident after the warm-up launch
  uint8_t* buf;
  int* err;
  cudaMalloc(&buf, buf_bytes);
  cudaMemset(buf, 1, buf_bytes);
  cudaMalloc(&err, 4);
  cudaMemset(err, 0, 4);
  int dev_clk = 0;
  cudaDeviceGetAttribute(&dev_clk, cudaDevAttrClockRate, 0);
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int chunks[] = {16384, 32768};
  const int grids[] = {36, 74, 98, 148};
  const size_t regions[] = {294912, 589824};  // 288 KB (own A rows) / 576 KB (a 256 x 2304 weight matrix)
  printf("max SM clock %d MHz\n", dev_clk / 1000);
  for (int mode = 0; mode < 4; ++mode)
    for (int chunk : chunks)
      for (int grid : grids) {
        const size_t region = regions[(mode == 1 || mode == 2) ? 1 : 0] / chunk * chunk;
        const int iters = 600;
        const int smem = SLOTS * chunk + SLOTS * 8 + 1024 + 64;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(32);
        cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = mode >= 2 ? 2 : 1;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
          cudaEventRecord(e0);
          cudaError_t rc = cudaLaunchKernelEx(&cfg, probe_kernel, (const uint8_t*)buf, buf_bytes, chunk, region, iters, mode, err);
          cudaEventRecord(e1);
          if (rc != cudaSuccess || cudaEventSynchronize(e1) != cudaSuccess) {
            printf("mode %d chunk %d grid %d: launch failed: %s\n", mode, chunk, grid, cudaGetErrorString(cudaGetLastError()));
            return 1;
          }
          float ms;
          cudaEventElapsedTime(&ms, e0, e1);
          if (rep > 0 && ms < best) best = ms;
        }
        int herr = 0;
        cudaMemcpy(&herr, err, 4, cudaMemcpyDeviceToHost);
        const double delivered = (double)grid * iters * chunk;  // bytes written into shared memory, all CTAs
        const double tbs = delivered / (best * 1e-3) / 1e12;
        printf("mode %d chunk %5d grid %3d: %7.1f us  delivered %6.2f TB/s  = %5.1f B/clk/SM at %d MHz  (timeouts %d)\n", mode, chunk,
               grid, best * 1e3, tbs, delivered / grid / (best * 1e-3 * dev_clk * 1e3), dev_clk / 1000, herr);
      }
  return 0;
}
