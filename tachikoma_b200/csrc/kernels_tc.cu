// tcgen05 (5th-gen tensor core) kernels for sm_100a: int8 GEMM with the fused QNN epilogue and the
// dense-int8 peak probe.  Warp-specialised: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA
// issuer (one elected thread issues tcgen05.mma kind::i8), warps 2..5 = epilogue (tcgen05.ld from
// TMEM -> requantize -> global).  Operands are staged by TMA into 128-byte-swizzled K-major tiles;
// accumulators (int32) live in TMEM.
#include <cstdio>
#include <mutex>

#include "kernels.h"
#include "tc_ptx.cuh"

namespace tkb {

using namespace ptx;

static constexpr int BM = 128;       // UMMA M (rows = TMEM lanes)
static constexpr int BK = 128;       // bytes of K per stage = one 128B swizzle row
static constexpr int UMMA_K = 32;    // int8 elements per tcgen05.mma

template <int BN, int STAGES>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK;
  static constexpr int B_BYTES = BN * BK;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFF + (2 * STAGES + 1) * 8 + 16 + 1024 /* alignment slack */;
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(192) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                      const __grid_constant__ CUtensorMap tmB, void* __restrict__ y,
                                                      int M, int N, int K, int w_batched, uint32_t idesc,
                                                      EpiParams e) {
  using L = GemmSmem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = (uint64_t*)(smem + L::BAR_OFF);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_ptr = (uint32_t*)(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM, b = blockIdx.z;
  const int num_kb = (K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<BN>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int st = kb % STAGES;
        mbar_wait(&empty[st], ((kb / STAGES) & 1) ^ 1);
        mbar_expect_tx(&full[st], L::STAGE_BYTES);
        tma_load_3d(smem + st * L::STAGE_BYTES, &tmA, &full[st], kb * BK, m0, b);
        tma_load_3d(smem + st * L::STAGE_BYTES + L::A_BYTES, &tmB, &full[st], kb * BK, n0, w_batched ? b : 0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int st = kb % STAGES;
        mbar_wait(&full[st], (kb / STAGES) & 1);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + st * L::STAGE_BYTES);
        const uint32_t b_addr = a_addr + L::A_BYTES;
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          // K-major, 128B swizzle: 8-row groups are 1024 B apart (SBO); advancing K inside the
          // swizzle row is a plain start-address offset.
          const uint64_t da = smem_desc(a_addr + k * UMMA_K, 16, 1024, 2);
          const uint64_t db = smem_desc(b_addr + k * UMMA_K, 16, 1024, 2);
          mma_i8(tmem_base, da, db, idesc, (kb | k) != 0);
        }
        tc_commit(&empty[st]);  // frees the smem slot once the MMAs above have read it
      }
      tc_commit(tmem_full);
    }
  } else {
    // epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31
    const int q = warp & 3;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int row = m0 + q * 32 + lane;
    const long long row_off = ((long long)b * M + row) * N;
    const bool vec8 = (e.out_dtype == kU8 || e.out_dtype == kI8) && (N & 15) == 0 && (((uintptr_t)y) & 15) == 0;
    const bool vec32 = e.out_dtype == kI32 && (N & 3) == 0 && (((uintptr_t)y) & 15) == 0;
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      if (n0 + c >= N) break;  // warp-uniform
      uint32_t r[16];
      tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + c, r);
      tmem_ld_wait();
      if (row < M) {
        int32_t v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int col = n0 + c + j;
          v[j] = col < N ? epilogue_one((int32_t)r[j], col, row_off + col, e) : 0;
        }
        if (vec8 && n0 + c + 16 <= N) {
          uint4 o;
          uint32_t* po = (uint32_t*)&o;
#pragma unroll
          for (int w = 0; w < 4; ++w)
            po[w] = (v[4 * w] & 0xFF) | ((v[4 * w + 1] & 0xFF) << 8) | ((v[4 * w + 2] & 0xFF) << 16) |
                    ((uint32_t)(v[4 * w + 3] & 0xFF) << 24);
          *(uint4*)((uint8_t*)y + row_off + n0 + c) = o;
        } else if (vec32 && n0 + c + 16 <= N) {
          int32_t* yp = (int32_t*)y + row_off + n0 + c;
#pragma unroll
          for (int w = 0; w < 4; ++w) *(int4*)(yp + 4 * w) = make_int4(v[4 * w], v[4 * w + 1], v[4 * w + 2], v[4 * w + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (n0 + c + j < N) store_from_i32(y, e.out_dtype, row_off + n0 + c + j, v[j]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<BN>(tmem_base);
}

// ---- host side -----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  });
  return fn;
}

// 3-D uint8 tensor map [batch][rows][k] (k innermost), box = {128 bytes of k, box_rows, 1}, 128B swizzle
static bool make_tmap_3d(CUtensorMap* tm, const void* base, long long k, long long rows, long long batch, int box_rows) {
  EncodeTiledFn fn = encode_tiled();
  if (!fn) return false;
  cuuint64_t dims[3] = {(cuuint64_t)k, (cuuint64_t)rows, (cuuint64_t)batch};
  cuuint64_t strides[2] = {(cuuint64_t)k, (cuuint64_t)(k * rows)};
  cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

bool tc_gemm_supported(int M, int N, int K) { return M > 0 && N > 0 && K > 0 && (K % 16) == 0; }

template <int BN, int STAGES>
static int launch_gemm_tc_t(const CUtensorMap& ta, const CUtensorMap& tb, void* y, int B, int M, int N, int K,
                            int w_batched, uint32_t idesc, const EpiParams& e, cudaStream_t s) {
  using L = GemmSmem<BN, STAGES>;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(gemm_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL) !=
        cudaSuccess)
      return -1;
    attr_set = true;
  }
  dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, B);
  gemm_tc_kernel<BN, STAGES><<<grid, 192, L::TOTAL, s>>>(ta, tb, y, M, N, K, w_batched, idesc, e);
  ++g_launches;
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int launch_gemm_tc(const void* x, const void* w, void* y, int B, int M, int N, int K, int w_batched, int x_dtype,
                   int w_dtype, const EpiParams& e, cudaStream_t s) {
  if (!tc_gemm_supported(M, N, K)) return -1;
  if ((((uintptr_t)x) | ((uintptr_t)w)) & 15) return -1;
  const int bn = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  CUtensorMap ta, tb;
  if (!make_tmap_3d(&ta, x, K, M, B, BM)) return -1;
  if (!make_tmap_3d(&tb, w, K, N, w_batched ? B : 1, bn)) return -1;
  const uint32_t idesc = idesc_i8(BM, bn, x_dtype == kI8, w_dtype == kI8);
  if (bn == 64) return launch_gemm_tc_t<64, 4>(ta, tb, y, B, M, N, K, w_batched, idesc, e, s);
  if (bn == 128) return launch_gemm_tc_t<128, 3>(ta, tb, y, B, M, N, K, w_batched, idesc, e, s);
  return launch_gemm_tc_t<256, 4>(ta, tb, y, B, M, N, K, w_batched, idesc, e, s);
}

// ---- dense int8 peak probe -------------------------------------------------------------------------
// One CTA per SM issues `iters` x 4 tcgen05.mma (128 x 256 x 32, int8) from resident (uninitialised)
// smem tiles into two TMEM accumulators; nothing is loaded or stored, so the time is the tensor
// pipe's issue floor.  This is the measured denominator for the conv/dense roofline.
__global__ void __launch_bounds__(128) probe_i8_kernel(int iters, uint32_t idesc) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t done;
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (BM * BK + 256 * BK) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x01010101u;
  if (threadIdx.x == 0) { mbar_init(&done, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc<512>(&tmem_ptr);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_ptr;
  if (warp == 0 && lane == 0) {
    const uint32_t a_addr = smem_u32(smem), b_addr = a_addr + BM * BK;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t da = smem_desc(a_addr + k * UMMA_K, 16, 1024, 2);
        const uint64_t db = smem_desc(b_addr + k * UMMA_K, 16, 1024, 2);
        mma_i8(tmem_base + (it & 1) * 256, da, db, idesc, 1);
      }
    }
    tc_commit(&done);
    mbar_wait(&done, 0);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem_base);
}

int probe_int8_peak(int iters, double* tops, cudaStream_t s) {
  const int smem = BM * BK + 256 * BK + 1024;
  if (cudaFuncSetAttribute(probe_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const uint32_t idesc = idesc_i8(128, 256, true, true);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  probe_i8_kernel<<<sms, 128, smem, s>>>(iters / 8 + 1, idesc);  // warm-up
  cudaEventRecord(e0, s);
  probe_i8_kernel<<<sms, 128, smem, s>>>(iters, idesc);
  cudaEventRecord(e1, s);
  g_launches += 2;
  if (cudaEventSynchronize(e1) != cudaSuccess) return -1;
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  const double ops = 2.0 * 128 * 256 * 32 * 4.0 * (double)iters * sms;
  *tops = ops / (ms * 1e-3) / 1e12;
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace tkb
