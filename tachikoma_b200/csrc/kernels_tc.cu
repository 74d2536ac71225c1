// tcgen05 (5th-gen tensor core) kernels for sm_100a: persistent int8 GEMM / implicit-GEMM convolution
// with the fused QNN epilogue, and the dense-int8 peak probe.
//
// One CTA per SM, looping over 128 x BN output tiles (static round-robin schedule).  Warp roles:
//   warp 0      TMA producer.  A tile: activations, either a plain K-major matrix (tiled tensor map) or
//               the NHWC activation tensor read through an IM2COL-mode tensor map (implicit GEMM: one
//               filter tap x one channel chunk per pipeline stage; out-of-image taps are zero-filled by
//               the TMA unit and corrected for the input zero point in the epilogue).  B tile: weights
//               [N][K] K-major.  Both land in swizzled shared memory (swizzle span = bytes of K per stage).
//   warp 1      TMEM allocator + MMA issuer: one thread issues tcgen05.mma.kind::i8 (u8/s8 x u8/s8 ->
//               s32) into one of two TMEM accumulator stages, tcgen05.commit frees the smem slot.
//   warps 2..5  epilogue: tcgen05.ld (TMEM -> registers, software-pipelined 32 columns at a time) ->
//               requantize / clip / residual qnn.add -> 16-byte global stores.  While they drain one
//               accumulator stage the MMA warp already fills the other.
// The arithmetic contract of the epilogue is in epilogue.cuh (reference: requantize.cc:195-273,
// intrin_rule.cc:165-250, add.cc:40-93).
#include <cstdlib>
#include <mutex>

#include "tc_kernel.cuh"

namespace tkb {

// ---- host side -----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

static void* driver_entry(const char* name) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qr;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
    return p;
  return nullptr;
}
static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = (EncodeTiledFn)driver_entry("cuTensorMapEncodeTiled");
  return fn;
}
static EncodeIm2colFn encode_im2col() {
  static EncodeIm2colFn fn = (EncodeIm2colFn)driver_entry("cuTensorMapEncodeIm2col");
  return fn;
}

static CUtensorMapSwizzle swizzle_for(int bk) {
  return bk == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : bk == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
}

// 3-D uint8 tensor map [batch][rows][k] (k innermost), box = {bk bytes of k, box_rows, 1}, swizzle span = bk
static bool make_tmap_3d(CUtensorMap* tm, const void* base, long long k, long long rows, long long batch, int box_rows,
                         int bk) {
  EncodeTiledFn fn = encode_tiled();
  if (!fn) return false;
  cuuint64_t dims[3] = {(cuuint64_t)k, (cuuint64_t)rows, (cuuint64_t)batch};
  cuuint64_t strides[2] = {(cuuint64_t)k, (cuuint64_t)(k * rows)};
  cuuint32_t box[3] = {(cuuint32_t)bk, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(bk), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// [outer][rows][k] uint8 map with explicit byte strides: stride_row may be SMALLER than k (overlapping rows)
static bool make_tmap_rows(CUtensorMap* tm, const void* base, long long k, long long rows, long long outer,
                           long long stride_row, long long stride_outer, int box_rows, int bk) {
  EncodeTiledFn fn = encode_tiled();
  if (!fn || (stride_row & 15) || (stride_outer & 15)) return false;
  cuuint64_t dims[3] = {(cuuint64_t)k, (cuuint64_t)rows, (cuuint64_t)outer};
  cuuint64_t strides[2] = {(cuuint64_t)stride_row, (cuuint64_t)stride_outer};
  cuuint32_t box[3] = {(cuuint32_t)bk, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(bk), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// IM2COL-mode map over the NHWC activation tensor {C, W, H, N}: the TMA unit walks 128 consecutive output
// pixels (q fastest, then p, then n) starting at the base pixel given to the instruction, stepping by the
// convolution stride inside the bounding box [lower corner, dim + upper corner); the filter tap is an
// instruction-time offset.  Elements outside the tensor are zero-filled.
static bool make_tmap_im2col(CUtensorMap* tm, const void* base, const ConvGeom& g, int bk) {
  EncodeIm2colFn fn = encode_im2col();
  if (!fn) return false;
  cuuint64_t dims[4] = {(cuuint64_t)g.C, (cuuint64_t)g.W, (cuuint64_t)g.H, (cuuint64_t)g.N};
  cuuint64_t strides[3] = {(cuuint64_t)g.C, (cuuint64_t)g.C * g.W, (cuuint64_t)g.C * g.W * g.H};
  int lower[2] = {-g.pl, -g.pt};
  int upper[2] = {g.pr - (g.KW - 1) * g.dw, g.pb - (g.KH - 1) * g.dh};
  cuuint32_t estr[4] = {1, (cuuint32_t)g.sw, (cuuint32_t)g.sh, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(base), dims, strides, lower, upper,
                  (cuuint32_t)bk, (cuuint32_t)BM, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(bk),
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return false;
  // Drivers up to 13.1 mis-encode im2col maps of tensors smaller than 128 KiB (same work-around as
  // cute/atom/copy_traits_sm90_im2col.hpp make_im2col_tma_copy_desc).
  int drv = 0;
  cudaDriverGetVersion(&drv);
  if (drv <= 13010 && (long long)g.C * g.W * g.H * g.N < 131072) reinterpret_cast<uint64_t*>(tm)[1] &= ~(1ull << 21);
  return true;
}

bool make_tmap_im2col_pub(CUtensorMap* tm, const void* base, const ConvGeom& g, int bk) { return make_tmap_im2col(tm, base, g, bk); }
bool make_tmap_3d_pub(CUtensorMap* tm, const void* base, long long k, long long rows, long long batch, int box_rows, int bk) {
  return make_tmap_3d(tm, base, k, rows, batch, box_rows, bk);
}

// 2-D uint8 map of an output matrix [rows][n] for TMA stores of 128 x 128 tiles out of a 128-byte-swizzled staging buffer
bool make_tmap_out_pub(CUtensorMap* tm, const void* base, long long n, long long rows) {
  EncodeTiledFn fn = encode_tiled();
  if (!fn || (n & 15)) return false;
  cuuint64_t dims[2] = {(cuuint64_t)n, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)n};
  cuuint32_t box[2] = {128, (cuuint32_t)BM};
  cuuint32_t estr[2] = {1, 1};
  return fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool tc_gemm_supported(int M, int N, int K) { return M > 0 && N > 0 && K > 0 && (K % 16) == 0; }

bool tc_implicit_conv_supported(const ConvGeom& g) {
  if (g.groups != 1 || (g.C % 32) != 0) return false;
  const int lo_w = -g.pl, lo_h = -g.pt, up_w = g.pr - (g.KW - 1) * g.dw, up_h = g.pb - (g.KH - 1) * g.dh;
  for (int v : {lo_w, lo_h, up_w, up_h})
    if (v < -128 || v > 127) return false;
  if ((g.KW - 1) * g.dw > 255 || (g.KH - 1) * g.dh > 255) return false;
  if (g.sw < 1 || g.sw > 8 || g.sh < 1 || g.sh > 8) return false;
  return true;
}

static int pick_bn(long long mtiles, int N, int K, bool fast_epi) {
  if (const char* f = getenv("TKB_BN")) {  // tuning experiments only
    const int v = atoi(f);
    if (v == 64 || v == 128 || v == 256) return v;
  }
  const int cand[3] = {256, 128, 64};
  int best = 64;
  long long best_cost = -1;
  const int sms = num_sms();
  for (int bn : cand) {
    if (bn > 64 && bn >= 2 * ((N + 63) / 64 * 64)) continue;  // mostly padding
    const long long tiles = mtiles * ((N + bn - 1) / bn);
    long long cost = ((tiles + sms - 1) / sms) * (bn + 32) * 4;
    // short-K layers (MobileNetV2's expand convolutions, K = 64 .. 160) are epilogue-dominated, and the 64-column kernels
    // have the faster epilogue (tile-split groups, saturating pack): measured 12-15 % faster than BN = 128 at equal waves
    if (bn == 64 && K <= 256 && fast_epi && N > 128 && N % 64 == 0) cost = cost * 3 / 4;
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = bn; }
  }
  return best;
}

static int pick_bk(const TcArgs& a) {
  if (a.implicit == 1) return a.g.C % 128 == 0 ? 128 : (a.g.C % 64 == 0 ? 64 : 32);
  return a.K >= 128 ? 128 : (a.K > 32 ? 64 : 32);
}

void tc_tile_config(const TcArgs& a, int* bn, int* bk) {
  if (a.implicit == 3) { tc_halo_tile_config(a.g, a.N, bn, bk); return; }
  const long long mtiles = a.implicit == 2 ? (long long)a.g.N * a.g.OH * ((a.g.OW + BM - 1) / BM) : (a.M + BM - 1) / BM;
  *bn = pick_bn(mtiles * a.B, a.N, a.K, a.f.enabled != 0 && a.B == 1);
  *bk = pick_bk(a);
}

// 16-byte chunk c of row n goes to chunk c ^ f(n) (the 128 / 64 / 32-byte TMA + UMMA swizzle patterns,
// Swizzle<3,4,3> / <2,4,3> / <1,4,3> over byte offsets with a row pitch of bk bytes).
size_t tc_pack_weights(const uint8_t* w, int N, int K, int bn, int bk, uint8_t* dst) {
  const int ntiles = (N + bn - 1) / bn, num_kb = (K + bk - 1) / bk;
  const size_t tile = (size_t)bn * bk, total = tile * ntiles * num_kb;
  if (!dst) return total;
  memset(dst, 0, total);
  for (int nt = 0; nt < ntiles; ++nt)
    for (int kb = 0; kb < num_kb; ++kb) {
      uint8_t* d = dst + ((size_t)nt * num_kb + kb) * tile;
      for (int r = 0; r < bn; ++r) {
        const int n = nt * bn + r;
        if (n >= N) break;
        const int x = bk == 128 ? (r & 7) : (bk == 64 ? ((r >> 1) & 3) : ((r >> 2) & 1));
        for (int c = 0; c < bk / 16; ++c) {
          const int k0 = kb * bk + c * 16;
          if (k0 >= K) break;
          const int len = K - k0 < 16 ? K - k0 : 16;
          memcpy(d + (size_t)r * bk + ((c ^ x) * 16), w + (size_t)n * K + k0, len);
        }
      }
    }
  return total;
}

int launch_tc(const TcArgs& a, cudaStream_t s) {
  if (a.implicit == 3) return launch_tc_halo(a, s);
  if ((((uintptr_t)a.a) | ((uintptr_t)a.w)) & 15) return -1;
  TcKernelParams p;
  memset(&p, 0, sizeof(p));
  p.y = a.y;
  p.sum = a.e.sum;
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.batches = a.B; p.w_batched = a.w_batched;
  p.e = a.e;
  p.f = a.f;
  if (const char* d = getenv("TKB_DBG")) p.dbg = atoi(d);
  p.delta = a.delta; p.cls_h = a.cls_h; p.cls_w = a.cls_w; p.ncls_w = a.ncls_w;
  const int bk = pick_bk(a);
  if (a.implicit == 2) {  // windowed rows over the unfolded tensor X3 (see launch_unfold_rows_nchw)
    const ConvGeom& g = a.g;
    if (a.K != g.KW * 32 || a.M != g.N * g.OH * g.OW) return -1;
    p.win = 1;
    p.win_rows = g.N * g.OH;
    p.win_qt = (g.OW + BM - 1) / BM;
    p.OW = g.OW; p.OH = g.OH;
  } else if (a.implicit) {
    const ConvGeom& g = a.g;
    if (!tc_implicit_conv_supported(g)) return -1;
    p.implicit = 1;
    p.C = g.C; p.KH = g.KH; p.KW = g.KW; p.OH = g.OH; p.OW = g.OW;
    p.sh = g.sh; p.sw = g.sw; p.dh = g.dh; p.dw = g.dw; p.pt = g.pt; p.pl = g.pl;
  } else {
    if (!tc_gemm_supported(a.M, a.N, a.K)) return -1;
  }
  p.bk = bk;
  const uint32_t layout = bk == 128 ? 2u : (bk == 64 ? 4u : 6u);
  p.desc_hi = (uint32_t)(smem_desc(0, 16, 8 * bk, layout) >> 32);
  p.idesc = ((a.a_dtype == kI8 ? 1u : 0u) << 7) | ((a.w_dtype == kI8 ? 1u : 0u) << 10);
  int bn, bk2;
  tc_tile_config(a, &bn, &bk2);
  if (a.w_packed) {
    if (a.pack_bn != bn || a.pack_bk != bk || a.w_batched) return -1;
    p.wpack = (const uint8_t*)a.w;
  }
  CUtensorMap ta, tb;
  if (a.implicit == 2) {
    const ConvGeom& g = a.g;
    const long long wp = g.W + g.pl + g.pr;
    if (!make_tmap_rows(&ta, a.a, a.K, g.OW, (long long)g.N * g.OH, (long long)g.sw * 32, wp * 32, BM, bk)) return -1;
  } else if (a.implicit) {
    if (!make_tmap_im2col(&ta, a.a, a.g, bk)) return -1;
  } else {
    if (!make_tmap_3d(&ta, a.a, a.K, a.M, a.B, BM, bk)) return -1;
  }
  if (a.w_packed) tb = ta;  // never dereferenced: B tiles come from p.wpack with bulk copies
  else if (!make_tmap_3d(&tb, a.w, a.K, a.N, a.w_batched ? a.B : 1, bn, bk)) return -1;
  int mode = kModeGeneric;
  // the saturating-pack epilogue needs more registers; at BN >= 128 (two TMEM chunks in flight) it spills under the
  // 168-register cap and is slower than the explicit clamps, so it is used for 64-column tiles only
  const bool sat_ok = bn == 64;
  if (a.f.enabled)
    mode = a.f.has_sum ? (a.f.sum_sat && sat_ok ? kModeFastSumSat : kModeFastSum) : (a.f.sat && sat_ok ? kModeFastSat : kModeFast);
  else if (a.e.mode == kEpiRaw) mode = kModeRaw;
  if (mode == kModeGeneric || mode == kModeRaw) return launch_tc_other(bn, mode, ta, tb, p, s);
  if (mode == kModeFastSat) return launch_t<64, kModeFastSat>(ta, tb, p, s);
  if (mode == kModeFastSumSat) return launch_t<64, kModeFastSumSat>(ta, tb, p, s);
#define TKB_GEMM(MODE_)                                                    \
  if (mode == MODE_) {                                                     \
    if (bn == 64) return launch_t<64, MODE_>(ta, tb, p, s);                \
    if (bn == 128) return launch_t<128, MODE_>(ta, tb, p, s);              \
    return launch_t<256, MODE_>(ta, tb, p, s);                             \
  }
  TKB_GEMM(kModeFast) TKB_GEMM(kModeFastSum)
#undef TKB_GEMM
  return -1;
}

// ---- dense int8 peak probe -------------------------------------------------------------------------
// One CTA per SM issues `iters` x 4 tcgen05.mma (128 x 256 x 32, int8) from resident smem tiles into two
// TMEM accumulators; nothing is loaded or stored, so the time is the tensor pipe's issue floor.  This is
// the measured denominator for the conv/dense roofline.
__global__ void __launch_bounds__(128) probe_i8_kernel(int iters, uint32_t idesc, int commit_every) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t done;
  __shared__ uint64_t kb_bar[2];
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (BM * 128 + 256 * 128) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x01010101u;
  if (threadIdx.x == 0) { mbar_init(&done, (commit_every & 16) ? 2 : 1); mbar_init(&kb_bar[0], 1); mbar_init(&kb_bar[1], 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc<512>(&tmem_ptr);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_ptr;
  const int issuers = (commit_every & 16) ? 2 : 1;
  if (warp < issuers) {
    // the whole warp runs the loop; one elected lane issues (elect.sync lets the compiler emit the
    // uniform-datapath UTCIMMA / UTCBAR directly instead of a per-active-thread serialisation loop)
    const uint32_t a_lo = ((smem_u32(smem) >> 4) & 0x3FFF) | (1u << 16);
    const uint32_t b_lo = (((smem_u32(smem) + BM * 128) >> 4) & 0x3FFF) | (1u << 16);
    const uint32_t hi = (uint32_t)(smem_desc(0, 16, 1024, 2) >> 32);
    // commit_every is a bit mask of per-iteration extras: 1 = tcgen05.commit to a scratch mbarrier,
    // 2 = tcgen05.fence::after_thread_sync, 4 = wait for the commit of the iteration two back (what a
    // TMA producer does before it refills a slot), 8 = wait for the commit of THIS iteration (no overlap)
    uint32_t phase[2] = {0, 0};
    for (int it = 0; it < iters; ++it) {
      if (commit_every & 2) tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          mma_i8(tmem_base + (issuers == 2 ? warp * 256 : (it & 1) * 256), ((uint64_t)hi << 32) | (a_lo + 2 * k), ((uint64_t)hi << 32) | (b_lo + 2 * k), idesc, 1);
        if (commit_every & 1) tc_commit(&kb_bar[it & 1]);
      }
      __syncwarp();
      if ((commit_every & 5) == 5 && it >= 2) { mbar_wait(&kb_bar[it & 1], phase[it & 1]); phase[it & 1] ^= 1; }
      if ((commit_every & 9) == 9) { mbar_wait(&kb_bar[it & 1], phase[it & 1]); phase[it & 1] ^= 1; }
    }
    if (elect_one()) {
      tc_commit(&done);
    }
    __syncwarp();
    mbar_wait(&done, 0);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tmem_base);
}

int probe_int8_shape(int iters, int n, int commit_every, double* tops, cudaStream_t s) {
  const int smem = BM * 128 + 256 * 128 + 1024;
  if (cudaFuncSetAttribute(probe_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
  const int sms = num_sms();
  const uint32_t idesc = idesc_i8(128, n, true, true);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  probe_i8_kernel<<<sms, 128, smem, s>>>(iters / 8 + 1, idesc, commit_every);  // warm-up
  cudaEventRecord(e0, s);
  probe_i8_kernel<<<sms, 128, smem, s>>>(iters, idesc, commit_every);
  cudaEventRecord(e1, s);
  g_launches += 2;
  if (cudaEventSynchronize(e1) != cudaSuccess) return -1;
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  const double ops = 2.0 * 128 * n * 32 * 4.0 * (double)iters * sms * ((commit_every & 16) ? 2 : 1);
  *tops = ops / (ms * 1e-3) / 1e12;
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int probe_int8_peak(int iters, double* tops, cudaStream_t s) { return probe_int8_shape(iters, 256, 0, tops, s); }

}  // namespace tkb
