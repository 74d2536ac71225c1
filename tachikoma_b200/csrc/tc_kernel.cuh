// Shared between kernels_tc.cu (folded "fast" epilogue instantiations, host-side tensor maps) and
// kernels_tc_generic.cu (generic epilogue instantiations): the persistent tcgen05 GEMM / implicit-GEMM
// convolution kernel.  See kernels_tc.cu for the design notes.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "kernels.h"
#include "tc_ptx.cuh"

namespace tkb {

using namespace ptx;

static constexpr int BM = 128;       // UMMA M (rows = TMEM lanes)
static constexpr int UMMA_K = 32;    // int8 elements per tcgen05.mma
static constexpr int MAX_STAGES = 12;
static constexpr int EPI_WARPS = 8;        // two GROUPS of four (one warp per TMEM lane quadrant): group g drains the
                                           // tiles whose accumulator lives in TMEM stage g, so two tiles are in the
                                           // epilogue at once (one tile's TMEM-load / barrier latency hides behind
                                           // the other's arithmetic)
static constexpr int EPI_GROUP_WARPS = 4;
static constexpr int EPI_GROUP_THREADS = EPI_GROUP_WARPS * 32;
static constexpr int EPI_THREADS = EPI_WARPS * 32;
static constexpr int TC_THREADS = 64 + EPI_THREADS;
// EG (template parameter of the single-layer kernels) = number of epilogue groups.  2: the default.  1: a 192-thread CTA
// (producer, issuer, four epilogue warps) launched TWO per SM, for layers of very many tiles with a handful of MMAs each
// (MobileNetV2's pointwise convolutions, K = 16 .. 96): there the limit is the ~860 cycles of mbarrier / commit round
// trips per tile on the producer and issuer threads (profiles/r02_notes.md section 13), and two CTAs per SM run two such
// chains side by side with the same eight epilogue warps per SM.  4 (sixteen epilogue warps, four accumulator stages)
// was measured and is no faster anywhere: 18 warps cap the kernel at 96 registers and the epilogue is not latency bound.
__host__ __device__ constexpr int tc_threads(int eg) { return 64 + eg * EPI_GROUP_THREADS; }
#define TKB_TC_BOUNDS(EG) __launch_bounds__(tc_threads(EG), (EG) == 1 ? 2 : 1)

struct TcKernelParams {
  void* y;
  const void* sum;        // residual tensor (same shape as y) or nullptr
  int M, N, K;            // GEMM view; implicit conv: K = KH*KW*C
  int batches, w_batched; // batched GEMM (qnn.batch_matmul): tensor-map z coordinate
  int bk;                 // bytes of K per pipeline stage: 128 / 64 / 32 (= swizzle span)
  int stages;
  int kps;                // k-blocks per pipeline stage: one mbarrier round trip covers kps * bk/32 MMAs
  uint32_t idesc;
  uint32_t desc_hi;       // upper 32 bits of the UMMA smem descriptors (SBO, version, swizzle mode)
  // implicit convolution
  int implicit, C, KH, KW, OH, OW, sh, sw, dh, dw, pt, pl;
  // windowed rows (A = overlapping windows of an unfolded tensor): tile = 128 pixels of ONE output row
  int win, win_rows, win_qt;  // rows = N*OH, qt = ceil(OW / 128)
  // zero-point correction of taps that fall into the padding (see runtime.cu border_classes())
  const int32_t* delta;   // [n_classes][N] or nullptr
  const uint8_t* cls_h;   // [OH]
  const uint8_t* cls_w;   // [OW]
  int ncls_w;
  // halo-reuse convolution (tc_halo_kernel): tile = hR output rows of one image, A = one (hR+KH-1) x hWp pixel
  // patch per channel chunk, taps = shifted UMMA descriptors into the patch, weights resident in shared memory
  int hR, hWp, hRows, hTpi, a_slot, a_stages;
  const uint8_t* wpack;   // packed weight tiles (tc_pack_weights) or nullptr: B comes through tmB
  long long* tl;          // TKB_TIMELINE: clock64 stamps of CTA 0 (see tl_stamp) or nullptr
  int dbg;                // timing experiments only (TKB_DBG): 1 = no epilogue work, 2 = no B loads, 4 = no A loads
  EpiParams e;            // generic epilogue (any rounding / compute dtype / float mode / raw int32)
  FastEpi f;              // host-folded fixed-point epilogue (default QNN configuration)
  int4 ctab[64];          // copy of f.tabA for layers with exactly 64 output channels: read as constant-bank operands by the
                          // CTAB epilogue (no shared-memory traffic for the per-channel constants)
  static constexpr bool kChain = false;  // single-layer launch: the residual was written by an EARLIER kernel (ld.global.nc is safe)
  static constexpr bool kPair = false;   // true in the cta_group::2 kernel: accumulator stages are released to the LEADER CTA's barrier
};

// Device timeline of CTA 0 (TKB_TIMELINE=1, tools/conv_probe.py): slot -> clock64.  0 entry, 1 prologue done,
// 2 griddepcontrol.wait passed, 3 first operands landed (MMA warp), 4 first tile issued, 5 all tiles issued,
// 6 first accumulator ready (epilogue), 7 first tile drained, 8 all tiles drained, 9 teardown; 16.. = issue of the
// first tile's k-groups (GEMM kernel) / of the CTA's tiles (halo kernel).
// Compiled in only with -DTKB_TIMELINE (TKB_BUILD_TIMELINE=1 python -m tachikoma_b200.build --force): the stamps
// cost ~2 % of a ResNet-18 step even when disabled at run time (hot-loop instructions, instruction cache).
template <class P>
__device__ __forceinline__ void tl_stamp(const P& p, int slot) {
  if (p.tl != nullptr && blockIdx.x == 0) p.tl[slot] = clock64();
}
#ifdef TKB_TIMELINE
#define TKB_TL_IF(cond, p, slot) do { if (cond) tl_stamp(p, slot); } while (0)
#else
#define TKB_TL_IF(cond, p, slot) do { } while (0)
#endif
inline long long* timeline_buf() {
  static long long* buf = [] {
    long long* b = nullptr;
#ifdef TKB_TIMELINE
    if (getenv("TKB_TIMELINE") && cudaMalloc(&b, 64 * sizeof(long long)) != cudaSuccess) b = nullptr;
#endif
    return b;
  }();
  if (buf) cudaMemset(buf, 0, 64 * sizeof(long long));
  return buf;
}
inline void timeline_print(const char* what, long long* buf, cudaStream_t s) {
  if (!buf) return;
  long long h[64];
  if (cudaStreamSynchronize(s) != cudaSuccess || cudaMemcpy(h, buf, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) return;
  fprintf(stderr, "timeline %s (cycles since entry):", what);
  for (int i = 1; i < 10; ++i) fprintf(stderr, " [%d]%lld", i, h[i] ? h[i] - h[0] : -1);
  fprintf(stderr, " | issue stamps (cycles since [3]):");
  for (int i = 16; i < 64 && h[i]; ++i) fprintf(stderr, " %lld", h[i] - h[3]);
  fprintf(stderr, "\n");
}

__device__ __forceinline__ void tma_load_im2col_4d(void* dst, const CUtensorMap* m, uint64_t* bar, int c, int w, int h,
                                                   int n, uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2], {%7, %8};" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}

// 32 lanes x 32 columns of 32-bit: thread i of the warp gets lane (base_lane + i), columns c..c+31
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
// tcgen05.wait::ld with a register dependency on the loaded values, so that no consumer of r[] can be
// scheduled above the wait.
__device__ __forceinline__ void tmem_ld_wait32(uint32_t (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                 "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                 "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}
template <int EG = 2>
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(EG * EPI_GROUP_THREADS) : "memory"); }
// named barrier of one epilogue group (ids 2 .. 2 + EG - 1)
__device__ __forceinline__ void epi_group_sync(int grp) {
  asm volatile("bar.sync %0, %1;" ::"r"(2 + grp), "n"(EPI_GROUP_THREADS) : "memory");
}

__device__ __forceinline__ uint4 ldg_nc_16(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
// Coherent 16-byte load (L2, not the non-coherent path): for data another CTA of the SAME launch has written and
// published with a release at gpu scope (chained layers: the residual of layer k is the output of layer k-2).
__device__ __forceinline__ uint4 ldg_cg_16(const void* p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void stg256(void* p, const uint32_t (&w)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]),
               "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
               : "memory");
}

// ---- epilogue -------------------------------------------------------------------------------------
// MODE selects the inlined hot path of the epilogue warps; everything unusual (ragged N tails, outputs
// whose rows are not 16-byte multiples, border rows that need the zero-point delta, the generic
// rounding / compute-dtype / float variants) goes element by element through one out-of-line function,
// which keeps the hot loop small enough to stay in the instruction cache.
// kModeFastSat / kModeFastSumSat: same arithmetic, the final clamp is the output dtype's own range and is done by
// the saturating pack (separate instantiations keep each kernel's hot loop small: instruction cache)
enum TcMode : int { kModeGeneric = 0, kModeFast = 1, kModeFastSum = 2, kModeRaw = 3, kModeFastSat = 5, kModeFastSumSat = 6 };
__host__ __device__ constexpr bool mode_is_fast(int m) { return (m & 3) == kModeFast || (m & 3) == kModeFastSum; }
__host__ __device__ constexpr bool mode_is_sum(int m) { return (m & 3) == kModeFastSum; }
__host__ __device__ constexpr bool mode_is_sat(int m) { return (m & 4) != 0; }

__device__ __forceinline__ int4 lds128(uint32_t addr) {
  int4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, int4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ int32_t fast_requant(int32_t a, const int4& t, int32_t lo, int32_t hi) {
  const long long c = ((long long)t.w << 32) | (unsigned long long)(uint32_t)t.z;
  const long long v = (long long)a * (long long)t.x + c;
  const int32_t y = (int32_t)(v >> 32) >> t.y;
  return min(max(y, lo), hi);
}
__device__ __forceinline__ int32_t fast_side(int32_t q, const FastSide& f) {
  const long long v = (long long)(q << f.ls) * (long long)f.M + f.C;
  return (int32_t)(v >> 32) >> f.s1;
}

// cvt.pack.sat: two int32 -> two saturated 8-bit values in the low half of d, the low half of c moved up:
// d = (c << 16) | (sat8(a) << 8) | sat8(b).  Saturation to the full range of the output dtype IS the final clamp
// whenever the clip bounds are the dtype's own range, which saves two min/max and the shift/or packing per element.
template <bool UNSIGNED>
__device__ __forceinline__ uint32_t pack4_sat8(int32_t v0, int32_t v1, int32_t v2, int32_t v3) {
  uint32_t t, d;
  if (UNSIGNED) {
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(v3), "r"(v2), "r"(0));
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(v1), "r"(v0), "r"(t));
  } else {
    asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(v3), "r"(v2), "r"(0));
    asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(v1), "r"(v0), "r"(t));
  }
  return d;
}
__device__ __forceinline__ int32_t fast_requant_raw(int32_t a, const int4& t) {
  const long long c = ((long long)t.w << 32) | (unsigned long long)(uint32_t)t.z;
  return (int32_t)(((long long)a * (long long)t.x + c) >> 32) >> t.y;
}

// one element, any configuration
template <class P>
static __device__ __noinline__ void epi_slow_elem(int32_t acc, int col, long long idx, const int32_t* drow,
                                           const P* p, int mode) {
  if (drow) acc = (int32_t)((uint32_t)acc + (uint32_t)__ldg(drow + col));
  int32_t v = 0;
  if (mode_is_fast(mode)) {
    const FastEpi& f = p->f;
    v = fast_requant(acc, __ldg(f.tabA + col), f.lo, f.hi);
    if (mode_is_sum(mode)) {
      const int32_t b = (int32_t)(((const volatile uint8_t*)p->sum)[idx] ^ (uint8_t)f.sum_xor);
      v = min(max(fast_side(v, f.lhs) + fast_side(b, f.rhs), f.sum_lo), f.sum_hi);
    }
  } else {
    if constexpr (!P::kChain) v = epilogue_one(acc, col, idx, p->e);
  }
  store_from_i32(p->y, p->e.out_dtype, idx, v);
}

// 32 columns of one output row.  r: raw TMEM accumulators; ncols: valid columns (1..32); col0: absolute
// first column; idx0: linear output index of (row, col0); tab_addr: shared address of this chunk's table.
// Per-channel constants {M, T - 32, C_lo, C_hi} of column j of the chunk: from the shared-memory table, or (CT >= 0:
// compile-time first column of the chunk) straight from the kernel parameters, i.e. as constant-bank operands.
template <int CT, class P>
__device__ __forceinline__ int4 tab_entry(const P& p, uint32_t tab_addr, int j) {
  if constexpr (CT >= 0) return p.ctab[CT + j];
  else return lds128(tab_addr + j * 16);
}

template <int MODE, bool AL8 = false, class P = TcKernelParams, int CT = -1>
__device__ __forceinline__ void epilogue_chunk(const uint32_t (&r)[32], const P& p, uint32_t tab_addr,
                                               bool row_valid, int ncols, int col0, long long idx0,
                                               const int32_t* drow, bool vec_ok) {
  if (!row_valid) return;
  // Rows whose window overlaps the padding add their border class's per-channel delta first (runtime.cu border_plan).
  // The fast modes do that inside the hot path (chained launches: a quarter to a half of the pixels of a 14x14 / 7x7
  // feature map are border pixels); the single-layer kernels keep sending those rows through the out-of-line path.
  // Ragged chunks (N = 16, 24, 144, ...: MobileNetV2's projection layers) stay on the vector path when the valid columns
  // are whole store vectors: all 32 values are computed (table entries beyond N are zero) and only the valid 16-byte
  // (8-byte for AL8 kernels) pieces are stored.  Everything else goes element by element through the out-of-line path.
  const bool ragged_ok = mode_is_fast(MODE) && ncols > 0 && (AL8 ? (ncols & 7) == 0 : (ncols & 15) == 0);
  const bool hot = MODE != kModeGeneric && (ncols == 32 || ragged_ok) && vec_ok && (drow == nullptr || (P::kChain && mode_is_fast(MODE)));
  if (!hot) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < ncols) epi_slow_elem((int32_t)r[j], col0 + j, idx0 + j, drow, &p, MODE);
    return;
  }
  if constexpr (MODE == kModeRaw) {
    int32_t* dst = (int32_t*)p.y + idx0;
    const int32_t* bias = p.e.bias;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      int4 b = make_int4(0, 0, 0, 0);
      if (bias) b = __ldg((const int4*)(bias + col0) + i);
      ((int4*)dst)[i] = make_int4((int32_t)(r[4 * i] + (uint32_t)b.x), (int32_t)(r[4 * i + 1] + (uint32_t)b.y),
                                  (int32_t)(r[4 * i + 2] + (uint32_t)b.z), (int32_t)(r[4 * i + 3] + (uint32_t)b.w));
    }
  } else if constexpr (mode_is_fast(MODE)) {
    const FastEpi& f = p.f;
    uint32_t sraw[8];
    if constexpr (mode_is_sum(MODE)) {
      const uint4 s0 = P::kChain ? ldg_cg_16((const uint8_t*)p.sum + idx0) : ldg_nc_16((const uint8_t*)p.sum + idx0);
      uint4 s1 = make_uint4(0, 0, 0, 0);  // a ragged chunk (16 valid columns) must not read past the row
      if (ncols > 16) s1 = P::kChain ? ldg_cg_16((const uint8_t*)p.sum + idx0 + 16) : ldg_nc_16((const uint8_t*)p.sum + idx0 + 16);
      sraw[0] = s0.x ^ f.sum_xor; sraw[1] = s0.y ^ f.sum_xor; sraw[2] = s0.z ^ f.sum_xor; sraw[3] = s0.w ^ f.sum_xor;
      sraw[4] = s1.x ^ f.sum_xor; sraw[5] = s1.y ^ f.sum_xor; sraw[6] = s1.z ^ f.sum_xor; sraw[7] = s1.w ^ f.sum_xor;
    }
    uint32_t w[8];
    // sat: 0 = explicit clamps, 1 = final clamp is the uint8 range, 2 = final clamp is the int8 range
    if constexpr (mode_is_sat(MODE)) {
      const int sat = mode_is_sum(MODE) ? f.sum_sat : f.sat;  // 1: uint8 range, 2: int8 range
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        int32_t v[4];
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int j = 4 * i + jj;
          int32_t y = fast_requant_raw((int32_t)r[j], tab_entry<CT>(p, tab_addr, j));
          if constexpr (mode_is_sum(MODE)) {
            y = min(max(y, f.lo), f.hi);
            const int32_t b = (int32_t)((sraw[i] >> (8 * jj)) & 0xFFu);
            y = fast_side(y, f.lhs) + fast_side(b, f.rhs);
          }
          v[jj] = y;
        }
        w[i] = sat == 1 ? pack4_sat8<true>(v[0], v[1], v[2], v[3]) : pack4_sat8<false>(v[0], v[1], v[2], v[3]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        uint32_t word = 0;
        int4 dv = make_int4(0, 0, 0, 0);
        if constexpr (P::kChain) {
          if (drow != nullptr) dv = __ldg((const int4*)(drow + col0) + i);
        }
        const int32_t dl[4] = {dv.x, dv.y, dv.z, dv.w};
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int j = 4 * i + jj;
          int32_t y = fast_requant((int32_t)(r[j] + (uint32_t)dl[jj]), tab_entry<CT>(p, tab_addr, j), f.lo, f.hi);
          if constexpr (mode_is_sum(MODE)) {
            const int32_t b = (int32_t)((sraw[i] >> (8 * jj)) & 0xFFu);
            y = min(max(fast_side(y, f.lhs) + fast_side(b, f.rhs), f.sum_lo), f.sum_hi);
          }
          word |= (uint32_t)(y & 0xFF) << (8 * jj);
        }
        w[i] = word;
      }
    }
    uint8_t* dst8 = (uint8_t*)p.y + idx0;
    if (ncols == 32 && (((uintptr_t)dst8) & 31) == 0) {
      // one 256-bit store per lane (sm_100: STG.256): the lanes of a warp write 32 different rows, so every store
      // instruction costs the LSU one wavefront per lane; two 128-bit stores cost twice that and write each 32-byte
      // sector in two halves (profiles/r02_notes.md section 17)
      stg256(dst8, w);
    } else if ((!AL8 || (((uintptr_t)dst8) & 15) == 0) && (ncols & 15) == 0) {
      uint4* dst = (uint4*)dst8;
      dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
      if (ncols == 32) dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
    } else {  // rows that are 8- but not 16-byte multiples (e.g. 1000 classes), ragged chunks of 8 / 24 columns
      uint2* dst = (uint2*)dst8;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (8 * i < ncols) dst[i] = make_uint2(w[2 * i], w[2 * i + 1]);
    }
  }
}

// Hand an accumulator stage back to the MMA issuer.  In the cta_group::2 kernel the issuer lives in the pair's leader CTA
// (rank 0): both CTAs' epilogue warps arrive on ITS barrier, through the cluster address space, with cluster-scope release.
template <class P>
__device__ __forceinline__ void acc_release(uint64_t* bar) {
  if constexpr (P::kPair) {
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(ra) : "r"(smem_u32(bar)));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(ra) : "memory");
  } else {
    mbar_arrive(bar);
  }
}

// Drain one accumulator stage: the calling epilogue warp owns TMEM lanes 32*q .. +31 and columns
// [cbeg0, cbeg0 + cspan) of the tile; `row` is the output row (pixel) of this lane.
template <int BN, int MODE, bool AL8 = false, class P = TcKernelParams, bool CTAB = false, int AS = 2, int EG = 2>
__device__ __forceinline__ void epi_drain_tile(const P& p, uint32_t tmem_base, uint32_t lt, int q, int cbeg0, int cspan,
                                               int lane, int row, bool row_valid, int b, int n0, uint32_t tab_base,
                                               bool vec_ok, uint64_t* tempty, uint64_t* tfull) {
  const long long row_off = ((long long)b * p.M + row) * p.N;
  const int32_t* drow = nullptr;
  if (p.delta != nullptr && row_valid) {
    const int qq = row % p.OW, pp = (row / p.OW) % p.OH;
    const int cls = (int)p.cls_h[pp] * p.ncls_w + (int)p.cls_w[qq];
    if (cls) drow = p.delta + (long long)cls * p.N;
  }
  const uint32_t as = lt % AS;  // AS accumulator stages of BN columns
  mbar_wait(&tfull[as], (lt / AS) & 1);
  tc_fence_after();
  TKB_TL_IF(lt == 0 && q == 2 && lane == 0 && cbeg0 == 0, p, 6);
  const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
  const int ncol_tile = min(BN, p.N - n0);
  const int cbeg = cbeg0, cend = min(cbeg + cspan, ncol_tile);
  if (cbeg >= cend) {  // nothing to drain for this warp (ragged N tail): just release the stage
    tc_fence_before();
    __syncwarp();
    if (lane == 0) acc_release<P>(&tempty[as]);
    return;
  }
  if (p.dbg & 1) {
    tc_fence_before();
    __syncwarp();
    if (lane == 0) acc_release<P>(&tempty[as]);
    return;
  }
  uint32_t ra[32], rb[32];
  if constexpr (CTAB) {
    // 64-column tile owned entirely by this warp, N == 64: both chunks with compile-time column numbers
    static_assert(BN == 64, "CTAB epilogue: 64-column tiles only");
    tmem_ld32(t_addr, ra);
    tmem_ld_wait32(ra);
    tmem_ld32(t_addr + 32, rb);
    epilogue_chunk<MODE, AL8, P, 0>(ra, p, 0, row_valid, 32, n0, row_off + n0, drow, vec_ok);
    tmem_ld_wait32(rb);
    tc_fence_before();
    __syncwarp();
    if (lane == 0) acc_release<P>(&tempty[as]);
    epilogue_chunk<MODE, AL8, P, 32>(rb, p, 0, row_valid, 32, n0 + 32, row_off + n0 + 32, drow, vec_ok);
    return;
  }
  tmem_ld32(t_addr + cbeg, ra);
#pragma unroll 1
  for (int c = cbeg;; c += 64) {
    tmem_ld_wait32(ra);
    const bool more_b = c + 32 < cend;
    if (more_b) tmem_ld32(t_addr + c + 32, rb);
    else {  // all TMEM reads of this warp are complete: hand the accumulator stage back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) acc_release<P>(&tempty[as]);
    }
    epilogue_chunk<MODE, AL8, P>(ra, p, tab_base + c * 16, row_valid, min(32, cend - c), n0 + c, row_off + n0 + c, drow, vec_ok);
    if (!more_b) break;
    tmem_ld_wait32(rb);
    const bool more_a = c + 64 < cend;
    if (more_a) tmem_ld32(t_addr + c + 64, ra);
    else {
      tc_fence_before();
      __syncwarp();
      if (lane == 0) acc_release<P>(&tempty[as]);
    }
    epilogue_chunk<MODE, AL8, P>(rb, p, tab_base + (c + 32) * 16, row_valid, min(32, cend - c - 32), n0 + c + 32,
                         row_off + n0 + c + 32, drow, vec_ok);
    if (!more_a) break;
  }
}

template <int BN, int MODE, bool SPLIT, int EG = 2>
__global__ void TKB_TC_BOUNDS(EG) tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA,
                                                         const __grid_constant__ CUtensorMap tmB,
                                                         const __grid_constant__ TcKernelParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int a_bytes = BM * p.bk, b_bytes = BN * p.bk, sub_bytes = a_bytes + b_bytes, stage_bytes = p.kps * sub_bytes;
  int4* tab_s = (int4*)(smem + p.stages * stage_bytes);  // one table of BN entries per epilogue group
  uint64_t* full = (uint64_t*)(tab_s + EG * BN);
  uint64_t* empty = full + MAX_STAGES;
  uint64_t* tfull = empty + MAX_STAGES;
  uint64_t* tempty = tfull + 4;
  uint32_t* tmem_ptr = (uint32_t*)(tempty + 4);
  // accumulator stages: one per epilogue group when the groups take alternate tiles, else two
  constexpr int AS = SPLIT ? EG : 2;
  static_assert(AS * BN <= 512, "TMEM columns");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mtiles = p.win ? p.win_rows * p.win_qt : (p.M + BM - 1) / BM, ntiles = (p.N + BN - 1) / BN;
  const int total_tiles = (p.dbg & 8) ? 0 : mtiles * ntiles * p.batches;  // (host: < 2^30) dbg 8: launch + prologue + teardown only
  const int cchunks = p.implicit ? p.C / p.bk : 0;
  const int num_kb = p.implicit ? p.KH * p.KW * cchunks : (p.K + p.bk - 1) / p.bk;

  // Programmatic dependent launch: let the next kernel of the stream start its own prologue as soon as SMs
  // free up; everything this kernel reads or writes in global memory comes after griddepcontrol.wait below.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  TKB_TL_IF(threadIdx.x == 0, p, 0);
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < p.stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < AS; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], SPLIT ? EPI_GROUP_WARPS : EG * EPI_GROUP_WARPS); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<AS * BN>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  TKB_TL_IF(threadIdx.x == 0, p, 1);
  // wait for the preceding kernel(s) of the stream to complete and flush (no-op without the PDL launch attribute)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  TKB_TL_IF(threadIdx.x == 0, p, 2);

  // Producer and MMA warps run their loops with the whole warp converged (all lanes poll the mbarriers) and
  // guard only the issuing instructions with elect.sync: the compiler then emits the uniform-datapath
  // UTMALDG / UTCIMMA / UTCBAR directly.  (An `if (lane == 0)` around the loop makes it wrap each of them
  // in a per-active-thread serialisation loop, which cost ~300 cycles per MMA in the first version.)
  if (warp == 0) {
    // ===== TMA producer =====
    int st = 0;
    uint32_t ph = 0;
    const bool ld_a = !(p.dbg & 4), ld_b = !(p.dbg & 2);
    const uint32_t tx_bytes = (uint32_t)((ld_a ? a_bytes : 0) + (ld_b ? b_bytes : 0));
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {  // 32-bit tile decode (64-bit divisions: ~200 cycles each)
      const int nt = tile % ntiles;
      const int t2 = tile / ntiles;
      const int mt = t2 % mtiles, b = t2 / mtiles;
      const int n0 = nt * BN, m0 = mt * BM;
      int cw = 0, chh = 0, cn = 0;
      if (p.implicit) {
        const int q0 = m0 % p.OW, t3 = m0 / p.OW;
        cw = q0 * p.sw - p.pl;
        chh = (t3 % p.OH) * p.sh - p.pt;
        cn = t3 / p.OH;
      }
      int r = 0, s = 0, cc = 0;
      for (int kb0 = 0; kb0 < num_kb; kb0 += p.kps) {
        const int nsub = min(p.kps, num_kb - kb0);
        mbar_wait(&empty[st], ph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&full[st], (uint32_t)nsub * tx_bytes);
          for (int j = 0; j < nsub; ++j) {
            uint8_t* sa = smem + st * stage_bytes + j * sub_bytes;
            const int kb = kb0 + j;
            if (ld_a) {
              if (p.implicit) {
                tma_load_im2col_4d(sa, &tmA, &full[st], cc * p.bk, cw, chh, cn, (uint16_t)(s * p.dw), (uint16_t)(r * p.dh));
                if (++cc == cchunks) { cc = 0; if (++s == p.KW) { s = 0; ++r; } }
              } else if (p.win) {
                tma_load_3d(sa, &tmA, &full[st], kb * p.bk, (mt % p.win_qt) * BM, mt / p.win_qt);
              } else {
                tma_load_3d(sa, &tmA, &full[st], kb * p.bk, m0, b);
              }
            }
            if (ld_b) {
              if (p.wpack) bulk_load(sa + a_bytes, p.wpack + ((size_t)nt * num_kb + kb) * b_bytes, (uint32_t)b_bytes, &full[st]);
              else tma_load_3d(sa + a_bytes, &tmB, &full[st], kb * p.bk, n0, p.w_batched ? b : 0);
            }
          }
        }
        // the tap / chunk counters live in the elected lane only; every lane needs them for the next group
        r = __shfl_sync(0xffffffffu, r, 0); s = __shfl_sync(0xffffffffu, s, 0); cc = __shfl_sync(0xffffffffu, cc, 0);
        if (++st == p.stages) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    int st = 0;
    uint32_t ph = 0, lt = 0;
    const int kper = p.bk / UMMA_K;
    const uint32_t smem_base = smem_u32(smem);
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const uint32_t as = lt % AS;
      mbar_wait(&tempty[as], ((lt / AS) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      for (int kb0 = 0; kb0 < num_kb; kb0 += p.kps) {
        const int nsub = min(p.kps, num_kb - kb0);
        mbar_wait(&full[st], ph);
        tc_fence_after();
        TKB_TL_IF(lt == 0 && kb0 == 0 && lane == 0, p, 3);
        if (elect_one()) {
          // K-major swizzled tiles: 8-row groups are 8*bk bytes apart (SBO, in desc_hi); advancing K inside the
          // swizzle row is a plain start-address offset (+32 B = +2 in the descriptor's 16-byte units).
          const uint64_t hi = (uint64_t)p.desc_hi << 32;
          uint32_t acc = kb0 != 0;
          for (int j = 0; j < nsub; ++j) {
            const uint32_t a_addr = smem_base + st * stage_bytes + j * sub_bytes;
            const uint32_t a_lo = ((a_addr >> 4) & 0x3FFF) | (1u << 16);
            const uint32_t b_lo = (((a_addr + a_bytes) >> 4) & 0x3FFF) | (1u << 16);
            for (int k = 0; k < kper; ++k) {
              mma_i8(d_tmem, hi | (a_lo + 2 * k), hi | (b_lo + 2 * k), p.idesc, acc);
              acc = 1;
            }
          }
          tc_commit(&empty[st]);  // frees the smem slot once the MMAs above have read it
          if (kb0 + nsub >= num_kb) tc_commit(&tfull[as]);
          TKB_TL_IF(lt == 0 && kb0 / p.kps < 48, p, 16 + kb0 / p.kps);
        }
        __syncwarp();
        if (++st == p.stages) { st = 0; ph ^= 1; }
      }
      TKB_TL_IF(lt == 0 && lane == 0, p, 4);
    }
    TKB_TL_IF(lane == 0, p, 5);
  } else {
    // ===== epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31 =====
    const int q = warp & 3, grp = (warp - 2) >> 2;
    const int et = threadIdx.x - 64 - grp * EPI_GROUP_THREADS;  // 0..EPI_GROUP_THREADS-1 within the group
    const int esz = dtype_size(p.e.out_dtype);
    // 8-bit outputs without a residual only need 8-byte row alignment (the chunk store picks 16- or 8-byte vectors)
    constexpr bool AL8 = BN == 64 && mode_is_fast(MODE) && !mode_is_sum(MODE);
    constexpr int row_align = AL8 ? 8 : 16;
    const bool vec_ok = ((long long)p.N * esz) % row_align == 0 && (((uintptr_t)p.y) & 15) == 0 &&
                        (p.sum == nullptr || (((uintptr_t)p.sum) & 15) == 0) &&
                        (MODE == kModeRaw ? (p.e.out_dtype == kI32 && (p.e.bias == nullptr || (((uintptr_t)p.e.bias) & 15) == 0))
                                          : (p.e.out_dtype == kU8 || p.e.out_dtype == kI8));
    const uint32_t tab_base = smem_u32(tab_s) + grp * BN * 16;
    int tab_n0 = -1;
    // tile split: group g takes the CTA's tiles g, g + EG, ...: exactly those whose accumulator is TMEM stage g;
    // column split: every group takes BN / EG columns (at least one 32-column chunk) of every tile
    constexpr int tstep = SPLIT ? EG : 1, cspan = SPLIT ? BN : (BN / EG < 32 ? 32 : BN / EG);
    const int cbeg0 = SPLIT ? 0 : grp * cspan;
    uint32_t lt = SPLIT ? grp : 0;
    for (int tile = blockIdx.x + (int)lt * gridDim.x; tile < total_tiles; tile += tstep * gridDim.x, lt += tstep) {
      const int nt = tile % ntiles;
      const int t2 = tile / ntiles;
      const int mt = t2 % mtiles, b = t2 / mtiles;
      const int n0 = nt * BN, m0 = mt * BM;
      if (mode_is_fast(MODE) && n0 != tab_n0) {
        // stage this n-tile's per-channel constants (the condition is uniform over the group's warps)
        epi_group_sync(grp);
        for (int i = et; i < BN; i += EPI_GROUP_THREADS)
          sts128(tab_base + i * 16, n0 + i < p.N ? __ldg(p.f.tabA + n0 + i) : make_int4(0, 0, 0, 0));
        epi_group_sync(grp);
        tab_n0 = n0;
      }
      int row = m0 + q * 32 + lane;
      bool row_valid = row < p.M;
      if (p.win) {
        const int qp = (mt % p.win_qt) * BM + q * 32 + lane;
        row_valid = qp < p.OW;
        row = (mt / p.win_qt) * p.OW + qp;
      }
      epi_drain_tile<BN, MODE, AL8, TcKernelParams, false, AS, EG>(p, tmem_base, lt, q, cbeg0, cspan, lane, row, row_valid, b, n0, tab_base, vec_ok,
                                                               tempty, tfull);
      TKB_TL_IF(lt == 0 && warp == 2 && lane == 0, p, 7);
    }
    TKB_TL_IF(warp == 2 && lane == 0, p, 8);
  }
  tc_fence_before();
  __syncthreads();
  TKB_TL_IF(threadIdx.x == 0, p, 9);
  if (warp == 1) tmem_dealloc<AS * BN>(tmem_base);
}

inline int num_sms() {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
    // experiment knob: persistent grids of fewer CTAs leave SMs to whatever runs beside them (trace transposes)
    if (const char* e = getenv("TKB_SMS")) { const int v = atoi(e); if (v > 0 && v < sms) sms = v; }
  }
  return sms;
}

// SPLIT (epilogue partition, template parameter of the kernels): true = the two epilogue groups take alternate
// TILES, all columns each (64-column tiles of CTAs that have several tiles: two tiles are in the epilogue at once);
// false = both groups work on every tile, half of its columns each.
template <int BN, int MODE, bool SPLIT, int EG = 2>
static int launch_ts(const CUtensorMap& ta, const CUtensorMap& tb, TcKernelParams& p, cudaStream_t s) {
  const int sub_bytes = (BM + BN) * p.bk;
  const int fixed = EG * BN * 16 + (2 * MAX_STAGES + 8) * 8 + 16 + 1024 /* alignment slack */;
  const int budget = (EG == 1 ? 110 : 227) * 1024 - fixed;  // EG == 1: two CTAs share the SM's shared memory
  const int num_kb = p.implicit ? p.KH * p.KW * (p.C / p.bk) : (p.K + p.bk - 1) / p.bk;
  // Group k-blocks so that one mbarrier round trip (~100+ cycles on the single issuing thread for each of
  // wait / elect / commit) covers enough MMAs: stages of up to 64 KB, at least 3 of them.
  int kps = (64 * 1024) / sub_bytes;
  kps = kps < 1 ? 1 : (kps > 8 ? 8 : kps);
  if (kps > num_kb) kps = num_kb;
  while (kps > 1 && budget / (kps * sub_bytes) < 3) --kps;
  p.kps = kps;
  const int stage_bytes = kps * sub_bytes;
  int stages = budget / stage_bytes;
  stages = stages > MAX_STAGES ? MAX_STAGES : stages;
  if (stages < 2) return -1;
  p.stages = stages;
  const int smem = stages * stage_bytes + fixed;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(tc_gemm_kernel<BN, MODE, SPLIT, EG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) !=
        cudaSuccess)
      return -1;
    attr_set = true;
  }
  const long long mt_host = p.win ? (long long)p.win_rows * p.win_qt : (p.M + BM - 1) / BM;
  const long long tiles = mt_host * ((p.N + BN - 1) / BN) * p.batches;
  if (tiles > 0x3fffffffLL) return -1;  // the kernel decodes tiles in 32-bit arithmetic
  const int ctas = (EG == 1 ? 2 : 1) * num_sms();
  const int grid = (int)(tiles < ctas ? tiles : ctas);
  p.idesc = idesc_i8(BM, BN, false, false) | (p.idesc & ((1u << 7) | (1u << 10)));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(tc_threads(EG));
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = getenv("TKB_NO_PDL") ? 0 : 1;
  p.tl = timeline_buf();
  if (cudaLaunchKernelEx(&cfg, tc_gemm_kernel<BN, MODE, SPLIT, EG>, ta, tb, p) != cudaSuccess) return -1;
  timeline_print("tc_gemm", p.tl, s);
  ++g_launches;
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// experiment knob TKB_EG: 1 = two small CTAs per SM wherever instantiated, 2 = never; default: by tiles and MMAs per tile
// (measured, MobileNetV2 batch 128: 16 -> 96 at 112x112 112.8 -> 93.4 us, 32 -> 16 62.5 -> 52.2 us, 96 -> 24 at 56x56
// 23.8 -> 20.3 us; layers with two column tiles or ~10 tiles per SM get slower, so they keep the one-CTA form)
inline int epi_groups_for(long long tiles_per_sm, int num_mma_per_tile, int ntiles) {
  static const int forced = [] { const char* e = getenv("TKB_EG"); return e ? atoi(e) : 0; }();
  if (forced == 1 || forced == 2) return forced;
  static const int max_mma = [] { const char* e = getenv("TKB_EG_MMA"); return e ? atoi(e) : 4; }();      // tuning knobs
  static const int min_tiles = [] { const char* e = getenv("TKB_EG_TILES"); return e ? atoi(e) : 16; }();
  return (tiles_per_sm >= min_tiles && num_mma_per_tile <= max_mma && ntiles == 1) ? 1 : 2;
}

template <int BN, int MODE>
static int launch_t(const CUtensorMap& ta, const CUtensorMap& tb, TcKernelParams& p, cudaStream_t s) {
  const long long mt_host = p.win ? (long long)p.win_rows * p.win_qt : (p.M + BM - 1) / BM;
  const long long tiles = mt_host * ((p.N + BN - 1) / BN) * p.batches;
  if constexpr (mode_is_fast(MODE) && BN <= 128) {
    const int num_kb = p.implicit ? p.KH * p.KW * (p.C / p.bk) : (p.K + p.bk - 1) / p.bk;
    // (the saturating-pack variants spill in this configuration: explicit clamps, same results)
    if (epi_groups_for(tiles / num_sms(), num_kb * (p.bk / UMMA_K), (p.N + BN - 1) / BN) == 1) return launch_ts<BN, (MODE & 3), false, 1>(ta, tb, p, s);
  }
  if constexpr (BN == 64 && mode_is_fast(MODE)) {
    if (tiles >= 2LL * num_sms()) return launch_ts<BN, MODE, true>(ta, tb, p, s);
  }
  return launch_ts<BN, MODE, false>(ta, tb, p, s);
}

// generic / raw epilogue instantiations live in their own translation unit (compile time)
int launch_tc_other(int bn, int mode, const CUtensorMap& ta, const CUtensorMap& tb, TcKernelParams& p, cudaStream_t s);

}  // namespace tkb
