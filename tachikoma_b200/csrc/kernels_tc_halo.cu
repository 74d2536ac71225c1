// Halo-reuse tcgen05 convolution for stride-1, undilated KHxKW kernels (the 3x3 body layers of ResNet-class
// networks; arithmetic: qnn.conv2d, src/relay/qnn/op/convolution.cc:677-770 of the reference).
//
// The implicit-GEMM kernel (tc_kernel.cuh) fetches every activation pixel once per filter tap: KH*KW times the
// tensor's bytes go through L2 -> shared memory, and for 64/128-channel layers that traffic, not the tensor
// pipe, bounds the kernel.  Here a tile is hR whole output rows of ONE image and the A operand is loaded ONCE
// per channel chunk as a (hR+KH-1) x (OW+KW-1) pixel patch (one tiled 4-D TMA box; rows / columns outside the
// image are zero-filled by the TMA unit and corrected by the border-class deltas, as in the im2col path).
// The patch is a dense pixel-major array with a pitch of bk bytes per pixel, i.e. exactly a K-major UMMA
// operand whose "rows" are patch pixels.  GEMM row m is patch pixel m (row m / hWp, column m % hWp), so the
// operand of filter tap (r, s) is the SAME array started (r*hWp + s) pixels later: one descriptor start
// address per tap, no data movement.  Rows whose column falls into the right-hand halo (m % hWp >= OW) are
// junk and are simply not stored.  The whole weight matrix stays resident in shared memory (bulk-copied once
// per CTA from the pre-swizzled image tc_pack_weights() makes), so the steady state streams nothing but
// patches: (hR+KH-1)/hR * hWp/OW bytes per output pixel-channel instead of KH*KW.
#include "tc_kernel.cuh"

namespace tkb {

static constexpr int MAX_A_STAGES = 8;

// KH_ = 0: any KH x KW / chunk size, loops with run-time bounds.  Otherwise KH_ x KW_ taps with BK_ bytes per pixel
// chunk, fully unrolled -- a straight line of UTCIMMAs with one uniform add per descriptor.  (The run-time loop
// costs ~110 cycles per MMA on the single issuing thread -- LDCU / R2UR / UIMAD chains between the MMAs -- which
// for N = 64 tiles is 3x the tensor pipe's own time per instruction; unrolled it is ~57.)
template <int BN, int MODE, int KH_, int KW_, int BK_, bool SPLIT, bool CTAB = false, int EG = 2>
__global__ void TKB_TC_BOUNDS(EG) tc_halo_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ TcKernelParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int cchunks = p.C / p.bk, taps = p.KH * p.KW, num_kb = taps * cchunks;
  const int b_bytes = BN * p.bk;
  uint8_t* smem_b = smem;                                // resident weights: num_kb blocks of BN x bk
  uint8_t* smem_a = smem + (size_t)num_kb * b_bytes;     // patch ring: a_stages slots of a_slot bytes
  int4* tab_s = (int4*)(smem_a + (size_t)p.a_stages * p.a_slot);
  uint64_t* a_full = (uint64_t*)(tab_s + BN);
  uint64_t* a_empty = a_full + MAX_A_STAGES;
  uint64_t* tfull = a_empty + MAX_A_STAGES;
  uint64_t* tempty = tfull + 4;
  uint64_t* b_full = tempty + 4;
  uint32_t* tmem_ptr = (uint32_t*)(b_full + 1);
  constexpr int AS = SPLIT ? EG : 2;  // accumulator stages (tc_kernel.cuh)
  static_assert(AS * BN <= 512, "TMEM columns");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nimg = p.M / (p.OH * p.OW);
  const int total_tiles = (p.dbg & 8) ? 0 : nimg * p.hTpi;  // host: < 2^30  // dbg 8: launch + prologue + teardown only

  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  TKB_TL_IF(threadIdx.x == 0, p, 0);
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    for (int i = 0; i < p.a_stages; ++i) { mbar_init(&a_full[i], 1); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < AS; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], SPLIT ? EPI_GROUP_WARPS : EG * EPI_GROUP_WARPS); }
    mbar_init(b_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<AS * BN>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  TKB_TL_IF(threadIdx.x == 0, p, 1);

  if (warp == 0) {
    // ===== producer: weights once (constants: they do not depend on the preceding kernel), then patches =====
    if (elect_one()) {
      mbar_expect_tx(b_full, (uint32_t)(num_kb * b_bytes));
      for (int kb = 0; kb < num_kb; ++kb)
        bulk_load(smem_b + (size_t)kb * b_bytes, p.wpack + (size_t)kb * b_bytes, (uint32_t)b_bytes, b_full);
    }
    __syncwarp();
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const uint32_t patch_bytes = (uint32_t)(p.hRows * p.hWp * p.bk);
    int st = 0;
    uint32_t ph = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int n = tile / p.hTpi, oh0 = (tile % p.hTpi) * p.hR;
      for (int cc = 0; cc < cchunks; ++cc) {
        mbar_wait(&a_empty[st], ph ^ 1);
        if (elect_one()) {
          if (p.dbg & 4) mbar_arrive(&a_full[st]);  // timing experiment: no patch traffic
          else {
            mbar_expect_tx(&a_full[st], patch_bytes);
            tma_load_4d(smem_a + (size_t)st * p.a_slot, &tmA, &a_full[st], cc * p.bk, -p.pl, oh0 - p.pt, n);
          }
        }
        __syncwarp();
        if (++st == p.a_stages) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    asm volatile("griddepcontrol.wait;" ::: "memory");
    TKB_TL_IF(lane == 0, p, 2);
    int st = 0;
    uint32_t ph = 0, lt = 0;
    const int kper = p.bk / UMMA_K;
    const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
    mbar_wait(b_full, 0);
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const uint32_t as = lt % AS;
      mbar_wait(&tempty[as], ((lt / AS) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      for (int cc = 0; cc < cchunks; ++cc) {
        mbar_wait(&a_full[st], ph);
        tc_fence_after();
        TKB_TL_IF(lt == 0 && cc == 0 && lane == 0, p, 3);
        if (elect_one()) {
          const uint32_t slot = a_base + st * p.a_slot;
          uint32_t acc = cc != 0;
          // The swizzle XOR is applied to absolute shared-memory address bits, so a start address that is not a
          // multiple of the 8-row pattern needs nothing else: measured on B200, the descriptor's "matrix base
          // offset" field must stay 0 here (setting it to (addr >> 7) & 7 gives wrong sums).
          const uint64_t hi = (uint64_t)p.desc_hi << 32;
          if constexpr (KH_ != 0) {
            const uint32_t a_lo0 = ((slot >> 4) & 0x3FFF) | (1u << 16);
            const uint32_t b_lo0 = (((b_base + (uint32_t)(cc * b_bytes)) >> 4) & 0x3FFF) | (1u << 16);
            const uint32_t b_step = (uint32_t)(cchunks * b_bytes) >> 4, row_step = (uint32_t)(p.hWp * BK_) >> 4;
            if (!(p.dbg & 32)) {
#pragma unroll
              for (int r = 0; r < KH_; ++r)
#pragma unroll
                for (int s = 0; s < KW_; ++s)
#pragma unroll
                  for (int k = 0; k < BK_ / UMMA_K; ++k) {
                    mma_i8(d_tmem, hi | (a_lo0 + r * row_step + s * (BK_ / 16) + 2 * k),
                           hi | (b_lo0 + (r * KW_ + s) * b_step + 2 * k), p.idesc, acc);
                    acc = 1;
                  }
            }
          } else {
            int t = 0;
            for (int r = 0; r < ((p.dbg & 32) ? 0 : p.KH); ++r)  // dbg 32: timing experiment without MMAs
              for (int s = 0; s < p.KW; ++s, ++t) {
                const uint32_t a_addr = slot + (uint32_t)((r * p.hWp + s) * p.bk);
                const uint32_t b_addr = b_base + (uint32_t)((t * cchunks + cc) * b_bytes);
                const uint32_t a_lo = ((a_addr >> 4) & 0x3FFF) | (1u << 16);
                const uint32_t b_lo = ((b_addr >> 4) & 0x3FFF) | (1u << 16);
                for (int k = 0; k < kper; ++k) {
                  mma_i8(d_tmem, hi | (a_lo + 2 * k), hi | (b_lo + 2 * k), p.idesc, acc);
                  acc = 1;
                }
              }
          }
          tc_commit(&a_empty[st]);
          if (cc == cchunks - 1) tc_commit(&tfull[as]);
          TKB_TL_IF(lt < 48 && cc == cchunks - 1, p, 16 + lt);  // per TILE here
        }
        __syncwarp();
        if (++st == p.a_stages) { st = 0; ph ^= 1; }
      }
      TKB_TL_IF(lt == 0 && lane == 0, p, 4);
    }
    TKB_TL_IF(lane == 0, p, 5);
  } else {
    // ===== epilogue =====
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int q = warp & 3, grp = (warp - 2) >> 2;  // group g drains the tiles of TMEM stage g (see tc_kernel.cuh)
    const int et = threadIdx.x - 64;
    const bool vec_ok = (p.N % 16) == 0 && (((uintptr_t)p.y) & 15) == 0 &&
                        (p.sum == nullptr || (((uintptr_t)p.sum) & 15) == 0) &&
                        (p.e.out_dtype == kU8 || p.e.out_dtype == kI8);
    const uint32_t tab_base = smem_u32(tab_s);
    if constexpr (!CTAB) {  // CTAB: the constants are kernel parameters (p.ctab), nothing to stage
      for (int i = et; i < BN; i += EG * EPI_GROUP_THREADS)
        sts128(tab_base + i * 16, i < p.N ? __ldg(p.f.tabA + i) : make_int4(0, 0, 0, 0));
      epi_bar_sync<EG>();
    }
    const int m = q * 32 + lane;
    const int prow = m / p.hWp, pcol = m - prow * p.hWp;
    const bool slot_valid = prow < p.hR && pcol < p.OW;
    constexpr int tstep = SPLIT ? EG : 1, cspan = SPLIT ? BN : (BN / EG < 32 ? 32 : BN / EG);
    const int cbeg0 = SPLIT ? 0 : grp * cspan;
    uint32_t lt = SPLIT ? grp : 0;
    for (int tile = blockIdx.x + (int)lt * gridDim.x; tile < total_tiles; tile += tstep * gridDim.x, lt += tstep) {
      const int n = tile / p.hTpi, oh = (tile % p.hTpi) * p.hR + prow;
      const bool row_valid = slot_valid && oh < p.OH;
      const int row = (n * p.OH + oh) * p.OW + pcol;
      epi_drain_tile<BN, MODE, false, TcKernelParams, CTAB, AS, EG>(p, tmem_base, lt, q, cbeg0, cspan, lane, row, row_valid, 0, 0, tab_base, vec_ok,
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

// ---- host side -------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess ||
        qr != cudaDriverEntryPointSuccess)
      f = nullptr;
    return (EncodeTiledFn)f;
  }();
  return fn;
}

struct HaloCfg { int bn, bk, R, Wp, rows, tpi, a_slot, a_stages, smem; };

static bool halo_config(const ConvGeom& g, int N, HaloCfg* c) {
  if (g.groups != 1 || g.sh != 1 || g.sw != 1 || g.dh != 1 || g.dw != 1) return false;
  if (g.C % 64 != 0 || g.KH * g.KW < 2 || g.KH > 7 || g.KW > 7) return false;
  if (N % 16 != 0 || N > 256) return false;
  c->bn = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
  c->bk = g.C % 128 == 0 ? 128 : 64;
  c->Wp = g.OW + g.KW - 1;
  if (c->Wp != g.W + g.pl + g.pr || g.OH + g.KH - 1 != g.H + g.pt + g.pb) return false;
  if (c->Wp > BM || c->Wp > 256) return false;
  c->R = BM / c->Wp;
  if (c->R > g.OH) c->R = g.OH;
  c->rows = c->R + g.KH - 1;
  if (c->rows > 256) return false;
  c->tpi = (g.OH + c->R - 1) / c->R;
  if ((long long)c->R * g.OW * 10 < (long long)BM * 6) return false;  // more than 40 % junk rows: not worth it
  // a slot must hold the patch and everything a 128-row operand started at the last tap can touch
  const int px = std::max(c->rows * c->Wp, BM + (g.KH - 1) * c->Wp + g.KW - 1);
  c->a_slot = (px * c->bk + 1023) / 1024 * 1024;
  const int num_kb = g.KH * g.KW * (g.C / c->bk);
  const long long b_total = (long long)num_kb * c->bn * c->bk;
  const int fixed = c->bn * 16 + (2 * MAX_A_STAGES + 9) * 8 + 16 + 1024;
  const long long room = 227 * 1024 - fixed - b_total;
  if (room < 2LL * c->a_slot) return false;
  c->a_stages = (int)std::min<long long>(MAX_A_STAGES, room / c->a_slot);
  c->smem = (int)(b_total + (long long)c->a_stages * c->a_slot + fixed);
  return true;
}

bool tc_halo_supported(const ConvGeom& g, int N) {
  HaloCfg c;
  return halo_config(g, N, &c);
}

void tc_halo_tile_config(const ConvGeom& g, int N, int* bn, int* bk) {
  HaloCfg c;
  if (halo_config(g, N, &c)) { *bn = c.bn; *bk = c.bk; }
}

template <int BN, int MODE, int KH_, int KW_, int BK_, bool SPLIT, bool CTAB = false, int EG = 2>
static int launch_halo_ts(const CUtensorMap& ta, TcKernelParams& p, int smem, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(tc_halo_kernel<BN, MODE, KH_, KW_, BK_, SPLIT, CTAB, EG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) !=
        cudaSuccess)
      return -1;
    attr_set = true;
  }
  const long long tiles = (long long)(p.M / (p.OH * p.OW)) * p.hTpi;
  if (tiles > 0x3fffffffLL) return -1;  // the kernel decodes tiles in 32-bit arithmetic
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(tiles < num_sms() ? tiles : num_sms()));
  cfg.blockDim = dim3(tc_threads(EG));
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = getenv("TKB_NO_PDL") ? 0 : 1;
  p.idesc = idesc_i8(BM, BN, false, false) | (p.idesc & ((1u << 7) | (1u << 10)));
  p.tl = timeline_buf();
  if (cudaLaunchKernelEx(&cfg, tc_halo_kernel<BN, MODE, KH_, KW_, BK_, SPLIT, CTAB, EG>, ta, p) != cudaSuccess) return -1;
  timeline_print("tc_halo", p.tl, s);
  ++g_launches;
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// epilogue partition (see launch_t in tc_kernel.cuh): tile split for 64-column tiles of multi-tile CTAs
template <int BN, int MODE, int KH_, int KW_, int BK_>
static int launch_halo_t(const CUtensorMap& ta, TcKernelParams& p, int smem, cudaStream_t s) {
  if constexpr (BN == 64) {
    const long long tiles = (long long)(p.M / (p.OH * p.OW)) * p.hTpi;
    if (tiles >= 2LL * num_sms()) {
      // exactly 64 output channels: per-channel constants through the parameter block (profiles/r02_notes.md section 7:
      // the table's broadcast LDS.128 were half of the shared-memory data pipe's load in these kernels)
      static const bool ctab_off = getenv("TKB_NO_CTAB") != nullptr;
      if (p.N == 64 && p.f.tab_host != nullptr && !ctab_off) {
        memcpy(p.ctab, p.f.tab_host, sizeof(p.ctab));
        return launch_halo_ts<BN, MODE, KH_, KW_, BK_, true, true>(ta, p, smem, s);
      }
      return launch_halo_ts<BN, MODE, KH_, KW_, BK_, true>(ta, p, smem, s);
    }
  }
  return launch_halo_ts<BN, MODE, KH_, KW_, BK_, false>(ta, p, smem, s);
}

int launch_tc_halo(const TcArgs& a, cudaStream_t s) {
  const ConvGeom& g = a.g;
  HaloCfg c;
  if (!halo_config(g, a.N, &c) || !a.f.enabled || !a.w_packed || a.pack_bn != c.bn || a.pack_bk != c.bk) return -1;
  if ((((uintptr_t)a.a) | ((uintptr_t)a.w)) & 15) return -1;
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return -1;
  TcKernelParams p;
  memset(&p, 0, sizeof(p));
  p.y = a.y;
  p.sum = a.e.sum;
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.batches = 1;
  p.e = a.e;
  p.f = a.f;
  if (const char* d = getenv("TKB_DBG")) p.dbg = atoi(d);
  p.delta = a.delta; p.cls_h = a.cls_h; p.cls_w = a.cls_w; p.ncls_w = a.ncls_w;
  p.C = g.C; p.KH = g.KH; p.KW = g.KW; p.OH = g.OH; p.OW = g.OW; p.pt = g.pt; p.pl = g.pl;
  p.bk = c.bk;
  p.hR = c.R; p.hWp = c.Wp; p.hRows = c.rows; p.hTpi = c.tpi; p.a_slot = c.a_slot; p.a_stages = c.a_stages;
  p.wpack = (const uint8_t*)a.w;
  const uint32_t layout = c.bk == 128 ? 2u : 4u;
  p.desc_hi = (uint32_t)(smem_desc(0, 16, 8 * c.bk, layout) >> 32);
  p.idesc = ((a.a_dtype == kI8 ? 1u : 0u) << 7) | ((a.w_dtype == kI8 ? 1u : 0u) << 10);
  // NHWC activation tensor as {C, W, H, N}; box = one channel chunk of a (rows x Wp) pixel patch
  CUtensorMap ta;
  cuuint64_t dims[4] = {(cuuint64_t)g.C, (cuuint64_t)g.W, (cuuint64_t)g.H, (cuuint64_t)g.N};
  cuuint64_t strides[3] = {(cuuint64_t)g.C, (cuuint64_t)g.C * g.W, (cuuint64_t)g.C * g.W * g.H};
  cuuint32_t box[4] = {(cuuint32_t)c.bk, (cuuint32_t)c.Wp, (cuuint32_t)c.rows, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  if (fn(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(a.a), dims, strides, box, estr,
         CU_TENSOR_MAP_INTERLEAVE_NONE, c.bk == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return -1;
  const bool sum = a.f.has_sum != 0;
  // saturating-pack epilogue (kModeFastSat): spills at BN = 128 without the residual, so 64-column tiles only there
  const bool sat = a.f.sat && c.bn == 64, sum_sat = a.f.sum_sat != 0;
  const bool k33 = g.KH == 3 && g.KW == 3, k14 = g.KH == 1 && g.KW == 4, k12 = g.KH == 1 && g.KW == 2;
#define TKB_HALO(COND_, BN_, KH_, KW_, BK_)                                                        \
  if ((COND_) && c.bn == BN_ && (BK_ == 0 || c.bk == BK_))                                         \
    return sum ? (sum_sat ? launch_halo_t<BN_, kModeFastSumSat, KH_, KW_, BK_>(ta, p, c.smem, s)  \
                              : launch_halo_t<BN_, kModeFastSum, KH_, KW_, BK_>(ta, p, c.smem, s))     \
               : (sat ? launch_halo_t<BN_, kModeFastSat, KH_, KW_, BK_>(ta, p, c.smem, s)          \
                          : launch_halo_t<BN_, kModeFast, KH_, KW_, BK_>(ta, p, c.smem, s));
  TKB_HALO(k33, 64, 3, 3, 64) TKB_HALO(k33, 64, 3, 3, 128) TKB_HALO(k33, 128, 3, 3, 64) TKB_HALO(k33, 128, 3, 3, 128)
  TKB_HALO(k14, 64, 1, 4, 64)  // the stride-2 7x7 stem as a 1x4 convolution over pixel pairs of the unfolded input
  TKB_HALO(k12, 64, 1, 2, 64)  // a stride-2 3x3 stem (MobileNetV2) in the same form: 1x2
  TKB_HALO(true, 64, 0, 0, 0) TKB_HALO(true, 128, 0, 0, 0) TKB_HALO(true, 256, 0, 0, 0)
#undef TKB_HALO
  return -1;
}

}  // namespace tkb
