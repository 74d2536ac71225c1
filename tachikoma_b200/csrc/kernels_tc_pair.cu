// cta_group::2 form of the persistent tcgen05 GEMM / implicit-GEMM convolution (tc_kernel.cuh): a cluster of TWO CTAs on
// the two SMs of a TPC computes one 256 x BN output tile.
//
// Why: a kind::i8 MMA reads (128 + N) x 32 operand bytes from shared memory per instruction; measured cycles per MMA of
// the single-CTA kernel (108 at N = 128, 168 at N = 256) track that operand rate, not the tensor pipe (N / 2 cycles).
// With cta_group::2 the MMA has M = 256: each CTA supplies its own 128 rows of A but only HALF of B (N / 2 rows), so a
// CTA's shared memory delivers (128 + N / 2) x 32 bytes per instruction: 8 KB instead of 12 KB at N = 256.
//
// Protocol (per CTA pair; rank 0 = leader):
//   * both CTAs' producer warps load their A half (128 rows; tiled or im2col-mode tensor map) and their B half (BN / 2
//     weight rows) into their OWN shared memory with cp.async.bulk.tensor...cta_group::2, signalling the LEADER's "full"
//     barrier; the leader's producer posts the expected bytes of both CTAs;
//   * the leader's MMA warp issues tcgen05.mma.cta_group::2 (UTCIMMA.2CTA), which reads both CTAs' shared memory at the
//     same offsets and writes both CTAs' TMEM (lanes = the CTA's 128 rows, columns = BN);
//   * tcgen05.commit.cta_group::2...multicast::cluster arrives on the "empty" (slot free) and "accumulator full" barriers
//     of BOTH CTAs; each CTA's eight epilogue warps drain their own TMEM with the shared epilogue code and release the
//     accumulator stage on the leader's barrier (16 arrivals per stage, acc_release<P>).
// Arithmetic and epilogue are those of the single-CTA kernel (reference: src/relay/qnn/op/convolution.cc:708-811,
// requantize.cc:195-273, src/target/intrin_rule.cc:165-195, add.cc:40-93).
#include "tc_kernel.cuh"

namespace tkb {

struct PairParams : TcKernelParams {
  static constexpr bool kPair = true;
};

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t leader_addr(const void* p) {  // cluster address of the same object in CTA rank 0
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(ra) : "r"(smem_u32(p)));
  return ra;
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait_cluster(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait_cluster(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {
      printf("tkb: pair mbarrier timeout block=%d thread=%d parity=%u\n", blockIdx.x, threadIdx.x, parity);
      __trap();
    }
  }
}
__device__ __forceinline__ void tma2_load_3d(void* dst, const CUtensorMap* m, uint32_t bar_cluster, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma2_load_im2col_4d(void* dst, const CUtensorMap* m, uint32_t bar_cluster, int c, int w, int h,
                                                    int n, uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w),
      "h"(off_h)
      : "memory");
}
__device__ __forceinline__ void mma2_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_commit2(uint64_t* bar) {  // arrives on `bar` of both CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)), "h"((uint16_t)3)
               : "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_alloc2(uint32_t* dst_smem) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "n"(COLS) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}

static constexpr int PM = 2 * BM;  // rows of a pair tile

template <int BN, int MODE>
__global__ void __launch_bounds__(TC_THREADS, 1) tc_pair_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ CUtensorMap tmB,
                                                                const __grid_constant__ PairParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  constexpr int BH = BN / 2;  // weight rows held by this CTA
  const int a_bytes = BM * p.bk, b_bytes = BH * p.bk, sub_bytes = a_bytes + b_bytes, stage_bytes = p.kps * sub_bytes;
  int4* tab_s = (int4*)(smem + p.stages * stage_bytes);
  uint64_t* full = (uint64_t*)(tab_s + 2 * BN);
  uint64_t* empty = full + MAX_STAGES;
  uint64_t* tfull = empty + MAX_STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_ptr = (uint32_t*)(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_rank();
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int mtiles = (p.M + PM - 1) / PM, ntiles = (p.N + BN - 1) / BN;
  const int total_tiles = (p.dbg & 8) ? 0 : mtiles * ntiles * p.batches;
  const int cchunks = p.implicit ? p.C / p.bk : 0;
  const int num_kb = p.implicit ? p.KH * p.KW * cchunks : (p.K + p.bk - 1) / p.bk;

  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < p.stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 2 * EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc2<2 * BN>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's barriers and TMEM exist before anything is sent to them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp == 0) {
    // ===== TMA producer (both CTAs): own A rows, own half of the weight rows, completion on the leader's barrier =====
    int st = 0;
    uint32_t ph = 0;
    const uint32_t tx_bytes = (uint32_t)(a_bytes + b_bytes);
    for (int tile = pair; tile < total_tiles; tile += npairs) {
      const int nt = tile % ntiles;
      const int t2 = tile / ntiles;
      const int mt = t2 % mtiles, b = t2 / mtiles;
      const int n0 = nt * BN + (int)rank * BH, m0 = mt * PM + (int)rank * BM;
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
          if (rank == 0) mbar_expect_tx(&full[st], 2u * (uint32_t)nsub * tx_bytes);  // both CTAs' bytes
          const uint32_t lbar = leader_addr(&full[st]);
          for (int j = 0; j < nsub; ++j) {
            uint8_t* sa = smem + st * stage_bytes + j * sub_bytes;
            const int kb = kb0 + j;
            if (p.implicit) {
              tma2_load_im2col_4d(sa, &tmA, lbar, cc * p.bk, cw, chh, cn, (uint16_t)(s * p.dw), (uint16_t)(r * p.dh));
              if (++cc == cchunks) { cc = 0; if (++s == p.KW) { s = 0; ++r; } }
            } else {
              tma2_load_3d(sa, &tmA, lbar, kb * p.bk, m0, b);
            }
            tma2_load_3d(sa + a_bytes, &tmB, lbar, kb * p.bk, n0, p.w_batched ? b : 0);
          }
        }
        r = __shfl_sync(0xffffffffu, r, 0); s = __shfl_sync(0xffffffffu, s, 0); cc = __shfl_sync(0xffffffffu, cc, 0);
        if (++st == p.stages) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the leader CTA only =====
    if (rank == 0) {
      int st = 0;
      uint32_t ph = 0, lt = 0;
      const int kper = p.bk / UMMA_K;
      const uint32_t smem_base = smem_u32(smem);
      for (int tile = pair; tile < total_tiles; tile += npairs, ++lt) {
        const uint32_t as = lt & 1;
        mbar_wait_cluster(&tempty[as], ((lt >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb0 = 0; kb0 < num_kb; kb0 += p.kps) {
          const int nsub = min(p.kps, num_kb - kb0);
          mbar_wait(&full[st], ph);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t hi = (uint64_t)p.desc_hi << 32;
            uint32_t acc = kb0 != 0;
            for (int j = 0; j < nsub; ++j) {
              const uint32_t a_addr = smem_base + st * stage_bytes + j * sub_bytes;
              const uint32_t a_lo = ((a_addr >> 4) & 0x3FFF) | (1u << 16);
              const uint32_t b_lo = (((a_addr + a_bytes) >> 4) & 0x3FFF) | (1u << 16);
              for (int k = 0; k < kper; ++k) {
                mma2_i8(d_tmem, hi | (a_lo + 2 * k), hi | (b_lo + 2 * k), p.idesc, acc);
                acc = 1;
              }
            }
            tc_commit2(&empty[st]);
            if (kb0 + nsub >= num_kb) tc_commit2(&tfull[as]);
          }
          __syncwarp();
          if (++st == p.stages) { st = 0; ph ^= 1; }
        }
      }
    }
  } else {
    // ===== epilogue (both CTAs): this CTA's 128 rows of the pair tile =====
    const int q = warp & 3, grp = (warp - 2) >> 2;
    const int et = threadIdx.x - 64 - grp * EPI_GROUP_THREADS;
    const int esz = dtype_size(p.e.out_dtype);
    const bool vec_ok = ((long long)p.N * esz) % 16 == 0 && (((uintptr_t)p.y) & 15) == 0 &&
                        (p.sum == nullptr || (((uintptr_t)p.sum) & 15) == 0) && (p.e.out_dtype == kU8 || p.e.out_dtype == kI8);
    const uint32_t tab_base = smem_u32(tab_s) + grp * BN * 16;
    int tab_n0 = -1;
    const int cbeg0 = grp * (BN / 2);
    uint32_t lt = 0;
    for (int tile = pair; tile < total_tiles; tile += npairs, ++lt) {
      const int nt = tile % ntiles;
      const int t2 = tile / ntiles;
      const int mt = t2 % mtiles, b = t2 / mtiles;
      const int n0 = nt * BN, m0 = mt * PM + (int)rank * BM;
      if (n0 != tab_n0) {
        epi_group_sync(grp);
        for (int i = et; i < BN; i += EPI_GROUP_THREADS)
          sts128(tab_base + i * 16, n0 + i < p.N ? __ldg(p.f.tabA + n0 + i) : make_int4(0, 0, 0, 0));
        epi_group_sync(grp);
        tab_n0 = n0;
      }
      const int row = m0 + q * 32 + lane;
      const bool row_valid = row < p.M;
      epi_drain_tile<BN, MODE, false, PairParams>(p, tmem_base, lt, q, cbeg0, BN / 2, lane, row, row_valid, b, n0, tab_base, vec_ok,
                                                  tempty, tfull);
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // no CTA leaves while the other may still signal it or read its shared memory
  if (warp == 1) tmem_dealloc2<2 * BN>(tmem_base);
}

// ---- host side -------------------------------------------------------------------------------------
bool make_tmap_im2col_pub(CUtensorMap* tm, const void* base, const ConvGeom& g, int bk);
bool make_tmap_3d_pub(CUtensorMap* tm, const void* base, long long k, long long rows, long long batch, int box_rows, int bk);

static int pair_bk(const TcArgs& a) {
  if (a.implicit == 1) return a.g.C % 128 == 0 ? 128 : (a.g.C % 64 == 0 ? 64 : 32);
  return a.K >= 128 ? 128 : (a.K > 32 ? 64 : 32);
}

// Pair tile width: 256 columns when that still gives every pair of SMs a tile, else 128.
int tc_pair_bn(const TcArgs& a) {
  if (const char* f = getenv("TKB_PAIR_BN")) { const int v = atoi(f); if (v == 128 || v == 256) return v; }
  const long long mt = (a.M + PM - 1) / PM * a.B;
  const int pairs = num_sms() / 2;
  if (a.N % 256 == 0 && mt * (a.N / 256) * 3 >= pairs * 2) return 256;
  return 128;
}

bool tc_pair_supported(const TcArgs& a) {
  if (!a.f.enabled || a.w_packed || a.w_batched) return false;
  if (a.implicit != 0 && a.implicit != 1) return false;
  if (a.N % 128 != 0 || a.K % 16 != 0 || a.M <= 0) return false;
  if (a.implicit == 1 && !tc_implicit_conv_supported(a.g)) return false;
  if (a.delta != nullptr) return true;  // border rows take the out-of-line path, as in the single-CTA kernel
  return true;
}

template <int BN, int MODE>
static int launch_pair_t(const CUtensorMap& ta, const CUtensorMap& tb, PairParams& p, cudaStream_t s) {
  const int sub_bytes = (BM + BN / 2) * p.bk;
  const int fixed = 2 * BN * 16 + (2 * MAX_STAGES + 4) * 8 + 16 + 1024;
  const int budget = 227 * 1024 - fixed;
  const int num_kb = p.implicit ? p.KH * p.KW * (p.C / p.bk) : (p.K + p.bk - 1) / p.bk;
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
    if (cudaFuncSetAttribute(tc_pair_kernel<BN, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) return -1;
    attr_set = true;
  }
  const long long tiles = (long long)((p.M + PM - 1) / PM) * ((p.N + BN - 1) / BN) * p.batches;
  if (tiles > 0x3fffffffLL) return -1;
  const int pairs = (int)std::min<long long>(tiles, num_sms() / 2);
  p.idesc = idesc_i8(PM, BN, false, false) | (p.idesc & ((1u << 7) | (1u << 10)));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(TC_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = getenv("TKB_NO_PDL") ? 1 : 2;
  if (cudaLaunchKernelEx(&cfg, tc_pair_kernel<BN, MODE>, ta, tb, p) != cudaSuccess) return -1;
  ++g_launches;
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int launch_tc_pair(const TcArgs& a, cudaStream_t s) {
  if (!tc_pair_supported(a) || ((((uintptr_t)a.a) | ((uintptr_t)a.w)) & 15)) return -1;
  PairParams p;
  memset((void*)&p, 0, sizeof(p));
  p.y = a.y;
  p.sum = a.e.sum;
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.batches = a.B; p.w_batched = 0;
  p.e = a.e;
  p.f = a.f;
  if (const char* d = getenv("TKB_DBG")) p.dbg = atoi(d);
  p.delta = a.delta; p.cls_h = a.cls_h; p.cls_w = a.cls_w; p.ncls_w = a.ncls_w;
  const int bk = pair_bk(a), bn = tc_pair_bn(a);
  if (a.implicit) {
    const ConvGeom& g = a.g;
    p.implicit = 1;
    p.C = g.C; p.KH = g.KH; p.KW = g.KW; p.OH = g.OH; p.OW = g.OW;
    p.sh = g.sh; p.sw = g.sw; p.dh = g.dh; p.dw = g.dw; p.pt = g.pt; p.pl = g.pl;
  }
  p.bk = bk;
  const uint32_t layout = bk == 128 ? 2u : (bk == 64 ? 4u : 6u);
  p.desc_hi = (uint32_t)(smem_desc(0, 16, 8 * bk, layout) >> 32);
  p.idesc = ((a.a_dtype == kI8 ? 1u : 0u) << 7) | ((a.w_dtype == kI8 ? 1u : 0u) << 10);
  CUtensorMap ta, tb;
  if (a.implicit) {
    if (!make_tmap_im2col_pub(&ta, a.a, a.g, bk)) return -1;
  } else {
    if (!make_tmap_3d_pub(&ta, a.a, a.K, a.M, a.B, BM, bk)) return -1;
  }
  if (!make_tmap_3d_pub(&tb, a.w, a.K, a.N, 1, bn / 2, bk)) return -1;
  const bool sum = a.f.has_sum != 0;
  if (bn == 256) return sum ? launch_pair_t<256, kModeFastSum>(ta, tb, p, s) : launch_pair_t<256, kModeFast>(ta, tb, p, s);
  return sum ? launch_pair_t<128, kModeFastSum>(ta, tb, p, s) : launch_pair_t<128, kModeFast>(ta, tb, p, s);
}

}  // namespace tkb
