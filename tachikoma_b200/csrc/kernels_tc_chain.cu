// Chained tcgen05 convolutions / GEMMs: ONE persistent launch walks the tiles of several consecutive layers.
//
// Why: a per-layer launch of the persistent GEMM kernel (tc_kernel.cuh) pays ~4.5 us of ramp / prologue / teardown,
// and the deep ResNet layers have fewer tiles than the chip has SMs, one tile per CTA, nothing overlapped (round 1:
// 21 launches, 351 us of tensor-core time that would take 52 us at the int8 peak).  Here the layers of a network stage
// form one work list of (layer, tile) items in topological order:
//   * items are CLAIMED dynamically (atomic counter) by the resident CTAs, so a CTA that finishes the last tiles of
//     layer k continues with the first tiles of layer k+1 without a kernel boundary, and the epilogue of one tile
//     always overlaps the MMAs of the next (two TMEM accumulator stages), also across layers;
//   * a tile of layer k+1 needs only layer k's rows of ITS images (plus the residual's): completion is tracked per
//     (layer, image) in global counters; the epilogue warps publish a finished tile with a gpu-scope release, the TMA
//     producer acquires the counters of the tile's images before it issues the tile's first activation load
//     (weights are constants and never wait);
//   * dynamic claiming keeps the scheme deadlock-free when not all CTAs are resident (a second stream running its
//     own chain): the lowest unfinished item is always owned by a running CTA.
// Every layer's output still goes to global memory (the per-operator trace needs each of them), which is also what
// makes the cross-CTA hand-over possible.  Arithmetic, tile shapes, operand staging and the folded epilogue are the
// single-layer kernel's (tc_kernel.cuh); reference arithmetic: src/relay/qnn/op/convolution.cc:708-811,
// requantize.cc:195-273, src/target/intrin_rule.cc:165-195, add.cc:40-93.
#include <algorithm>
#include <vector>

#include "tc_kernel.cuh"

namespace tkb {

static constexpr int QD = 8;           // depth of the claimed-item queue between the producer warp and the other roles
static constexpr int ITEM_SHIFT = 20;  // work-list entry = (layer << ITEM_SHIFT) | tile

// Per-layer parameters, in the kernel's parameter (constant) space.  Member names mirror TcKernelParams where the
// shared epilogue code (epi_drain_tile / epilogue_chunk) reads them.
struct ChainEpiStub {
  int out_dtype;
  const int32_t* bias;  // unused (fast epilogues fold the bias); present for the shared raw-mode code path
};
struct ChainSeg {
  void* y;
  const void* sum;
  const uint8_t* wpack;
  const int32_t* delta;
  const uint8_t* cls_h;
  const uint8_t* cls_w;
  int* flags;               // [nimg] arrivals of epilogue warps on this layer's images
  const int* depA_flags;    // producer of the A operand inside this launch (nullptr: an earlier kernel)
  const int* depA_tgt;      // [nimg] arrivals that mean "image complete"
  const int* depS_flags;    // producer of the residual inside this launch
  const int* depS_tgt;
  long long* tl;            // timeline stamps (always nullptr here; the shared epilogue code reads the member)
  int M, N, K;
  int implicit, bk, kps, num_kb, cchunks;
  int C, KH, KW, OH, OW, sh, sw, dh, dw, pt, pl;
  int mtiles, ntiles, item_begin, item_end, rows_per_img, ncls_w, dbg;
  uint32_t idesc, desc_hi;
  ChainEpiStub e;
  FastEpi f;
  static constexpr bool kChain = true;  // residuals may have been written by this launch: coherent loads
  static constexpr bool kPair = false;
};
struct ChainParams {
  alignas(64) CUtensorMap tm[TKB_CHAIN_MAX_SEG];   // A operand (im2col-mode or tiled)
  alignas(64) CUtensorMap tmY[TKB_CHAIN_MAX_SEG];  // output [M][N] bytes, box 128 x 128, 128-byte swizzle (TMA store)
  ChainSeg seg[TKB_CHAIN_MAX_SEG];
  int nseg, total_items;
  int* counter;      // next unclaimed position of the work list (zeroed with the flags before the launch)
  const int* items;  // work list: (layer << ITEM_SHIFT) | tile, in an order that is topological for the dependencies
};

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add(int* p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

// Debug statistics (TKB_CHAIN_STATS=1 prints them): [0] items that found a dependency incomplete, [1] cycles spent
// waiting for dependencies (summed over CTAs), [2] items, [3] cycles the producer waited for a free pipeline stage.
__device__ unsigned long long g_chain_stats[8];

// Bounded spin on a completion counter: a protocol bug must surface as a CUDA error, never as a hung GPU.
__device__ __forceinline__ void wait_flag(const int* flag, int target) {
  if (ld_acquire(flag) >= target) return;
  const long long t0 = clock64();
  while (ld_acquire(flag) < target) {
    __nanosleep(32);
    if (clock64() - t0 > 4000000000LL) {
      printf("tkb: chain dependency timeout block=%d thread=%d have=%d want=%d\n", blockIdx.x, threadIdx.x, ld_acquire(flag), target);
      __trap();
    }
  }
  atomicAdd(&g_chain_stats[0], 1ull);
  atomicAdd(&g_chain_stats[1], (unsigned long long)(clock64() - t0));
}

// Warp roles of the chained kernel: 0 = claim / dependency wait / TMA producer, 1 = MMA issuer, 2 = publisher,
// 3..18 = sixteen epilogue warps (4 column groups x 4 TMEM lane quadrants; <= 96 registers each, 4-5 resident warps per
// scheduler instead of 2.5: the epilogue is issue-latency bound).  The epilogue warps write a finished tile into a
// swizzled shared-memory staging buffer; the publisher stores it with ONE TMA store (full 128-byte lines instead of
// 16-byte pieces scattered over 128 rows), waits for that store's completion and only then bumps the completion
// counters of the tile's images.  No epilogue warp ever executes a gpu-scope fence: the first version of this kernel
// (generic stores + fence + release-add per warp) spent more time in MEMBAR / ERRBAR / L1 invalidation than in arithmetic.
static constexpr int CH_EPI_WARPS = 16;
static constexpr int CH_EPI_THREADS = CH_EPI_WARPS * 32;
static constexpr int CH_THREADS = 96 + CH_EPI_THREADS;

template <int BN>
struct ChainCfg {
  static constexpr int KS = 128;                       // bytes of K per pipeline stage (kps = KS / bk k-blocks)
  static constexpr int STAGE_BYTES = (BM + BN) * KS;
  static constexpr int OUT_BYTES = BM * BN;            // one staged output tile (8-bit)
  static constexpr int TAB_BYTES = BN * 16;            // per-channel epilogue constants of one n-tile (one queue slot)
  static constexpr int FIXED = 2 * OUT_BYTES + QD * TAB_BYTES + (2 * MAX_STAGES + 6 + 3 * QD) * 8 + QD * 4 + 16 + 1024;
  static constexpr int STAGES = (227 * 1024 - FIXED) / STAGE_BYTES > MAX_STAGES ? MAX_STAGES : (227 * 1024 - FIXED) / STAGE_BYTES;
  static constexpr int SMEM = STAGES * STAGE_BYTES + FIXED;
};

// All MMAs of one full pipeline stage, unrolled: kps = KS / BK k-blocks of BK / 32 instructions each.  (The run-time
// form of this loop costs ~110 cycles per MMA on the single issuing thread, see kernels_tc_halo.cu.)
template <int BN, int BK>
__device__ __forceinline__ void issue_stage(uint32_t d_tmem, uint32_t stage_addr, uint64_t hi, uint32_t idesc, uint32_t acc) {
  constexpr int KS = ChainCfg<BN>::KS;
  constexpr uint32_t SUB = (BM + BN) * BK, A_BYTES = BM * BK;
  const uint32_t a_lo0 = ((stage_addr >> 4) & 0x3FFF) | (1u << 16);
#pragma unroll
  for (int j = 0; j < KS / BK; ++j)
#pragma unroll
    for (int k = 0; k < BK / UMMA_K; ++k) {
      mma_i8(d_tmem, hi | (a_lo0 + ((j * SUB) >> 4) + 2 * k), hi | (a_lo0 + ((j * SUB + A_BYTES) >> 4) + 2 * k), idesc, acc);
      acc = 1;
    }
}

__device__ __forceinline__ void ch_epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(CH_EPI_THREADS) : "memory"); }

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, uint32_t src_smem, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(reinterpret_cast<uint64_t>(m)),
               "r"(src_smem), "r"(c0), "r"(c1)
               : "memory");
}

// 32 columns of one output row: folded requantize (+ residual qnn.add), packed to bytes and written to the staging tile
// (sa0 / sa1: shared addresses of the row's two 16-byte pieces).  Same arithmetic as epilogue_chunk's fast modes.
template <bool SUM>
__device__ __forceinline__ void chain_epi_chunk(const uint32_t (&r)[32], const ChainSeg& sg, uint32_t tab_addr, int col0,
                                                const uint8_t* sum_row, const int32_t* drow, uint32_t sa0, uint32_t sa1) {
  const FastEpi& f = sg.f;
  uint32_t sraw[8];
  if constexpr (SUM) {
    const uint4 s0 = ldg_cg_16(sum_row), s1 = ldg_cg_16(sum_row + 16);
    sraw[0] = s0.x ^ f.sum_xor; sraw[1] = s0.y ^ f.sum_xor; sraw[2] = s0.z ^ f.sum_xor; sraw[3] = s0.w ^ f.sum_xor;
    sraw[4] = s1.x ^ f.sum_xor; sraw[5] = s1.y ^ f.sum_xor; sraw[6] = s1.z ^ f.sum_xor; sraw[7] = s1.w ^ f.sum_xor;
  }
  uint32_t w[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    uint32_t word = 0;
    int4 dv = make_int4(0, 0, 0, 0);
    if (drow != nullptr) dv = __ldg((const int4*)(drow + col0) + i);  // border pixel: zero-point delta of its padding taps
    const int32_t dl[4] = {dv.x, dv.y, dv.z, dv.w};
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int j = 4 * i + jj;
      int32_t y = fast_requant((int32_t)(r[j] + (uint32_t)dl[jj]), lds128(tab_addr + j * 16), f.lo, f.hi);
      if constexpr (SUM) {
        const int32_t b = (int32_t)((sraw[i] >> (8 * jj)) & 0xFFu);
        y = min(max(fast_side(y, f.lhs) + fast_side(b, f.rhs), f.sum_lo), f.sum_hi);
      }
      word |= (uint32_t)(y & 0xFF) << (8 * jj);
    }
    w[i] = word;
  }
  sts128(sa0, make_int4((int)w[0], (int)w[1], (int)w[2], (int)w[3]));
  sts128(sa1, make_int4((int)w[4], (int)w[5], (int)w[6], (int)w[7]));
}

template <int BN>
__global__ void __launch_bounds__(CH_THREADS, 1) tc_chain_kernel(const __grid_constant__ ChainParams cp) {
  using Cfg = ChainCfg<BN>;
  static_assert(BN == 128, "the staging tile is 128 bytes wide (one 128-byte swizzle span per output row)");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* out_s = smem + Cfg::STAGES * Cfg::STAGE_BYTES;  // two staged output tiles (1024-byte aligned)
  uint8_t* tab_s = out_s + 2 * Cfg::OUT_BYTES;             // ring: the per-channel constants of each queued item's n-tile
  uint64_t* full = (uint64_t*)(tab_s + QD * Cfg::TAB_BYTES);
  uint64_t* empty = full + MAX_STAGES;
  uint64_t* tfull = empty + MAX_STAGES;
  uint64_t* tempty = tfull + 2;
  uint64_t* out_free = tempty + 2;   // staging buffer b may be overwritten (its TMA store has completed)
  uint64_t* q_full = out_free + 2;   // item queue: slot written by the producer ...
  uint64_t* q_empty = q_full + QD;   // ... and released by the MMA warp, the publisher and the 16 epilogue warps
  uint64_t* stored = q_empty + QD;   // per slot: the 16 epilogue warps have written the item's tile to the staging buffer
  int* q_item = (int*)(stored + QD);
  uint32_t* tmem_ptr = (uint32_t*)(q_item + QD);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < cp.nseg; ++i) { prefetch_tmap(&cp.tm[i]); prefetch_tmap(&cp.tmY[i]); }
    for (int i = 0; i < Cfg::STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], CH_EPI_WARPS); mbar_init(&out_free[i], 1); }
    for (int i = 0; i < QD; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 2 + CH_EPI_WARPS); mbar_init(&stored[i], CH_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<2 * BN>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  // everything below reads or writes global memory the preceding kernels of the stream produce (activations, the
  // zeroed counters): wait for them (no-op without the PDL launch attribute)
  asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp == 0) {
    // ===== claim + dependency wait + TMA producer =====
    int st = 0;
    uint32_t ph = 0, qn = 0;
    // position in the work list and its entry, fetched one item ahead: the atomic's and the load's round trips hide
    // behind the current item's loads
    int next_entry = -1;
    if (lane == 0) {
      const int idx = atomicAdd(cp.counter, 1);
      if (idx < cp.total_items) next_entry = __ldg(cp.items + idx);
    }
    for (;;) {
      const int entry = __shfl_sync(0xffffffffu, next_entry, 0);
      const bool fin = entry < 0;
      if (lane == 0 && !fin) {
        const int idx = atomicAdd(cp.counter, 1);
        next_entry = idx < cp.total_items ? __ldg(cp.items + idx) : -1;
      }
      const int si = fin ? 0 : entry >> ITEM_SHIFT;
      const ChainSeg& sg = cp.seg[si];
      const int tile = entry & ((1 << ITEM_SHIFT) - 1);
      const int nt = tile % sg.ntiles, mt = tile / sg.ntiles;
      const int m0 = mt * BM;
      const uint32_t slot = qn % QD;
      mbar_wait(&q_empty[slot], ((qn / QD) & 1) ^ 1);
      if (lane == 0) {
        q_item[slot] = entry;
        if (fin) mbar_arrive(&q_full[slot]);
        else {  // the n-tile's per-channel constants travel with the queue slot (constants: no dependency to wait for)
          mbar_expect_tx(&q_full[slot], (uint32_t)Cfg::TAB_BYTES);
          bulk_load(tab_s + slot * Cfg::TAB_BYTES, sg.f.tabA + nt * BN, (uint32_t)Cfg::TAB_BYTES, &q_full[slot]);
        }
      }
      __syncwarp();
      ++qn;
      if (fin) break;
      // the tile's images must be complete in the producing layer(s) of this launch
      if ((sg.depA_flags != nullptr || sg.depS_flags != nullptr) && !(sg.dbg & 16)) {  // dbg 16: timing experiment, no waits
        const int n_lo = m0 / sg.rows_per_img, n_hi = min(sg.M - 1, m0 + BM - 1) / sg.rows_per_img;
        for (int n = n_lo + lane; n <= n_hi; n += 32) {
          if (sg.depA_flags) wait_flag(sg.depA_flags + n, sg.depA_tgt[n]);
          if (sg.depS_flags) wait_flag(sg.depS_flags + n, sg.depS_tgt[n]);
        }
        __syncwarp();
        fence_proxy_async();  // order the acquire above before the async-proxy (TMA) reads below
      }
      const int a_bytes = BM * sg.bk, b_bytes = BN * sg.bk, sub_bytes = a_bytes + b_bytes;
      int cw = 0, chh = 0, cn = 0;
      if (sg.implicit) {
        const int q0 = m0 % sg.OW, t3 = m0 / sg.OW;
        cw = q0 * sg.sw - sg.pl;
        chh = (t3 % sg.OH) * sg.sh - sg.pt;
        cn = t3 / sg.OH;
      }
      int r = 0, s = 0, cc = 0;
      for (int kb0 = 0; kb0 < sg.num_kb; kb0 += sg.kps) {
        const int nsub = min(sg.kps, sg.num_kb - kb0);
        mbar_wait(&empty[st], ph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&full[st], (uint32_t)(nsub * sub_bytes));
          for (int j = 0; j < nsub; ++j) {
            uint8_t* sa = smem + st * Cfg::STAGE_BYTES + j * sub_bytes;
            const int kb = kb0 + j;
            if (sg.implicit) {
              tma_load_im2col_4d(sa, &cp.tm[si], &full[st], cc * sg.bk, cw, chh, cn, (uint16_t)(s * sg.dw), (uint16_t)(r * sg.dh));
              if (++cc == sg.cchunks) { cc = 0; if (++s == sg.KW) { s = 0; ++r; } }
            } else {
              tma_load_3d(sa, &cp.tm[si], &full[st], kb * sg.bk, m0, 0);
            }
            bulk_load(sa + a_bytes, sg.wpack + ((size_t)nt * sg.num_kb + kb) * b_bytes, (uint32_t)b_bytes, &full[st]);
          }
        }
        r = __shfl_sync(0xffffffffu, r, 0); s = __shfl_sync(0xffffffffu, s, 0); cc = __shfl_sync(0xffffffffu, cc, 0);
        if (++st == Cfg::STAGES) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    int st = 0;
    uint32_t ph = 0, lt = 0, qn = 0;
    const uint32_t smem_base = smem_u32(smem);
    for (;; ++lt) {
      const uint32_t slot = qn % QD;
      mbar_wait(&q_full[slot], (qn / QD) & 1);
      const int item = q_item[slot];
      __syncwarp();
      if (lane == 0) mbar_arrive(&q_empty[slot]);
      ++qn;
      if (item < 0) break;
      const ChainSeg& sg = cp.seg[item >> ITEM_SHIFT];
      const uint32_t as = lt & 1;
      mbar_wait(&tempty[as], ((lt >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      const uint64_t hi = (uint64_t)sg.desc_hi << 32;
      const int kper = sg.bk / UMMA_K, sub_bytes = (BM + BN) * sg.bk, a_bytes = BM * sg.bk;
      for (int kb0 = 0; kb0 < sg.num_kb; kb0 += sg.kps) {
        const int nsub = min(sg.kps, sg.num_kb - kb0);
        mbar_wait(&full[st], ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t stage_addr = smem_base + st * Cfg::STAGE_BYTES;
          const uint32_t acc = kb0 != 0;
          if (nsub * sg.bk == Cfg::KS && sg.bk == 128) issue_stage<BN, 128>(d_tmem, stage_addr, hi, sg.idesc, acc);
          else if (nsub * sg.bk == Cfg::KS && sg.bk == 64) issue_stage<BN, 64>(d_tmem, stage_addr, hi, sg.idesc, acc);
          else {
            uint32_t a2 = acc;
            for (int j = 0; j < nsub; ++j) {
              const uint32_t a_addr = stage_addr + j * sub_bytes;
              const uint32_t a_lo = ((a_addr >> 4) & 0x3FFF) | (1u << 16);
              const uint32_t b_lo = (((a_addr + a_bytes) >> 4) & 0x3FFF) | (1u << 16);
              for (int k = 0; k < kper; ++k) {
                mma_i8(d_tmem, hi | (a_lo + 2 * k), hi | (b_lo + 2 * k), sg.idesc, a2);
                a2 = 1;
              }
            }
          }
          tc_commit(&empty[st]);
          if (kb0 + nsub >= sg.num_kb) tc_commit(&tfull[as]);
        }
        __syncwarp();
        if (++st == Cfg::STAGES) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 2) {
    // ===== publisher: TMA-store the staged tile, wait until the store has completed (the bytes are in L2, the point
    // of coherence for every SM), then bump the completion counters of the tile's images =====
    uint32_t qn = 0, lt = 0;
    for (;; ++lt) {
      const uint32_t slot = qn % QD;
      const uint32_t par = (qn / QD) & 1;
      mbar_wait(&q_full[slot], par);
      const int item = q_item[slot];
      ++qn;
      if (item < 0) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&q_empty[slot]);
        break;
      }
      const int si = item >> ITEM_SHIFT;
      const ChainSeg& sg = cp.seg[si];
      const int tile = item & ((1 << ITEM_SHIFT) - 1);
      const int n0 = (tile % sg.ntiles) * BN, m0 = (tile / sg.ntiles) * BM;
      const uint32_t buf = lt & 1;
      mbar_wait(&stored[slot], par);
      if (lane == 0) {
        const long long tp0 = clock64();
        if (!(sg.dbg & 2)) {  // dbg 2: timing experiment, nothing stored
          tma_store_2d(&cp.tmY[si], smem_u32(out_s + buf * Cfg::OUT_BYTES), n0, m0);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        }
        mbar_arrive(&out_free[buf]);
        const long long tp1 = clock64();
        if (!(sg.dbg & 32)) {  // dbg 32: timing experiment, nothing published (needs 16 as well)
          fence_proxy_async();
          asm volatile("fence.acq_rel.gpu;" ::: "memory");
          const int n_lo = m0 / sg.rows_per_img, n_hi = min(sg.M - 1, m0 + BM - 1) / sg.rows_per_img;
          for (int n = n_lo; n <= n_hi; ++n)
            asm volatile("red.relaxed.gpu.global.add.s32 [%0], %1;" ::"l"(sg.flags + n), "r"(1) : "memory");
        }
        if (sg.dbg & 512) {  // statistics: [4] store issue -> complete, [5] fences + counter updates
          atomicAdd(&g_chain_stats[2], 1ull);
          atomicAdd(&g_chain_stats[4], (unsigned long long)(tp1 - tp0));
          atomicAdd(&g_chain_stats[5], (unsigned long long)(clock64() - tp1));
        }
        mbar_arrive(&q_empty[slot]);
      }
      __syncwarp();
    }
  } else {
    // ===== epilogue: a warp owns TMEM lanes 32 * (warp % 4) .. + 31 (hardware rule) and 32 of the tile's 128 columns =====
    const int q = warp & 3, cg = (warp - 3) >> 2;
    const uint32_t tab_base = smem_u32(tab_s);
    uint32_t lt = 0, qn = 0;
    // this thread's two 16-byte pieces of staging row (q * 32 + lane): piece index XOR (row & 7), the 128-byte swizzle
    const int srow = q * 32 + lane;
    const uint32_t sa_off0 = (uint32_t)(srow * 128 + (((2 * cg) ^ (srow & 7)) << 4));
    const uint32_t sa_off1 = (uint32_t)(srow * 128 + (((2 * cg + 1) ^ (srow & 7)) << 4));
    for (;; ++lt) {
      const uint32_t slot = qn % QD;
      mbar_wait(&q_full[slot], (qn / QD) & 1);
      const int item = q_item[slot];
      ++qn;
      if (item < 0) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&q_empty[slot]);
        break;
      }
      const ChainSeg& sg = cp.seg[item >> ITEM_SHIFT];
      const int tile = item & ((1 << ITEM_SHIFT) - 1);
      const int nt = tile % sg.ntiles, mt = tile / sg.ntiles;
      const int n0 = nt * BN, m0 = mt * BM;
      const int c = cg * 32;  // first column of this warp inside the tile
      const uint32_t tab_addr = tab_base + slot * Cfg::TAB_BYTES + (uint32_t)(c * 16);  // this item's constants: see the producer
      const int row = m0 + srow;
      const bool row_valid = row < sg.M;
      const int32_t* drow = nullptr;
      if (sg.delta != nullptr && row_valid) {
        const int qq = row % sg.OW, pp = (row / sg.OW) % sg.OH;
        const int cls = (int)sg.cls_h[pp] * sg.ncls_w + (int)sg.cls_w[qq];
        if (cls) drow = sg.delta + (long long)cls * sg.N;
      }
      const uint32_t as = lt & 1;
      mbar_wait(&tfull[as], (lt >> 1) & 1);
      tc_fence_after();
      const long long te0 = clock64();
      uint32_t r[32];
      const bool work = row_valid && !(sg.dbg & 1);
      // all lanes take part in the TMEM load (.sync.aligned); rows past M hold garbage that is never stored
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + c, r);
      tmem_ld_wait32(r);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);  // the accumulator stage goes back to the MMA warp
      mbar_wait(&out_free[as], ((lt >> 1) & 1) ^ 1);  // the TMA store of the tile that used this staging buffer is done
      if (work) {
        const uint32_t so = smem_u32(out_s + as * Cfg::OUT_BYTES);
        const uint8_t* sum_row = (const uint8_t*)sg.sum + (long long)row * sg.N + n0 + c;
        if (sg.f.has_sum) chain_epi_chunk<true>(r, sg, tab_addr, n0 + c, sum_row, drow, so + sa_off0, so + sa_off1);
        else chain_epi_chunk<false>(r, sg, tab_addr, n0 + c, sum_row, drow, so + sa_off0, so + sa_off1);
      }
      // generic-proxy writes to shared memory -> visible to the async proxy (the publisher's TMA store)
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&stored[slot]);
        mbar_arrive(&q_empty[slot]);  // only now: the slot also holds the table this warp has been reading
        if ((sg.dbg & 512) && warp == 3) atomicAdd(&g_chain_stats[3], (unsigned long long)(clock64() - te0));  // accumulator ready -> tile staged
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<2 * BN>(tmem_base);
}

// ---- host side -------------------------------------------------------------------------------------
bool make_tmap_im2col_pub(CUtensorMap* tm, const void* base, const ConvGeom& g, int bk);
bool make_tmap_3d_pub(CUtensorMap* tm, const void* base, long long k, long long rows, long long batch, int box_rows, int bk);
bool make_tmap_out_pub(CUtensorMap* tm, const void* base, long long n, long long rows);

int tc_chain_bk(const TcArgs& a) {
  if (a.implicit == 1) return a.g.C % 128 == 0 ? 128 : (a.g.C % 64 == 0 ? 64 : 32);
  return a.K % 128 == 0 ? 128 : (a.K % 64 == 0 ? 64 : 32);
}

bool tc_chain_supported(const TcArgs& a, int bn) {
  if (!a.f.enabled || a.B != 1 || a.w_batched) return false;
  if (a.implicit != 0 && a.implicit != 1) return false;
  if (bn != 128) return false;
  if (a.N % 128 != 0 || a.M < 1) return false;  // whole 128-byte output lines (TMA store of 128 x 128 staging tiles)
  if (a.e.out_dtype != kU8 && a.e.out_dtype != kI8) return false;
  const int bk = tc_chain_bk(a);
  if (a.implicit == 1) {
    if (!tc_implicit_conv_supported(a.g) || a.g.C % bk) return false;
  } else {
    if (a.K % bk) return false;
  }
  return true;
}

int tc_chain_flag_ints(const TcChainLayer* layers, int n) {
  int total = 1;  // [0] = the claim counter
  for (int i = 0; i < n; ++i) total += (layers[i].a.M + layers[i].rows_per_img - 1) / layers[i].rows_per_img;
  return total;
}

size_t tc_chain_state_bytes() { return sizeof(ChainParams); }

int tc_chain_item_count(const TcChainLayer* layers, int n, int bn) {
  long long total = 0;
  for (int i = 0; i < n; ++i)
    total += (long long)((layers[i].a.M + BM - 1) / BM) * ((layers[i].a.N + bn - 1) / bn);
  return total > 0x3fffffff ? -1 : (int)total;
}

// Order of the work list.  Every CTA takes the next unclaimed entry when it becomes free and waits if the entry's
// images are not complete yet, so the order decides how much of the machine idles: layer by layer, the deep layers
// (fewer tiles than SMs) run one dependent wave after the other.  The list is therefore produced by simulating a
// list scheduler with `workers` CTAs that always starts the DEEPEST layer that has a ready tile (depth first: the
// first images race ahead through the layers, so tiles of several layers -- of different images -- are in flight
// together and a stalled entry is rare).  The order is topological by construction: an entry is emitted only after
// every tile it depends on.  Costs are a model (cycles): they shape the order, never correctness.
static void chain_schedule(const TcChainLayer* layers, int n, int bn, int workers, std::vector<int>* order) {
  struct L { int mtiles, ntiles, tiles, rpi, nimg, M, da, ds; long long cost; };
  std::vector<L> ls(n);
  for (int i = 0; i < n; ++i) {
    const TcArgs& a = layers[i].a;
    L& l = ls[i];
    l.M = a.M; l.mtiles = (a.M + BM - 1) / BM; l.ntiles = (a.N + bn - 1) / bn; l.tiles = l.mtiles * l.ntiles;
    l.rpi = layers[i].rows_per_img; l.nimg = (a.M + l.rpi - 1) / l.rpi;
    l.da = layers[i].dep_a; l.ds = layers[i].dep_sum;
    l.cost = std::max<long long>((long long)(a.K / 32) * 150, 3000);  // MMA phase (shared-memory bound) vs epilogue floor
  }
  const long long latency = 8000;  // finish of a tile -> its bytes visible to a consumer's loads (epilogue, store, poll)
  std::vector<std::vector<long long>> done(n);   // per (layer, image): time the image is complete (-1: tiles outstanding)
  std::vector<std::vector<int>> left(n);         // per (layer, image): tiles not yet scheduled
  std::vector<std::vector<long long>> fin(n);    // per (layer, image): latest finish of its scheduled tiles
  for (int i = 0; i < n; ++i) {
    done[i].assign(ls[i].nimg, -1);
    fin[i].assign(ls[i].nimg, 0);
    left[i].assign(ls[i].nimg, 0);
    for (int img = 0; img < ls[i].nimg; ++img) {
      const long long r0 = (long long)img * ls[i].rpi, r1 = std::min<long long>(ls[i].M, r0 + ls[i].rpi) - 1;
      left[i][img] = (int)((r1 / BM - r0 / BM + 1) * ls[i].ntiles);
    }
  }
  std::vector<int> cur(n, 0);
  std::vector<long long> freeat(workers, 0);
  long long total = 0;
  for (auto& l : ls) total += l.tiles;
  order->clear();
  order->reserve((size_t)total);
  auto ready_time = [&](int i, int tile) -> long long {  // -1: a dependency has unscheduled tiles
    const L& l = ls[i];
    const int m0 = (tile / l.ntiles) * BM;
    const int n_lo = m0 / l.rpi, n_hi = std::min(l.M - 1, m0 + BM - 1) / l.rpi;
    long long t = 0;
    for (int d : {l.da, l.ds}) {
      if (d < 0) continue;
      for (int img = n_lo; img <= n_hi; ++img) {
        if (done[d][img] < 0) return -1;
        t = std::max(t, done[d][img] + latency);
      }
    }
    return t;
  };
  for (long long k = 0; k < total; ++k) {
    int w = 0;
    for (int j = 1; j < workers; ++j) if (freeat[j] < freeat[w]) w = j;
    const long long now = freeat[w];
    int best = -1;
    long long best_ready = 0;
    for (int i = n - 1; i >= 0; --i) {  // deepest layer first
      if (cur[i] >= ls[i].tiles) continue;
      const long long rt = ready_time(i, cur[i]);
      if (rt < 0) continue;
      if (rt <= now) { best = i; best_ready = rt; break; }
      if (best < 0 || rt < best_ready) { best = i; best_ready = rt; }
    }
    if (best < 0) {  // cannot happen for a DAG in layer order: the shallowest unfinished layer is always schedulable
      for (int i = 0; i < n; ++i) if (cur[i] < ls[i].tiles) { best = i; best_ready = now; break; }
    }
    const L& l = ls[best];
    const int tile = cur[best]++;
    const long long start = std::max(now, best_ready), end = start + l.cost;
    freeat[w] = end;
    order->push_back((best << ITEM_SHIFT) | tile);
    const int m0 = (tile / l.ntiles) * BM;
    const int n_lo = m0 / l.rpi, n_hi = std::min(l.M - 1, m0 + BM - 1) / l.rpi;
    for (int img = n_lo; img <= n_hi; ++img) {
      fin[best][img] = std::max(fin[best][img], end);
      if (--left[best][img] == 0) done[best][img] = fin[best][img];
    }
  }
}

// Fill `state` (host memory, tc_chain_state_bytes() long) for the layers; `flags` is device memory of
// tc_chain_flag_ints() ints (zeroed by the caller before every launch), `tgt_dev` device memory of the same size and
// `items_dev` device memory of tc_chain_item_count() ints, both filled by the caller from *tgt_host / *items_host.
int tc_chain_build(const TcChainLayer* layers, int n, int bn, int* flags, int* tgt_dev, std::vector<int>* tgt_host,
                   const int* items_dev, std::vector<int>* items_host, void* state) {
  if (n < 1 || n > TKB_CHAIN_MAX_SEG) return -1;
  ChainParams* cp = new (state) ChainParams();
  memset((void*)cp, 0, sizeof(ChainParams));
  cp->nseg = n;
  cp->counter = flags;
  cp->items = items_dev;
  tgt_host->assign(1, 0);
  std::vector<int> off(n);
  int item = 0;
  for (int i = 0; i < n; ++i) {
    const TcArgs& a = layers[i].a;
    if (!tc_chain_supported(a, bn)) return -1;
    if ((((uintptr_t)a.a) | ((uintptr_t)a.w) | ((uintptr_t)a.y)) & 15) return -1;
    ChainSeg& sg = cp->seg[i];
    const int bk = tc_chain_bk(a);
    sg.y = a.y; sg.sum = a.e.sum; sg.wpack = (const uint8_t*)a.w;
    sg.delta = a.delta; sg.cls_h = a.cls_h; sg.cls_w = a.cls_w; sg.ncls_w = a.ncls_w;
    sg.M = a.M; sg.N = a.N; sg.K = a.K;
    sg.implicit = a.implicit;
    sg.bk = bk;
    const int KS = 128;
    sg.kps = KS / bk;
    if (a.implicit) {
      const ConvGeom& g = a.g;
      sg.C = g.C; sg.KH = g.KH; sg.KW = g.KW; sg.OH = g.OH; sg.OW = g.OW;
      sg.sh = g.sh; sg.sw = g.sw; sg.dh = g.dh; sg.dw = g.dw; sg.pt = g.pt; sg.pl = g.pl;
      sg.cchunks = g.C / bk;
      sg.num_kb = g.KH * g.KW * sg.cchunks;
      if (!make_tmap_im2col_pub(&cp->tm[i], a.a, g, bk)) return -1;
    } else {
      sg.OH = 1; sg.OW = a.M;
      sg.num_kb = a.K / bk;
      if (!make_tmap_3d_pub(&cp->tm[i], a.a, a.K, a.M, 1, BM, bk)) return -1;
    }
    if (a.pack_bn != bn || a.pack_bk != bk || !a.w_packed) return -1;
    if (!make_tmap_out_pub(&cp->tmY[i], a.y, a.N, a.M)) return -1;
    sg.mtiles = (a.M + BM - 1) / BM;
    sg.ntiles = (a.N + bn - 1) / bn;
    sg.item_begin = item;
    item += sg.mtiles * sg.ntiles;
    sg.item_end = item;
    sg.rows_per_img = layers[i].rows_per_img;
    const uint32_t layout = bk == 128 ? 2u : (bk == 64 ? 4u : 6u);
    sg.desc_hi = (uint32_t)(smem_desc(0, 16, 8 * bk, layout) >> 32);
    sg.idesc = idesc_i8(BM, bn, a.a_dtype == kI8, a.w_dtype == kI8);
    sg.e.out_dtype = a.e.out_dtype;
    sg.e.bias = nullptr;
    sg.f = a.f;
    if (const char* d = getenv("TKB_DBG")) sg.dbg = atoi(d);
    // completion counters: every tile that touches an image arrives once
    const int nimg = (a.M + sg.rows_per_img - 1) / sg.rows_per_img;
    off[i] = (int)tgt_host->size();
    for (int img = 0; img < nimg; ++img) {
      const long long r0 = (long long)img * sg.rows_per_img, r1 = std::min<long long>(a.M, r0 + sg.rows_per_img) - 1;
      tgt_host->push_back((int)((r1 / BM - r0 / BM + 1) * sg.ntiles));  // one arrival (the publisher's) per tile
    }
    sg.flags = flags + off[i];
  }
  cp->total_items = item;
  for (int i = 0; i < n; ++i)
    if (cp->seg[i].mtiles * cp->seg[i].ntiles >= (1 << ITEM_SHIFT)) return -1;
  if (getenv("TKB_CHAIN_ORDER") && atoi(getenv("TKB_CHAIN_ORDER")) == 0) {  // A/B: plain layer order
    items_host->clear();
    for (int i = 0; i < n; ++i)
      for (int t = 0; t < cp->seg[i].mtiles * cp->seg[i].ntiles; ++t) items_host->push_back((i << ITEM_SHIFT) | t);
  } else {
    chain_schedule(layers, n, bn, num_sms(), items_host);
  }
  if ((int)items_host->size() != item) return -1;
  for (int i = 0; i < n; ++i) {
    ChainSeg& sg = cp->seg[i];
    const int da = layers[i].dep_a, ds = layers[i].dep_sum;
    auto nimg_of = [&](int j) { return (layers[j].a.M + layers[j].rows_per_img - 1) / layers[j].rows_per_img; };
    if (da >= i || ds >= i) return -1;
    if (da >= 0) {
      if (nimg_of(da) != nimg_of(i)) return -1;  // image index must mean the same thing on both sides
      sg.depA_flags = flags + off[da]; sg.depA_tgt = tgt_dev + off[da];
    }
    if (ds >= 0) {
      if (nimg_of(ds) != nimg_of(i)) return -1;
      sg.depS_flags = flags + off[ds]; sg.depS_tgt = tgt_dev + off[ds];
    }
  }
  return 0;
}

template <int BN>
static int launch_chain_t(const ChainParams& cp, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(tc_chain_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) return -1;
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)std::min(cp.total_items, num_sms()));
  cfg.blockDim = dim3(CH_THREADS);
  cfg.dynamicSmemBytes = ChainCfg<BN>::SMEM;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = getenv("TKB_NO_PDL") ? 0 : 1;
  if (cudaLaunchKernelEx(&cfg, tc_chain_kernel<BN>, cp) != cudaSuccess) return -1;
  ++g_launches;
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

void tc_chain_stats(unsigned long long* out, bool reset) {
  cudaMemcpyFromSymbol(out, g_chain_stats, sizeof(unsigned long long) * 8);
  if (reset) {
    unsigned long long z[8] = {0};
    cudaMemcpyToSymbol(g_chain_stats, z, sizeof(z));
  }
}

int launch_tc_chain(const void* state, int bn, cudaStream_t s) {
  const ChainParams& cp = *(const ChainParams*)state;
  if (cp.total_items <= 0) return 0;
  if (bn == 128) return launch_chain_t<128>(cp, s);
  return -1;
}

}  // namespace tkb
