// CUDA-core kernels: complete-coverage implementations of every operator on the QNN hot path.
// They are the path for shapes the tcgen05 kernels do not take (grouped / depthwise / dilated
// convolutions, non-zero weight zero points, K not a multiple of 16) and for the HBM-bound
// elementwise operators.  Arithmetic follows epilogue.cuh (bit-exact TVM integer lowering).
#include <cstdio>

#include "kernels.h"

namespace tkb {

thread_local long long g_launches = 0;

// Programmatic dependent launch for the CUDA-core kernels that sit between tcgen05 layers (stem unfold, max pool, casts,
// average pool, depthwise / pointwise convolutions, qnn.add): the kernel lets its successor start launching at once and
// waits for its predecessor's memory before it touches any tensor, so launch latency and the neighbour's tail overlap as
// they do between the tensor-core kernels (profiles/r02_notes.md section 17: worth ~2.3 us per boundary there).
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <typename... KArgs, typename... Args>
static void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  static const bool off = getenv("TKB_NO_PDL") != nullptr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = off ? 0 : 1;
  cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

static inline int blocks_for(long long n, int threads) {
  long long b = (n + threads - 1) / threads;
  const long long cap = 148LL * 32;  // grid-stride beyond 32 CTAs per SM
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

template <bool A_UNSIGNED, bool W_UNSIGNED>
__device__ __forceinline__ int dp4a_mixed(uint32_t a, uint32_t w, int c) {
  int d;
  if (A_UNSIGNED && W_UNSIGNED) asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(w), "r"(c));
  else if (A_UNSIGNED) asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(w), "r"(c));
  else if (W_UNSIGNED) asm("dp4a.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(w), "r"(c));
  else asm("dp4a.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(w), "r"(c));
  return d;
}

// Direct convolution, one thread per output element (n, oy, ox, o), o fastest.
// acc = sum over in-image taps of (a - zp_a)(w - zp_w)  (src/relay/qnn/op/convolution.cc:708-811;
// out-of-image taps read zp_a and contribute 0, :227-253).
template <bool AU, bool WU>
__global__ void conv2d_simt_kernel(const uint8_t* __restrict__ x, const uint8_t* __restrict__ w, void* __restrict__ y,
                                   ConvGeom g, EpiParams e) {
  const int Cg = g.C / g.groups, Og = g.O / g.groups;
  const long long total = (long long)g.N * g.OH * g.OW * g.O;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int o = (int)(idx % g.O);
    long long t = idx / g.O;
    const int ox = (int)(t % g.OW);
    t /= g.OW;
    const int oy = (int)(t % g.OH);
    const int n = (int)(t / g.OH);
    const int grp = o / Og;
    const int32_t zw = g.zp_w ? g.zp_w[o] : g.zp_w0;
    int raw = 0, sum_a = 0, sum_w = 0, taps = 0;
    for (int r = 0; r < g.KH; ++r) {
      const int iy = oy * g.sh + r * g.dh - g.pt;
      if (iy < 0 || iy >= g.H) continue;
      for (int s = 0; s < g.KW; ++s) {
        const int ix = ox * g.sw + s * g.dw - g.pl;
        if (ix < 0 || ix >= g.W) continue;
        const uint8_t* ap = x + (((long long)n * g.H + iy) * g.W + ix) * g.C + grp * Cg;
        const uint8_t* wp = w + (((long long)o * g.KH + r) * g.KW + s) * Cg;
        int c = 0;
        if ((Cg & 3) == 0 && ((((uintptr_t)ap) | ((uintptr_t)wp)) & 3) == 0) {
          for (; c < Cg; c += 4) {
            const uint32_t av = *(const uint32_t*)(ap + c), wv = *(const uint32_t*)(wp + c);
            raw = dp4a_mixed<AU, WU>(av, wv, raw);
            sum_a = dp4a_mixed<AU, false>(av, 0x01010101u, sum_a);
            sum_w = dp4a_mixed<false, WU>(0x01010101u, wv, sum_w);
          }
        } else {
          for (; c < Cg; ++c) {
            const int av = AU ? (int)ap[c] : (int)(int8_t)ap[c];
            const int wv = WU ? (int)wp[c] : (int)(int8_t)wp[c];
            raw += av * wv;
            sum_a += av;
            sum_w += wv;
          }
        }
        taps += Cg;
      }
    }
    // sum (a-za)(w-zw) = raw - za*sum_w - zw*sum_a + taps*za*zw   (int32 ring arithmetic)
    const uint32_t acc = (uint32_t)raw - (uint32_t)g.zp_a * (uint32_t)sum_w - (uint32_t)zw * (uint32_t)sum_a +
                         (uint32_t)taps * (uint32_t)g.zp_a * (uint32_t)zw;
    store_from_i32(y, e.out_dtype, idx, epilogue_one((int32_t)acc, o, idx, e));
  }
}

void launch_conv2d_simt(const void* x, const void* w, void* y, const ConvGeom& g, const EpiParams& e,
                        cudaStream_t s) {
  const long long total = (long long)g.N * g.OH * g.OW * g.O;
  const int th = 256, bl = blocks_for(total, th);
  const uint8_t* xp = (const uint8_t*)x;
  const uint8_t* wp = (const uint8_t*)w;
  const bool au = g.a_dtype == kU8, wu = g.w_dtype == kU8;
  if (au && wu) conv2d_simt_kernel<true, true><<<bl, th, 0, s>>>(xp, wp, y, g, e);
  else if (au) conv2d_simt_kernel<true, false><<<bl, th, 0, s>>>(xp, wp, y, g, e);
  else if (wu) conv2d_simt_kernel<false, true><<<bl, th, 0, s>>>(xp, wp, y, g, e);
  else conv2d_simt_kernel<false, false><<<bl, th, 0, s>>>(xp, wp, y, g, e);
  ++g_launches;
}

// Depthwise KH x KW convolution (groups == C == O, one input channel per filter: MobileNet-class layers; arithmetic
// src/relay/qnn/op/convolution.cc:272-413 of the reference) on dense channels-last 8-bit tensors.  One thread owns
// 16 consecutive channels of one output pixel: per tap one 128-bit activation load and one 128-bit load of the
// tap-major weights [KH*KW][C] (L1-resident), 16 integer MACs.  Taps outside the image read the zero point, so
// that  acc = sum_taps x' * w  and the per-channel constant  corr[c] = -zp_a * sum_taps w[c]  (int32 ring, added
// by the epilogue) give exactly  sum_{in-image taps} (x - zp_a) * w .  3-D grid: no 64-bit index arithmetic.
// Epilogue: the host-folded FastEpi form when available (corr is then already inside its table), else the
// generic epilogue_one.
template <bool AU, bool WU>
__global__ void __launch_bounds__(256) dwconv_nhwc_kernel(const uint8_t* __restrict__ x, const uint8_t* __restrict__ w,
                                                          const int32_t* __restrict__ corr, void* __restrict__ y,
                                                          ConvGeom g, EpiParams e, FastEpi f) {
  const int cg = g.C >> 4;
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= g.OW * cg) return;
  const int ox = t / cg, c0 = (t - ox * cg) * 16;
  const int oy = blockIdx.y, n = blockIdx.z;
  const uint32_t zp4 = (uint32_t)(g.zp_a & 0xFF) * 0x01010101u;
  int32_t acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0;
  const uint8_t* img = x + (long long)n * g.H * g.W * g.C + c0;
  for (int r = 0; r < g.KH; ++r) {
    const int iy = oy * g.sh + r * g.dh - g.pt;
    const bool yin = iy >= 0 && iy < g.H;
    for (int q = 0; q < g.KW; ++q) {
      const int ix = ox * g.sw + q * g.dw - g.pl;
      uint4 a = make_uint4(zp4, zp4, zp4, zp4);
      if (yin && ix >= 0 && ix < g.W) a = __ldg((const uint4*)(img + ((long long)iy * g.W + ix) * g.C));
      const uint4 b = __ldg((const uint4*)(w + (long long)(r * g.KW + q) * g.C + c0));
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const uint32_t ab = (aw[i >> 2] >> (8 * (i & 3))) & 0xFFu, bb = (bw[i >> 2] >> (8 * (i & 3))) & 0xFFu;
        const int av = AU ? (int)ab : (int)(int8_t)ab, bv = WU ? (int)bb : (int)(int8_t)bb;
        acc[i] += av * bv;
      }
    }
  }
  const long long idx0 = (((long long)n * g.OH + oy) * g.OW + ox) * g.C + c0;
  const bool out8 = e.out_dtype == kU8 || e.out_dtype == kI8;
  if (f.enabled && !f.has_sum && out8) {
    uint32_t o4[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t word = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int4 tb = __ldg(f.tabA + c0 + 4 * i + j);
        const long long cc = ((long long)tb.w << 32) | (unsigned long long)(uint32_t)tb.z;
        const int32_t v = (int32_t)(((long long)acc[4 * i + j] * (long long)tb.x + cc) >> 32) >> tb.y;
        word |= (uint32_t)(min(max(v, f.lo), f.hi) & 0xFF) << (8 * j);
      }
      o4[i] = word;
    }
    *(uint4*)((uint8_t*)y + idx0) = make_uint4(o4[0], o4[1], o4[2], o4[3]);
  } else {
#pragma unroll 1
    for (int i = 0; i < 16; ++i) {
      const int32_t a = (int32_t)((uint32_t)acc[i] + (corr ? (uint32_t)corr[c0 + i] : 0u));
      store_from_i32(y, e.out_dtype, idx0 + i, epilogue_one(a, c0 + i, idx0 + i, e));
    }
  }
}

// 3 x 3 depthwise, stride S in {1, 2}: one thread owns 4 consecutive channels (one 32-bit word per pixel) of TWO
// vertically adjacent outputs.  For every input row the three neighbouring pixel words are byte-transposed (PRMT)
// into one word per channel holding that channel's three taps, and dp4a against the row-packed weights
// wrow[r][c] = {w(r,0), w(r,1), w(r,2), 0} does 3 MACs per instruction; the transposed rows are shared by the two
// outputs.  ~16 instructions per output element instead of ~37 for the 16-channel kernel above.
template <bool AU, bool WU, int S>
__global__ void __launch_bounds__(256) dwconv3x3_nhwc_kernel(const uint8_t* __restrict__ x, const uint32_t* __restrict__ wrow,
                                                             const int32_t* __restrict__ corr, void* __restrict__ y,
                                                             ConvGeom g, EpiParams e, FastEpi f) {
  constexpr int ROWS = S + 3;
  const int cq = g.C >> 2;
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= g.OW * cq) return;
  const int ox = t / cq, c0 = (t - ox * cq) * 4;
  const int oy0 = blockIdx.y * 2, n = blockIdx.z;
  const uint32_t zp4 = (uint32_t)(g.zp_a & 0xFF) * 0x01010101u;
  const uint8_t* img = x + (long long)n * g.H * g.W * g.C + c0;
  int xoff[3];
  bool xin[3];
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    const int ix = ox * S - g.pl + q;
    xin[q] = ix >= 0 && ix < g.W;
    xoff[q] = ix * g.C;
  }
  uint32_t tr[ROWS][4];
#pragma unroll
  for (int j = 0; j < ROWS; ++j) {
    const int iy = oy0 * S - g.pt + j;
    const bool yin = iy >= 0 && iy < g.H;
    const uint8_t* row = img + (long long)iy * g.W * g.C;
    uint32_t pw[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) pw[q] = (yin && xin[q]) ? __ldg((const uint32_t*)(row + xoff[q])) : zp4;
    const uint32_t a = __byte_perm(pw[0], pw[1], 0x5140), b = __byte_perm(pw[0], pw[1], 0x7362);
    const uint32_t c2 = __byte_perm(pw[2], pw[2], 0x5140), d = __byte_perm(pw[2], pw[2], 0x7362);
    tr[j][0] = __byte_perm(a, c2, 0x5410);
    tr[j][1] = __byte_perm(a, c2, 0x7632);
    tr[j][2] = __byte_perm(b, d, 0x5410);
    tr[j][3] = __byte_perm(b, d, 0x7632);
  }
  int32_t acc[2][4];
#pragma unroll
  for (int c = 0; c < 4; ++c) acc[0][c] = acc[1][c] = 0;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const uint4 wv = __ldg((const uint4*)(wrow + (long long)r * g.C + c0));
    const uint32_t ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      acc[0][c] = dp4a_mixed<AU, WU>(tr[r][c], ww[c], acc[0][c]);
      acc[1][c] = dp4a_mixed<AU, WU>(tr[S + r][c], ww[c], acc[1][c]);
    }
  }
  const bool fast = f.enabled && !f.has_sum && (e.out_dtype == kU8 || e.out_dtype == kI8);
#pragma unroll
  for (int o = 0; o < 2; ++o) {
    const int oy = oy0 + o;
    if (oy >= g.OH) break;
    const long long idx0 = (((long long)n * g.OH + oy) * g.OW + ox) * g.C + c0;
    if (fast) {
      uint32_t word = 0;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int4 tb = __ldg(f.tabA + c0 + c);
        const long long cc = ((long long)tb.w << 32) | (unsigned long long)(uint32_t)tb.z;
        const int32_t v = (int32_t)(((long long)acc[o][c] * (long long)tb.x + cc) >> 32) >> tb.y;
        word |= (uint32_t)(min(max(v, f.lo), f.hi) & 0xFF) << (8 * c);
      }
      *(uint32_t*)((uint8_t*)y + idx0) = word;
    } else {
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        const int32_t a2 = (int32_t)((uint32_t)acc[o][c] + (corr ? (uint32_t)corr[c0 + c] : 0u));
        store_from_i32(y, e.out_dtype, idx0 + c, epilogue_one(a2, c0 + c, idx0 + c, e));
      }
    }
  }
}

// 3 x 3 depthwise, 16 channels per thread: 128-bit loads and stores (north_star: "coalesced 128-bit vectorised loads").
// One thread owns 16 consecutive channels of NV vertically adjacent outputs of one column.  It walks the NV*S + 2
// input rows once: per row three 128-bit loads (left / centre / right pixel; neighbouring threads of the block read the
// same lines, so two of the three hit L1), the same PRMT byte transposition + dp4a as the 4-channel kernel above per
// 32-bit word, accumulated straight into the outputs the row belongs to (no row buffer: 16 * NV accumulators).  The
// per-channel epilogue constants are loaded once and used for all NV outputs.  Taps outside the image read the zero
// point (see dwconv_nhwc_kernel for the identity that makes this exact).
// cvt.pack.sat: four int32 -> four saturated 8-bit values in one word (byte 0 = v0)
template <bool UNSIGNED>
__device__ __forceinline__ uint32_t dw_pack4_sat8(int32_t v0, int32_t v1, int32_t v2, int32_t v3) {
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

// (round 2, profiles/r02_notes.md section 18) The first version walked the rows one after the other -- three loads, wait,
// compute -- with 168 registers (64 accumulators + 48 weight registers), i.e. 12 warps per SM each paying a full memory
// latency per row: 1.2-1.9 TB/s.  Now the weights live in shared memory and ALL rows are requested before the first is
// used, so a thread pays the latency once.
template <bool AU, bool WU, int S, int NV>
__global__ void __launch_bounds__(128) dwconv3x3_v16_kernel(const uint8_t* __restrict__ x, const uint32_t* __restrict__ wrow,
                                                            void* __restrict__ y, ConvGeom g, FastEpi f) {
  constexpr int ROWS = (NV - 1) * S + 3;
  // Shared memory, both arrays laid out [item][channel group] so that the lanes of a warp (consecutive channel groups)
  // read consecutive 16-byte words: the per-thread form (16 table entries 256 bytes apart from the next lane's) cost 32
  // L1 wavefronts per load instruction and made the kernel L1-bound (ncu: L1/TEX throughput 87-89 %, DRAM 10-20 %).
  //   dw_ws [3 rows][4 words][cg]: row-packed weights, one 32-bit lane = one channel's {w(r,0), w(r,1), w(r,2), 0}
  //   dw_tab[16 channels][cg]    : folded requantize constants {M, T - 32, C_lo, C_hi}
  extern __shared__ uint4 dw_ws[];
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // (the constants below do not depend on the predecessor)
  const int cg = g.C >> 4;
  int4* dw_tab = (int4*)(dw_ws + 12 * cg);
  for (int i = threadIdx.x; i < 12 * cg; i += 128) {   // i = (r * 4 + k) * cg + group
    const int grp = i % cg, rk = i / cg;
    dw_ws[i] = __ldg((const uint4*)wrow + (rk >> 2) * (g.C >> 2) + grp * 4 + (rk & 3));
  }
  for (int i = threadIdx.x; i < 16 * cg; i += 128) {   // i = e * cg + group
    const int grp = i % cg, e = i / cg;
    dw_tab[i] = __ldg(f.tabA + grp * 16 + e);
  }
  asm volatile("griddepcontrol.wait;" ::: "memory");
  __syncthreads();
  const int t = blockIdx.x * 128 + threadIdx.x;
  if (t >= g.OW * cg) return;
  const int ox = t / cg, grp0 = t - ox * cg, c0 = grp0 * 16;
  const int oy0 = blockIdx.y * NV, n = blockIdx.z;
  const uint32_t zp4 = (uint32_t)(g.zp_a & 0xFF) * 0x01010101u;
  const uint8_t* img = x + (long long)n * g.H * g.W * g.C + c0;
  int xoff[3];
  bool xin[3];
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    const int ix = ox * S - g.pl + q;
    xin[q] = ix >= 0 && ix < g.W;
    xoff[q] = ix * g.C;
  }
  uint4 p[ROWS][3];
#pragma unroll
  for (int j = 0; j < ROWS; ++j) {
    const int iy = oy0 * S - g.pt + j;
    const bool yin = iy >= 0 && iy < g.H;
    const uint8_t* row = img + (long long)iy * g.W * g.C;
#pragma unroll
    for (int q = 0; q < 3; ++q) p[j][q] = (yin && xin[q]) ? __ldg((const uint4*)(row + xoff[q])) : make_uint4(zp4, zp4, zp4, zp4);
  }
  int32_t acc[NV][16];
#pragma unroll
  for (int o = 0; o < NV; ++o)
#pragma unroll
    for (int c = 0; c < 16; ++c) acc[o][c] = 0;
  const uint4* wt = dw_ws + grp0;
#pragma unroll
  for (int j = 0; j < ROWS; ++j) {
    const uint32_t p0[4] = {p[j][0].x, p[j][0].y, p[j][0].z, p[j][0].w}, p1[4] = {p[j][1].x, p[j][1].y, p[j][1].z, p[j][1].w},
                   p2[4] = {p[j][2].x, p[j][2].y, p[j][2].z, p[j][2].w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // 4 channels: three pixel words -> one word per channel holding its three taps
      const uint32_t a = __byte_perm(p0[k], p1[k], 0x5140), b = __byte_perm(p0[k], p1[k], 0x7362);
      const uint32_t c2 = __byte_perm(p2[k], p2[k], 0x5140), d = __byte_perm(p2[k], p2[k], 0x7362);
      const uint32_t tr[4] = {__byte_perm(a, c2, 0x5410), __byte_perm(a, c2, 0x7632), __byte_perm(b, d, 0x5410), __byte_perm(b, d, 0x7632)};
#pragma unroll
      for (int o = 0; o < NV; ++o) {
        const int r = j - o * S;  // filter row this input row is for output o (compile time)
        if (r < 0 || r > 2) continue;
        const uint4 w4 = wt[(r * 4 + k) * cg];
        const uint32_t ww[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[o][4 * k + c] = dp4a_mixed<AU, WU>(tr[c], ww[c], acc[o][4 * k + c]);
      }
    }
  }
  // folded requantize + clamp (FastEpi, epilogue.cuh); the zero-point correction is already inside the table.
  // f.sat (1: the clamp is the uint8 range, 2: the int8 range): the saturating pack does the clamp, 2.5 instead of 6
  // instructions per output.
  uint32_t out[NV][4];
  if (f.sat) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      int32_t v[NV][4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int4 tb = dw_tab[(4 * k + c) * cg + grp0];
        const long long cc = ((long long)tb.w << 32) | (unsigned long long)(uint32_t)tb.z;
#pragma unroll
        for (int o = 0; o < NV; ++o) v[o][c] = (int32_t)(((long long)acc[o][4 * k + c] * (long long)tb.x + cc) >> 32) >> tb.y;
      }
#pragma unroll
      for (int o = 0; o < NV; ++o) out[o][k] = f.sat == 1 ? dw_pack4_sat8<true>(v[o][0], v[o][1], v[o][2], v[o][3]) : dw_pack4_sat8<false>(v[o][0], v[o][1], v[o][2], v[o][3]);
    }
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
      for (int o = 0; o < NV; ++o) out[o][k] = 0;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int4 tb = dw_tab[(4 * k + c) * cg + grp0];
        const long long cc = ((long long)tb.w << 32) | (unsigned long long)(uint32_t)tb.z;
#pragma unroll
        for (int o = 0; o < NV; ++o) {
          const int32_t v = (int32_t)(((long long)acc[o][4 * k + c] * (long long)tb.x + cc) >> 32) >> tb.y;
          out[o][k] |= (uint32_t)(min(max(v, f.lo), f.hi) & 0xFF) << (8 * c);
        }
      }
    }
  }
#pragma unroll
  for (int o = 0; o < NV; ++o) {
    const int oy = oy0 + o;
    if (oy >= g.OH) break;
    *(uint4*)((uint8_t*)y + (((long long)n * g.OH + oy) * g.OW + ox) * g.C + c0) = make_uint4(out[o][0], out[o][1], out[o][2], out[o][3]);
  }
}

// ---- pointwise (1x1, stride 1, unpadded) convolution with very few input channels (8 .. 32): CUDA-core dp4a kernel -----------
// MobileNetV2's 32 -> 16, 16 -> 96 and 24 -> 144 layers at 112x112 / 56x56 have one to two tcgen05 MMAs per 128-pixel tile; on
// the tensor-core kernel they are bound by ~860 cycles of barrier round trips per tile (profiles/r02_notes.md section 13), at
// a fifth of the memory roofline.  Here a thread owns 4 consecutive pixels x CB consecutive output channels: the pixels' input
// bytes live in registers, weights and the folded requantize table are staged once per block in shared memory as
// [item][channel block] (consecutive lanes = consecutive channel blocks read consecutive 16-byte words), and the CB outputs
// of a pixel leave as one 8- or 16-byte store, the lanes of a warp covering contiguous bytes of the NHWC output.
template <bool AU, bool WU, int KB, int CB>
__global__ void __launch_bounds__(256) pwconv_kernel(const uint8_t* __restrict__ x, const uint32_t* __restrict__ w,
                                                     uint8_t* __restrict__ y, long long M, int N, FastEpi f) {
  constexpr int KW = KB / 4, KW4 = (KW + 3) / 4, PX = 4;  // input words per pixel, weight uint4 per channel, pixels per thread
  extern __shared__ uint4 pw_smem[];
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // (the constants below do not depend on the predecessor)
  const int NB = N / CB;
  uint4* ws = pw_smem;                           // [CB][KW4][NB]
  int4* tab = (int4*)(pw_smem + CB * KW4 * NB);  // [CB][NB]
  for (int i = threadIdx.x; i < CB * KW4 * NB; i += 256) {
    const int nb = i % NB, r = i / NB, j = r % KW4, c = r / KW4;
    ws[i] = __ldg((const uint4*)w + (long long)(nb * CB + c) * KW4 + j);
  }
  for (int i = threadIdx.x; i < CB * NB; i += 256) {
    const int nb = i % NB, c = i / NB;
    tab[i] = __ldg(f.tabA + nb * CB + c);
  }
  asm volatile("griddepcontrol.wait;" ::: "memory");
  __syncthreads();
  // A thread's 4 pixels are PS = 32 apart (pixel slot q -> pixels (q / 32) * 128 + q % 32 + 32 * i): for a given i the
  // lanes of a warp then read / write neighbouring pixels, i.e. a contiguous stretch of memory per instruction.  (With 4
  // consecutive pixels per thread every 16-byte load of a warp touched 32 different cache lines: L1-bound at 72 %.)
  constexpr int PS = 32;
  const long long slots = (M + PS * PX - 1) / (PS * PX) * PS, total = slots * NB;
  for (long long t = (long long)blockIdx.x * 256 + threadIdx.x; t < total; t += (long long)gridDim.x * 256) {
    const int nb = (int)(t % NB);
    const long long q = t / NB;
    const long long p0 = (q / PS) * (PS * PX) + (q % PS);
    uint32_t xin[PX][KW];
#pragma unroll
    for (int p = 0; p < PX; ++p) {
      const bool ok = p0 + (long long)p * PS < M;
      const uint8_t* xp = x + (p0 + (long long)p * PS) * KB;
      if constexpr (KB % 16 == 0) {
#pragma unroll
        for (int j = 0; j < KB / 16; ++j) {
          const uint4 v = ok ? __ldg((const uint4*)xp + j) : make_uint4(0, 0, 0, 0);
          xin[p][4 * j] = v.x; xin[p][4 * j + 1] = v.y; xin[p][4 * j + 2] = v.z; xin[p][4 * j + 3] = v.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < KB / 8; ++j) {
          const uint2 v = ok ? __ldg((const uint2*)xp + j) : make_uint2(0, 0);
          xin[p][2 * j] = v.x; xin[p][2 * j + 1] = v.y;
        }
      }
    }
    uint32_t out[PX][CB / 4];
#pragma unroll
    for (int c4 = 0; c4 < CB / 4; ++c4) {
      int32_t v[PX][4];
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        const int c = 4 * c4 + cc;
        uint32_t ww[4 * KW4];
#pragma unroll
        for (int j = 0; j < KW4; ++j) {
          const uint4 q = ws[(c * KW4 + j) * NB + nb];
          ww[4 * j] = q.x; ww[4 * j + 1] = q.y; ww[4 * j + 2] = q.z; ww[4 * j + 3] = q.w;
        }
        const int4 tb = tab[c * NB + nb];
        const long long add = ((long long)tb.w << 32) | (unsigned long long)(uint32_t)tb.z;
#pragma unroll
        for (int p = 0; p < PX; ++p) {
          int32_t acc = 0;
#pragma unroll
          for (int k = 0; k < KW; ++k) acc = dp4a_mixed<AU, WU>(xin[p][k], ww[k], acc);
          v[p][cc] = (int32_t)(((long long)acc * (long long)tb.x + add) >> 32) >> tb.y;
        }
      }
#pragma unroll
      for (int p = 0; p < PX; ++p) {
        if (f.sat == 1) out[p][c4] = dw_pack4_sat8<true>(v[p][0], v[p][1], v[p][2], v[p][3]);
        else if (f.sat == 2) out[p][c4] = dw_pack4_sat8<false>(v[p][0], v[p][1], v[p][2], v[p][3]);
        else {
          uint32_t word = 0;
#pragma unroll
          for (int cc = 0; cc < 4; ++cc) word |= (uint32_t)(min(max(v[p][cc], f.lo), f.hi) & 0xFF) << (8 * cc);
          out[p][c4] = word;
        }
      }
    }
#pragma unroll
    for (int p = 0; p < PX; ++p) {
      if (p0 + (long long)p * PS >= M) break;
      uint8_t* dst = y + (p0 + (long long)p * PS) * N + nb * CB;
      if constexpr (CB == 16) *(uint4*)dst = make_uint4(out[p][0], out[p][1], out[p][2], out[p][3]);
      else *(uint2*)dst = make_uint2(out[p][0], out[p][1]);
    }
  }
}

bool pwconv_supported(int K, int N) { return (K == 8 || K == 16 || K == 24 || K == 32) && N % 8 == 0 && N >= 8 && N <= 1024; }

// x: [M][K] bytes (NHWC pixels), w: [N][ceil(K/16)*16] bytes (rows zero-padded), y: [M][N] bytes
int launch_pwconv(const void* x, const void* w, void* y, long long M, int N, int K, int a_dtype, int w_dtype, const FastEpi& f,
                  cudaStream_t s) {
  if (!pwconv_supported(K, N) || !f.enabled || f.has_sum || M <= 0) return -1;
  if ((((uintptr_t)x) | ((uintptr_t)y) | ((uintptr_t)w)) & 15) return -1;
  const int CB = N % 16 == 0 ? 16 : 8, NB = N / CB, KW4 = (K / 4 + 3) / 4;
  const size_t smem = (size_t)(CB * KW4 * NB + CB * NB) * 16;
  if (smem > 48 * 1024) return -1;
  const long long total = ((M + 127) / 128) * 32 * NB, blocks = (total + 255) / 256;
  const unsigned grid = (unsigned)(blocks < 148 * 8 ? blocks : 148 * 8);
  const bool au = a_dtype == kU8, wu = w_dtype == kU8;
#define TKB_PW(AU_, WU_, KB_, CB_)                                                                                        \
  if (au == AU_ && wu == WU_ && K == KB_ && CB == CB_) {                                                                  \
    launch_pdl(pwconv_kernel<AU_, WU_, KB_, CB_>, dim3(grid), dim3(256), smem, s, (const uint8_t*)x, (const uint32_t*)w, (uint8_t*)y, M, N, f); \
    ++g_launches;                                                                                                          \
    return cudaGetLastError() == cudaSuccess ? 0 : -1;                                                                     \
  }
#define TKB_PW_K(AU_, WU_) TKB_PW(AU_, WU_, 8, 16) TKB_PW(AU_, WU_, 8, 8) TKB_PW(AU_, WU_, 16, 16) TKB_PW(AU_, WU_, 16, 8) \
  TKB_PW(AU_, WU_, 24, 16) TKB_PW(AU_, WU_, 24, 8) TKB_PW(AU_, WU_, 32, 16) TKB_PW(AU_, WU_, 32, 8)
  TKB_PW_K(true, false) TKB_PW_K(true, true) TKB_PW_K(false, false) TKB_PW_K(false, true)
#undef TKB_PW_K
#undef TKB_PW
  return -1;
}

bool dwconv_nhwc_supported(const ConvGeom& g) {
  return g.groups == g.C && g.O == g.C && (g.C % 16) == 0 && g.N <= 65535 && g.OH <= 65535 && g.zp_w == nullptr &&
         g.zp_w0 == 0 && g.KH * g.KW <= 49;
}

void launch_dwconv_nhwc(const void* x, const void* w_tap, const void* w_row3, const int32_t* corr, void* y,
                        const ConvGeom& g, const EpiParams& e, const FastEpi& f, cudaStream_t s) {
  if ((long long)g.N * g.OH * g.OW == 0) return;
  if (w_row3 && g.KH == 3 && g.KW == 3 && g.dh == 1 && g.dw == 1 && g.sh == g.sw && (g.sh == 1 || g.sh == 2) &&
      f.enabled && !f.has_sum && (e.out_dtype == kU8 || e.out_dtype == kI8) && !getenv("TKB_DW_OLD") &&
      ((((uintptr_t)x) | ((uintptr_t)y) | ((uintptr_t)w_row3)) & 15) == 0) {
    const uint8_t* xp = (const uint8_t*)x;
    const uint32_t* wp = (const uint32_t*)w_row3;
    const bool au = g.a_dtype == kU8, wu = g.w_dtype == kU8;
    constexpr int NV1 = 4, NV2 = 2;  // outputs per thread: stride 1 walks 6 rows for 4 outputs, stride 2 walks 5 for 2
    const int nv = g.sh == 1 ? NV1 : NV2;
    dim3 grid((unsigned)((g.OW * (g.C / 16) + 127) / 128), (unsigned)((g.OH + nv - 1) / nv), (unsigned)g.N);
#define TKB_DW16(AU_, WU_)                                                                                 \
  if (au == AU_ && wu == WU_) {                                                                            \
    if (g.sh == 1) launch_pdl(dwconv3x3_v16_kernel<AU_, WU_, 1, NV1>, grid, dim3(128), (size_t)28 * g.C, s, xp, wp, y, g, f); \
    else launch_pdl(dwconv3x3_v16_kernel<AU_, WU_, 2, NV2>, grid, dim3(128), (size_t)28 * g.C, s, xp, wp, y, g, f);          \
  }
    TKB_DW16(true, true) TKB_DW16(true, false) TKB_DW16(false, true) TKB_DW16(false, false)
#undef TKB_DW16
    ++g_launches;
    return;
  }
  if (w_row3 && g.KH == 3 && g.KW == 3 && g.dh == 1 && g.dw == 1 && g.sh == g.sw && (g.sh == 1 || g.sh == 2) &&
      g.OH <= 131070) {
    dim3 grid((unsigned)((g.OW * (g.C / 4) + 255) / 256), (unsigned)((g.OH + 1) / 2), (unsigned)g.N);
    const uint8_t* xp = (const uint8_t*)x;
    const uint32_t* wp = (const uint32_t*)w_row3;
    const bool au = g.a_dtype == kU8, wu = g.w_dtype == kU8;
#define TKB_DW3(AU_, WU_)                                                                                   \
  if (au == AU_ && wu == WU_) {                                                                             \
    if (g.sh == 1) dwconv3x3_nhwc_kernel<AU_, WU_, 1><<<grid, 256, 0, s>>>(xp, wp, corr, y, g, e, f);       \
    else dwconv3x3_nhwc_kernel<AU_, WU_, 2><<<grid, 256, 0, s>>>(xp, wp, corr, y, g, e, f);                 \
  }
    TKB_DW3(true, true) TKB_DW3(true, false) TKB_DW3(false, true) TKB_DW3(false, false)
#undef TKB_DW3
    ++g_launches;
    return;
  }
  dim3 grid((unsigned)((g.OW * (g.C / 16) + 255) / 256), (unsigned)g.OH, (unsigned)g.N);
  const uint8_t* xp = (const uint8_t*)x;
  const uint8_t* wp = (const uint8_t*)w_tap;
  const bool au = g.a_dtype == kU8, wu = g.w_dtype == kU8;
  if (au && wu) dwconv_nhwc_kernel<true, true><<<grid, 256, 0, s>>>(xp, wp, corr, y, g, e, f);
  else if (au) dwconv_nhwc_kernel<true, false><<<grid, 256, 0, s>>>(xp, wp, corr, y, g, e, f);
  else if (wu) dwconv_nhwc_kernel<false, true><<<grid, 256, 0, s>>>(xp, wp, corr, y, g, e, f);
  else dwconv_nhwc_kernel<false, false><<<grid, 256, 0, s>>>(xp, wp, corr, y, g, e, f);
  ++g_launches;
}

// GEMM fallback (qnn.dense dense.cc:160-205, qnn.batch_matmul batch_matmul.cc:161-200): one warp per
// output row-tile would be faster; this is the correctness path for odd K / non-zero weight zp.
template <bool AU, bool WU>
__global__ void gemm_simt_kernel(const uint8_t* __restrict__ x, const uint8_t* __restrict__ w, void* __restrict__ y,
                                 int B, int M, int N, int K, int w_batched, int32_t zpx, int32_t zpw, EpiParams e) {
  const long long total = (long long)B * M * N;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(idx % N);
    const long long bm = idx / N;
    const int b = (int)(bm / M);
    const uint8_t* ap = x + bm * K;
    const uint8_t* wp = w + ((long long)(w_batched ? b : 0) * N + n) * K;
    int raw = 0, sum_a = 0, sum_w = 0;
    int k = 0;
    if ((K & 3) == 0 && ((((uintptr_t)ap) | ((uintptr_t)wp)) & 3) == 0) {
      for (; k < K; k += 4) {
        const uint32_t av = *(const uint32_t*)(ap + k), wv = *(const uint32_t*)(wp + k);
        raw = dp4a_mixed<AU, WU>(av, wv, raw);
        sum_a = dp4a_mixed<AU, false>(av, 0x01010101u, sum_a);
        sum_w = dp4a_mixed<false, WU>(0x01010101u, wv, sum_w);
      }
    } else {
      for (; k < K; ++k) {
        const int av = AU ? (int)ap[k] : (int)(int8_t)ap[k];
        const int wv = WU ? (int)wp[k] : (int)(int8_t)wp[k];
        raw += av * wv;
        sum_a += av;
        sum_w += wv;
      }
    }
    const uint32_t acc = (uint32_t)raw - (uint32_t)zpx * (uint32_t)sum_w - (uint32_t)zpw * (uint32_t)sum_a +
                         (uint32_t)K * (uint32_t)zpx * (uint32_t)zpw;
    store_from_i32(y, e.out_dtype, idx, epilogue_one((int32_t)acc, n, idx, e));
  }
}

void launch_gemm_simt(const void* x, const void* w, void* y, int B, int M, int N, int K, int w_batched,
                      int x_dtype, int w_dtype, int32_t zpx, int32_t zpw, const EpiParams& e, cudaStream_t s) {
  const long long total = (long long)B * M * N;
  const int th = 256, bl = blocks_for(total, th);
  const uint8_t* xp = (const uint8_t*)x;
  const uint8_t* wp = (const uint8_t*)w;
  const bool au = x_dtype == kU8, wu = w_dtype == kU8;
  if (au && wu) gemm_simt_kernel<true, true><<<bl, th, 0, s>>>(xp, wp, y, B, M, N, K, w_batched, zpx, zpw, e);
  else if (au) gemm_simt_kernel<true, false><<<bl, th, 0, s>>>(xp, wp, y, B, M, N, K, w_batched, zpx, zpw, e);
  else if (wu) gemm_simt_kernel<false, true><<<bl, th, 0, s>>>(xp, wp, y, B, M, N, K, w_batched, zpx, zpw, e);
  else gemm_simt_kernel<false, false><<<bl, th, 0, s>>>(xp, wp, y, B, M, N, K, w_batched, zpx, zpw, e);
  ++g_launches;
}

__global__ void requantize_kernel(const void* __restrict__ x, int x_dtype, void* __restrict__ y, int y_dtype,
                                  long long total, int C, long long inner, RqParams p) {
  pdl_enter();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)((i / inner) % C);
    int32_t v;
    if (x_dtype == kI64) v = (int32_t)((const long long*)x)[i];
    else v = load_as_i32(x, x_dtype, i);
    store_from_i32(y, y_dtype, i, requantize_one(v, c, p));
  }
}

void launch_requantize(const void* x, int x_dtype, void* y, int y_dtype, long long outer, int C, long long inner,
                       const RqParams& p, cudaStream_t s) {
  const long long total = outer * C * inner;
  if (total == 0) return;
  launch_pdl(requantize_kernel, dim3(blocks_for(total, 256)), dim3(256), 0, s, x, x_dtype, y, y_dtype, total, C, inner, p);
  ++g_launches;
}

// qnn.add, HBM-bound: 16 bytes per thread per operand for the 8-bit dtypes.
__global__ void qnn_add_kernel(const void* __restrict__ a, const void* __restrict__ b, void* __restrict__ y,
                               int dtype, long long n, EpiParams e) {
  pdl_enter();
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  if ((dtype == kU8 || dtype == kI8) && (n & 15) == 0 &&
      ((((uintptr_t)a) | ((uintptr_t)b) | ((uintptr_t)y)) & 15) == 0) {
    const long long nv = n >> 4;
    for (long long i = tid; i < nv; i += stride) {
      const uint4 va = ((const uint4*)a)[i], vb = ((const uint4*)b)[i];
      uint4 vo;
      const uint32_t* pa = (const uint32_t*)&va;
      const uint32_t* pb = (const uint32_t*)&vb;
      uint32_t* po = (uint32_t*)&vo;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        uint32_t o = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint32_t ba = (pa[w] >> (8 * k)) & 0xFF, bb = (pb[w] >> (8 * k)) & 0xFF;
          const int32_t xa = dtype == kU8 ? (int32_t)ba : (int32_t)(int8_t)ba;
          const int32_t xb = dtype == kU8 ? (int32_t)bb : (int32_t)(int8_t)bb;
          o |= ((uint32_t)qnn_add_one(xa, xb, 0, e) & 0xFF) << (8 * k);
        }
        po[w] = o;
      }
      ((uint4*)y)[i] = vo;
    }
    return;
  }
  for (long long i = tid; i < n; i += stride)
    store_from_i32(y, dtype, i, qnn_add_one(load_as_i32(a, dtype, i), load_as_i32(b, dtype, i), 0, e));
}

// 8-bit qnn.add through two 256-entry tables: both operands are bytes, so each side's requantize (add.cc:40-93,
// the exact requantize_one) is tabulated once per block in shared memory and the per-element work is two lookups,
// one ring add and the clamp -- the kernel is then HBM-bound instead of bound by two int64 fixed-point multiplies.
__global__ void __launch_bounds__(256) qnn_add_lut_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b,
                                                          uint4* __restrict__ y, int dtype, long long nv, EpiParams e) {
  pdl_enter();
  __shared__ int32_t tl[256], tr[256];
  {
    const int i = threadIdx.x;
    const int32_t v = dtype == kU8 ? i : (int32_t)(int8_t)i;
    tl[i] = e.lhs_identity ? v : requantize_one(v, 0, e.lhs);
    const int32_t qb = e.rhs_identity ? v : requantize_one(v, 0, e.rhs);
    tr[i] = e.sum_sub ? (int32_t)((uint32_t)e.sum_zp_out - (uint32_t)qb) : (int32_t)((uint32_t)qb - (uint32_t)e.sum_zp_out);
  }
  __syncthreads();
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < nv; i += (long long)gridDim.x * 256) {
    const uint4 va = __ldg(a + i), vb = __ldg(b + i);
    const uint32_t pa[4] = {va.x, va.y, va.z, va.w}, pb[4] = {vb.x, vb.y, vb.z, vb.w};
    uint32_t po[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      uint32_t o = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int32_t s = (int32_t)((uint32_t)tl[(pa[w] >> (8 * k)) & 0xFF] + (uint32_t)tr[(pb[w] >> (8 * k)) & 0xFF]);
        o |= ((uint32_t)min(max(s, e.sum_qmin), e.sum_qmax) & 0xFF) << (8 * k);
      }
      po[w] = o;
    }
    y[i] = make_uint4(po[0], po[1], po[2], po[3]);
  }
}

void launch_qnn_add(const void* a, const void* b, void* y, int dtype, long long n, const EpiParams& e,
                    cudaStream_t s) {
  if (n == 0) return;
  if ((dtype == kU8 || dtype == kI8) && (n & 15) == 0 && n >= 4096 &&
      ((((uintptr_t)a) | ((uintptr_t)b) | ((uintptr_t)y)) & 15) == 0) {
    const long long nv = n >> 4;
    const int nb = (int)std::min<long long>((nv + 255) / 256, 148LL * 8);
    launch_pdl(qnn_add_lut_kernel, dim3(nb), dim3(256), 0, s, (const uint4*)a, (const uint4*)b, (uint4*)y, dtype, nv, e);
    ++g_launches;
    return;
  }
  const long long work = (dtype == kU8 || dtype == kI8) && (n & 15) == 0 ? n >> 4 : n;
  launch_pdl(qnn_add_kernel, dim3(blocks_for(work, 256)), dim3(256), 0, s, a, b, y, dtype, n, e);
  ++g_launches;
}

// ---- qnn.mul (src/relay/qnn/op/mul.cc:66-101, per-tensor form) ----------------------------------------------------
// Q' = (int32(a) - zp_a) * (int32(b) - zp_b) in the int32 ring, requantized from scale_a * scale_b (float32 product,
// computed on the host) to the output qparams with the default UPWARD / int64 lowering; output dtype = input dtype.
// 8-bit tensors: 16 elements per thread, 128-bit loads and stores.
__device__ __forceinline__ int32_t qnn_mul_one(int32_t a, int32_t b, int32_t zpa, int32_t zpb, const RqParams& rq) {
  const int32_t qa = (int32_t)((uint32_t)a - (uint32_t)zpa), qb = (int32_t)((uint32_t)b - (uint32_t)zpb);
  return requantize_one((int32_t)((uint32_t)qa * (uint32_t)qb), 0, rq);
}

__global__ void __launch_bounds__(256) qnn_mul8_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ y,
                                                       int dtype, long long nv, int32_t zpa, int32_t zpb, RqParams rq) {
  pdl_enter();
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < nv; i += (long long)gridDim.x * 256) {
    const uint4 va = __ldg(a + i), vb = __ldg(b + i);
    const uint32_t pa[4] = {va.x, va.y, va.z, va.w}, pb[4] = {vb.x, vb.y, vb.z, vb.w};
    uint32_t po[4];
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      uint32_t o = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t ba = (pa[w] >> (8 * k)) & 0xFF, bb = (pb[w] >> (8 * k)) & 0xFF;
        const int32_t xa = dtype == kU8 ? (int32_t)ba : (int32_t)(int8_t)ba, xb = dtype == kU8 ? (int32_t)bb : (int32_t)(int8_t)bb;
        o |= ((uint32_t)qnn_mul_one(xa, xb, zpa, zpb, rq) & 0xFF) << (8 * k);
      }
      po[w] = o;
    }
    y[i] = make_uint4(po[0], po[1], po[2], po[3]);
  }
}

__global__ void qnn_mul_kernel(const void* __restrict__ a, const void* __restrict__ b, void* __restrict__ y, int dtype,
                               long long n, int32_t zpa, int32_t zpb, RqParams rq) {
  pdl_enter();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    store_from_i32(y, dtype, i, qnn_mul_one(load_as_i32(a, dtype, i), load_as_i32(b, dtype, i), zpa, zpb, rq));
}

void launch_qnn_mul(const void* a, const void* b, void* y, int dtype, long long n, int32_t zpa, int32_t zpb,
                    const RqParams& rq, cudaStream_t s) {
  if (n == 0) return;
  if ((dtype == kU8 || dtype == kI8) && (n & 15) == 0 && ((((uintptr_t)a) | ((uintptr_t)b) | ((uintptr_t)y)) & 15) == 0) {
    const long long nv = n >> 4;
    const int nb = (int)std::min<long long>((nv + 255) / 256, 148LL * 8);
    launch_pdl(qnn_mul8_kernel, dim3(nb), dim3(256), 0, s, (const uint4*)a, (const uint4*)b, (uint4*)y, dtype, nv, zpa, zpb, rq);
  } else {
    launch_pdl(qnn_mul_kernel, dim3(blocks_for(n, 256)), dim3(256), 0, s, a, b, y, dtype, n, zpa, zpb, rq);
  }
  ++g_launches;
}

// ---- qnn.quantize / qnn.dequantize (src/relay/qnn/op/quantize.cc:113-150, dequantize.cc:95-127) ----------------------
// float32 arithmetic exactly as the reference lowers it: x / scale (IEEE division), tir.round = llvm.round (half away from
// zero), + float(zp), clip to the output dtype's range, cast; and float32(int32(x) - zp) * scale.  Layout
// [outer][C][inner] with per-tensor (n == 1) or per-channel (n == C) parameters; four elements per thread where the
// inner extent allows (128-bit float accesses, 32-bit packed 8-bit accesses).
struct QParams {
  const float* scale;
  const int32_t* zp;
  float scale0;
  int32_t zp0;
  int n_scale, n_zp;
};

__device__ __forceinline__ int32_t quantize_one(float x, float sc, int32_t zp, int32_t lo, int32_t hi) {
  float v = roundf(__fdiv_rn(x, sc));
  v = __fadd_rn(v, (float)zp);
  v = fminf(fmaxf(v, (float)lo), (float)hi);
  return (int32_t)v;
}

__global__ void quantize_kernel(const float* __restrict__ x, void* __restrict__ y, int y_dtype, long long total, int C,
                                long long inner, QParams q, int32_t lo, int32_t hi, int vec4) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  if (vec4) {  // inner % 4 == 0: the four elements share a channel
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < total / 4; v += stride) {
      const long long i = v * 4;
      const int c = (int)((i / inner) % C);
      const float sc = q.n_scale == 1 ? q.scale0 : q.scale[c];
      const int32_t zp = q.n_zp == 1 ? q.zp0 : q.zp[c];
      const float4 f = __ldg((const float4*)x + v);
      const int32_t r0 = quantize_one(f.x, sc, zp, lo, hi), r1 = quantize_one(f.y, sc, zp, lo, hi);
      const int32_t r2 = quantize_one(f.z, sc, zp, lo, hi), r3 = quantize_one(f.w, sc, zp, lo, hi);
      if (y_dtype == kU8 || y_dtype == kI8)
        ((uint32_t*)y)[v] = (uint32_t)(r0 & 0xFF) | ((uint32_t)(r1 & 0xFF) << 8) | ((uint32_t)(r2 & 0xFF) << 16) | ((uint32_t)(r3 & 0xFF) << 24);
      else if (y_dtype == kI32) ((int4*)y)[v] = make_int4(r0, r1, r2, r3);
      else { store_from_i32(y, y_dtype, i, r0); store_from_i32(y, y_dtype, i + 1, r1); store_from_i32(y, y_dtype, i + 2, r2); store_from_i32(y, y_dtype, i + 3, r3); }
    }
    return;
  }
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = (int)((i / inner) % C);
    store_from_i32(y, y_dtype, i, quantize_one(x[i], q.n_scale == 1 ? q.scale0 : q.scale[c], q.n_zp == 1 ? q.zp0 : q.zp[c], lo, hi));
  }
}

__global__ void dequantize_kernel(const void* __restrict__ x, int x_dtype, float* __restrict__ y, long long total, int C,
                                  long long inner, QParams q, int vec4) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  if (vec4) {
    for (long long v = blockIdx.x * (long long)blockDim.x + threadIdx.x; v < total / 4; v += stride) {
      const long long i = v * 4;
      const int c = (int)((i / inner) % C);
      const float sc = q.n_scale == 1 ? q.scale0 : q.scale[c];
      const uint32_t zp = (uint32_t)(q.n_zp == 1 ? q.zp0 : q.zp[c]);
      int32_t t[4];
      if (x_dtype == kU8 || x_dtype == kI8) {
        const uint32_t w = __ldg((const uint32_t*)x + v);
#pragma unroll
        for (int k = 0; k < 4; ++k) t[k] = x_dtype == kU8 ? (int32_t)((w >> (8 * k)) & 0xFF) : (int32_t)(int8_t)((w >> (8 * k)) & 0xFF);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) t[k] = load_as_i32(x, x_dtype, i + k);
      }
      ((float4*)y)[v] = make_float4(__fmul_rn((float)(int32_t)((uint32_t)t[0] - zp), sc), __fmul_rn((float)(int32_t)((uint32_t)t[1] - zp), sc),
                                    __fmul_rn((float)(int32_t)((uint32_t)t[2] - zp), sc), __fmul_rn((float)(int32_t)((uint32_t)t[3] - zp), sc));
    }
    return;
  }
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = (int)((i / inner) % C);
    const int32_t t = (int32_t)((uint32_t)load_as_i32(x, x_dtype, i) - (uint32_t)(q.n_zp == 1 ? q.zp0 : q.zp[c]));
    y[i] = __fmul_rn((float)t, q.n_scale == 1 ? q.scale0 : q.scale[c]);
  }
}

static QParams make_qparams(const float* scale, const int32_t* zp, float scale0, int32_t zp0, int n_scale, int n_zp) {
  QParams q;
  q.scale = scale; q.zp = zp; q.scale0 = scale0; q.zp0 = zp0; q.n_scale = n_scale; q.n_zp = n_zp;
  return q;
}

void launch_quantize(const float* x, void* y, int y_dtype, long long outer, int C, long long inner, const float* scale,
                     const int32_t* zp, float scale0, int32_t zp0, int n_scale, int n_zp, int32_t lo, int32_t hi, cudaStream_t s) {
  const long long total = outer * C * inner;
  if (total == 0) return;
  const int vec4 = (inner % 4 == 0) && ((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0;
  quantize_kernel<<<blocks_for(vec4 ? total / 4 : total, 256), 256, 0, s>>>(x, y, y_dtype, total, C, inner,
                                                                          make_qparams(scale, zp, scale0, zp0, n_scale, n_zp), lo, hi, vec4);
  ++g_launches;
}

void launch_dequantize(const void* x, int x_dtype, float* y, long long outer, int C, long long inner, const float* scale,
                       const int32_t* zp, float scale0, int32_t zp0, int n_scale, int n_zp, cudaStream_t s) {
  const long long total = outer * C * inner;
  if (total == 0) return;
  const int vec4 = (inner % 4 == 0) && ((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0;
  dequantize_kernel<<<blocks_for(vec4 ? total / 4 : total, 256), 256, 0, s>>>(x, x_dtype, y, total, C, inner,
                                                                            make_qparams(scale, zp, scale0, zp0, n_scale, n_zp), vec4);
  ++g_launches;
}

// ---- qnn.concatenate (src/relay/qnn/op/concatenate.cc:138-213): one piece [outer][Ci][inner] into [outer][Ctot][inner] at
// channel offset coff, requantized to the output qparams when its own differ (rq.skip == 0), same dtype.  16-byte copies
// when Ci * inner and the offsets allow.
__global__ void concat_piece_kernel(const void* __restrict__ x, void* __restrict__ y, int dtype, long long outer, long long row_in,
                                    long long row_out, long long off, RqParams rq, int requant, int vec16) {
  pdl_enter();
  const long long stride = (long long)gridDim.x * blockDim.x;
  if (vec16) {  // rows and offsets are multiples of 16 bytes, no requantize
    const long long rv = row_in * dtype_size(dtype) / 16, ov = row_out * dtype_size(dtype) / 16, offv = off * dtype_size(dtype) / 16;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < outer * rv; i += stride) {
      const long long o = i / rv, r = i - o * rv;
      ((uint4*)y)[o * ov + offv + r] = __ldg((const uint4*)x + i);
    }
    return;
  }
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < outer * row_in; i += stride) {
    const long long o = i / row_in, r = i - o * row_in;
    int32_t v = load_as_i32(x, dtype, i);
    if (requant) v = requantize_one(v, 0, rq);
    store_from_i32(y, dtype, o * row_out + off + r, v);
  }
}

void launch_concat_piece(const void* x, void* y, int dtype, long long outer, long long row_in, long long row_out, long long off,
                         const RqParams& rq, int requant, cudaStream_t s) {
  if (outer * row_in == 0) return;
  const int es = dtype_size(dtype);
  const int vec16 = !requant && (row_in * es) % 16 == 0 && (row_out * es) % 16 == 0 && (off * es) % 16 == 0 &&
                    ((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0;
  const long long work = vec16 ? outer * row_in * es / 16 : outer * row_in;
  launch_pdl(concat_piece_kernel, dim3(blocks_for(work, 256)), dim3(256), 0, s, x, y, dtype, outer, row_in, row_out, off, rq, requant, vec16);
  ++g_launches;
}

// take(table, indices) with 8-bit indices (bit pattern read as uint8) and a table of at most 256 entries of 1, 2 or 4
// bytes: the form the table-lookup unary QNN operators legalize to (python/tvm/relay/qnn/op/canonicalizations.py:106-160).
// HBM-bound: 16 indices per thread, the table staged in shared memory.
template <typename T>
__global__ void lut8_kernel(const uint8_t* __restrict__ x, const T* __restrict__ table, int entries, T* __restrict__ y,
                            long long n, int vec16) {
  pdl_enter();
  __shared__ T tab[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) tab[i] = i < entries ? table[i] : T(0);
  __syncthreads();
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
  if (vec16) {
    for (long long v = tid; v < n / 16; v += stride) {
      const uint4 in = __ldg((const uint4*)x + v);
      const uint32_t w[4] = {in.x, in.y, in.z, in.w};
      if constexpr (sizeof(T) == 1) {
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          o[k] = (uint32_t)tab[w[k] & 0xFF] | ((uint32_t)tab[(w[k] >> 8) & 0xFF] << 8) |
                 ((uint32_t)tab[(w[k] >> 16) & 0xFF] << 16) | ((uint32_t)tab[w[k] >> 24] << 24);
        ((uint4*)y)[v] = make_uint4(o[0], o[1], o[2], o[3]);
      } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) y[v * 16 + k] = tab[(w[k >> 2] >> (8 * (k & 3))) & 0xFF];
      }
    }
  } else {
    for (long long i = tid; i < n; i += stride) y[i] = tab[x[i]];
  }
}

void launch_lut8(const void* x, const void* table, int entries, int elem, void* y, long long n, cudaStream_t s) {
  if (n == 0) return;
  const int vec16 = (n % 16 == 0) && ((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0;
  const int bl = blocks_for(vec16 ? n / 16 : n, 256);
  if (elem == 1) launch_pdl(lut8_kernel<uint8_t>, dim3(bl), dim3(256), 0, s, (const uint8_t*)x, (const uint8_t*)table, entries, (uint8_t*)y, n, vec16);
  else if (elem == 2) launch_pdl(lut8_kernel<uint16_t>, dim3(bl), dim3(256), 0, s, (const uint8_t*)x, (const uint16_t*)table, entries, (uint16_t*)y, n, vec16);
  else launch_pdl(lut8_kernel<uint32_t>, dim3(bl), dim3(256), 0, s, (const uint8_t*)x, (const uint32_t*)table, entries, (uint32_t*)y, n, vec16);
  ++g_launches;
}

// qnn.leaky_relu (src/relay/qnn/op/leaky_relu.cc:70-139), element by element on any integer dtype up to 32 bits:
//   q = RequantizeOrUpcast(int32(x));  x < zp_in ? fixed_point_multiply(q, alpha) + fixed_point_multiply(zp_out, 1 - alpha) : q;
//   saturate to the dtype.  `rq.skip_mul`-style identity is passed as `identity`.
__global__ void qnn_leaky_relu_kernel(const void* __restrict__ x, void* __restrict__ y, int dtype, long long n, RqParams rq,
                                      int identity, int32_t zp_in, int32_t m_alpha, int32_t s_alpha, int32_t scaled_z,
                                      int32_t lo, int32_t hi) {
  pdl_enter();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int32_t v = load_as_i32(x, dtype, i);
    const int32_t q = identity ? v : requantize_one(v, 0, rq);
    int32_t r = q;
    if (v < zp_in) r = (int32_t)((uint32_t)fixed_point_multiply_upward(q, m_alpha, s_alpha) + (uint32_t)scaled_z);
    store_from_i32(y, dtype, i, min(max(r, lo), hi));
  }
}

void launch_qnn_leaky_relu(const void* x, void* y, int dtype, long long n, const RqParams& rq, int identity, int32_t zp_in,
                           int32_t m_alpha, int32_t s_alpha, int32_t scaled_z, int32_t lo, int32_t hi, cudaStream_t s) {
  if (n == 0) return;
  launch_pdl(qnn_leaky_relu_kernel, dim3(blocks_for(n, 256)), dim3(256), 0, s, x, y, dtype, n, rq, identity, zp_in, m_alpha, s_alpha, scaled_z, lo, hi);
  ++g_launches;
}

// fixed_point_multiply of ONE value on the device (used once at init for the leaky-relu zero-point term, so that host
// and device cannot disagree on the rounding)
__global__ void fixed_point_multiply_kernel(int32_t x, int32_t m, int32_t s, int32_t* out) { *out = fixed_point_multiply_upward(x, m, s); }
int32_t device_fixed_point_multiply(int32_t x, int32_t m, int32_t s) {
  int32_t* d = nullptr;
  int32_t h = 0;
  if (cudaMalloc(&d, 4) != cudaSuccess) return 0;
  fixed_point_multiply_kernel<<<1, 1>>>(x, m, s, d);
  cudaMemcpy(&h, d, 4, cudaMemcpyDeviceToHost);
  cudaFree(d);
  return h;
}

__global__ void clip_kernel(const void* __restrict__ x, void* __restrict__ y, int dtype, long long n, double lo,
                            double hi) {
  pdl_enter();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    if (dtype == kF32) {
      float v = ((const float*)x)[i];
      ((float*)y)[i] = fminf(fmaxf(v, (float)lo), (float)hi);
    } else {
      // relay clip on an integer tensor clamps in that dtype (bounds already intersected on the host)
      int32_t v = load_as_i32(x, dtype, i);
      const int32_t l = (int32_t)lo, h = (int32_t)hi;
      store_from_i32(y, dtype, i, v < l ? l : (v > h ? h : v));
    }
  }
}

void launch_clip(const void* x, void* y, int dtype, long long n, double lo, double hi, cudaStream_t s) {
  if (n == 0) return;
  launch_pdl(clip_kernel, dim3(blocks_for(n, 256)), dim3(256), 0, s, x, y, dtype, n, lo, hi);
  ++g_launches;
}

__global__ void cast_kernel(const void* __restrict__ x, int xd, void* __restrict__ y, int yd, long long n) {
  pdl_enter();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    if (xd == kF32) {
      float v = ((const float*)x)[i];
      if (yd == kF32) ((float*)y)[i] = v;
      else store_from_i32(y, yd, i, (int32_t)v);
    } else {
      int32_t v = load_as_i32(x, xd, i);
      if (yd == kF32) ((float*)y)[i] = (float)v;
      else store_from_i32(y, yd, i, v);  // truncating, like LLVM trunc / sext / zext
    }
  }
}

// 8-bit -> 32-bit integer widening, 4 elements per thread (32-bit load, 128-bit store)
template <bool SIGNED>
__global__ void cast_8_to_32_kernel(const uint32_t* __restrict__ x, int4* __restrict__ y, unsigned n4) {
  pdl_enter();
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  const uint32_t w = __ldg(x + i);
  if (SIGNED) y[i] = make_int4((int8_t)(w & 0xFF), (int8_t)((w >> 8) & 0xFF), (int8_t)((w >> 16) & 0xFF), (int8_t)(w >> 24));
  else y[i] = make_int4((int)(w & 0xFF), (int)((w >> 8) & 0xFF), (int)((w >> 16) & 0xFF), (int)(w >> 24));
}
// 32-bit -> 8-bit truncation (LLVM trunc), 4 elements per thread
__global__ void cast_32_to_8_kernel(const int4* __restrict__ x, uint32_t* __restrict__ y, unsigned n4) {
  pdl_enter();
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  const int4 v = __ldg(x + i);
  y[i] = ((uint32_t)v.x & 0xFF) | (((uint32_t)v.y & 0xFF) << 8) | (((uint32_t)v.z & 0xFF) << 16) | ((uint32_t)v.w << 24);
}

void launch_cast(const void* x, int x_dtype, void* y, int y_dtype, long long n, cudaStream_t s) {
  if (n == 0) return;
  const bool x8 = x_dtype == kU8 || x_dtype == kI8, y8 = y_dtype == kU8 || y_dtype == kI8;
  if ((n & 3) == 0 && n / 4 < (1LL << 31) && ((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0) {
    const unsigned n4 = (unsigned)(n / 4), nb = (n4 + 255) / 256;
    if (x8 && y_dtype == kI32) {
      if (x_dtype == kI8) launch_pdl(cast_8_to_32_kernel<true>, dim3(nb), dim3(256), 0, s, (const uint32_t*)x, (int4*)y, n4);
      else launch_pdl(cast_8_to_32_kernel<false>, dim3(nb), dim3(256), 0, s, (const uint32_t*)x, (int4*)y, n4);
      ++g_launches;
      return;
    }
    if (x_dtype == kI32 && y8) {
      launch_pdl(cast_32_to_8_kernel, dim3(nb), dim3(256), 0, s, (const int4*)x, (uint32_t*)y, n4);
      ++g_launches;
      return;
    }
  }
  launch_pdl(cast_kernel, dim3(blocks_for(n, 256)), dim3(256), 0, s, x, x_dtype, y, y_dtype, n);
  ++g_launches;
}

__global__ void bias_add_kernel(const void* __restrict__ x, const void* __restrict__ b, void* __restrict__ y,
                                int dtype, long long total, int C, long long inner) {
  pdl_enter();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)((i / inner) % C);
    const uint32_t v = (uint32_t)load_as_i32(x, dtype, i) + (uint32_t)load_as_i32(b, dtype, c);
    store_from_i32(y, dtype, i, (int32_t)v);
  }
}

void launch_bias_add(const void* x, const void* b, void* y, int dtype, long long outer, int C, long long inner,
                     cudaStream_t s) {
  const long long total = outer * C * inner;
  if (total == 0) return;
  launch_pdl(bias_add_kernel, dim3(blocks_for(total, 256)), dim3(256), 0, s, x, b, y, dtype, total, C, inner);
  ++g_launches;
}

__global__ void add_same_kernel(const void* __restrict__ a, const void* __restrict__ b, void* __restrict__ y,
                                int dtype, long long n) {
  pdl_enter();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    store_from_i32(y, dtype, i, (int32_t)((uint32_t)load_as_i32(a, dtype, i) + (uint32_t)load_as_i32(b, dtype, i)));
}

void launch_add_same(const void* a, const void* b, void* y, int dtype, long long n, cudaStream_t s) {
  if (n == 0) return;
  launch_pdl(add_same_kernel, dim3(blocks_for(n, 256)), dim3(256), 0, s, a, b, y, dtype, n);
  ++g_launches;
}

struct PoolArgs {
  int N, C, H, W, KH, KW, sh, sw, pt, pl, pb, pr, OH, OW, is_avg, cip, dtype;
  long long xs[4], ys[4];
};

// nn.max_pool2d / nn.avg_pool2d / global pooling on integer tensors (include/tvm/topi/nn/pooling.h).  The window sum is
// reduced IN THE TENSOR DTYPE (tvm::sum(temp(indices), daxis), :599): it wraps at 8 / 16 / 32 bits.  Windowed pooling
// divides by an int32 element count (:612-648; the 8/16-bit sum is promoted to int32 first); adaptive / global
// pooling (is_avg == 2) casts the divisor to the tensor dtype as well (:384-389).  Division truncates toward zero.
// Thread order follows the output's fastest-varying dimension so that stores coalesce for both
// channels-last and NCHW tensors.
__global__ void pool2d_kernel(const void* __restrict__ x, void* __restrict__ y, PoolArgs a) {
  pdl_enter();
  const long long total = (long long)a.N * a.C * a.OH * a.OW;
  const bool c_fast = a.ys[1] == 1;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    int n, c, oy, ox;
    long long t = idx;
    if (c_fast) { c = (int)(t % a.C); t /= a.C; ox = (int)(t % a.OW); t /= a.OW; oy = (int)(t % a.OH); n = (int)(t / a.OH); }
    else { ox = (int)(t % a.OW); t /= a.OW; oy = (int)(t % a.OH); t /= a.OH; c = (int)(t % a.C); n = (int)(t / a.C); }
    const int y0 = oy * a.sh - a.pt, x0 = ox * a.sw - a.pl;
    const int y1 = y0 + a.KH - 1, x1 = x0 + a.KW - 1;
    int32_t best = a.dtype == kU8 ? 0 : a.dtype == kI8 ? -128 : a.dtype == kI16 ? -32768 : INT32_MIN;
    uint32_t sum = 0;
    int cnt = 0;
    for (int iy = max(y0, 0); iy <= min(y1, a.H - 1); ++iy)
      for (int ix = max(x0, 0); ix <= min(x1, a.W - 1); ++ix) {
        const int32_t v = load_as_i32(x, a.dtype, n * a.xs[0] + c * a.xs[1] + iy * a.xs[2] + ix * a.xs[3]);
        best = max(best, v);
        sum = (uint32_t)cast_trunc((int32_t)(sum + (uint32_t)v), a.dtype);
        ++cnt;
      }
    int32_t r;
    if (!a.is_avg) r = best;
    else {
      int div;
      if (a.is_avg == 2) {  // adaptive pool: divisor multiplied up in the tensor dtype (host rejects a zero divisor)
        div = cast_trunc(cast_trunc(a.KH, a.dtype) * cast_trunc(a.KW, a.dtype), a.dtype);
      } else if (a.cip) {
        const int ye = min(y1, a.H + a.pb - 1), xe = min(x1, a.W + a.pr - 1);
        div = (ye - y0 + 1) * (xe - x0 + 1);
      } else div = cnt > 1 ? cnt : 1;
      r = (int32_t)sum / div;  // C division truncates toward zero
    }
    store_from_i32(y, a.dtype, n * a.ys[0] + c * a.ys[1] + oy * a.ys[2] + ox * a.ys[3], r);
  }
}

// 8-bit max pooling on channels-last tensors: one thread owns 16 consecutive channels of one output
// pixel (128-bit loads / stores, byte-wise SIMD max), threads of a warp cover consecutive channel groups.
template <bool UNSIGNED>
__global__ void maxpool8_nhwc_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ y, PoolArgs a) {
  pdl_enter();
  const int cg = a.C >> 4;
  const long long total = (long long)a.N * a.OH * a.OW * cg;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(idx % cg);
    long long t = idx / cg;
    const int ox = (int)(t % a.OW); t /= a.OW;
    const int oy = (int)(t % a.OH);
    const int n = (int)(t / a.OH);
    const int y0 = max(oy * a.sh - a.pt, 0), y1 = min(oy * a.sh - a.pt + a.KH - 1, a.H - 1);
    const int x0 = max(ox * a.sw - a.pl, 0), x1 = min(ox * a.sw - a.pl + a.KW - 1, a.W - 1);
    const uint32_t init = UNSIGNED ? 0u : 0x80808080u;
    uint4 m = make_uint4(init, init, init, init);
    for (int iy = y0; iy <= y1; ++iy) {
      const uint8_t* row = x + (((long long)n * a.H + iy) * a.W) * a.C + g * 16;
      for (int ix = x0; ix <= x1; ++ix) {
        const uint4 v = __ldg((const uint4*)(row + (long long)ix * a.C));
        if (UNSIGNED) { m.x = __vmaxu4(m.x, v.x); m.y = __vmaxu4(m.y, v.y); m.z = __vmaxu4(m.z, v.z); m.w = __vmaxu4(m.w, v.w); }
        else { m.x = __vmaxs4(m.x, v.x); m.y = __vmaxs4(m.y, v.y); m.z = __vmaxs4(m.z, v.z); m.w = __vmaxs4(m.w, v.w); }
      }
    }
    *(uint4*)(y + (((long long)n * a.OH + oy) * a.OW + ox) * a.C + g * 16) = m;
  }
}

// 3x3 / stride 2 specialisation (ResNet stem pool): one thread owns 16 channels of TWO vertically adjacent
// outputs (5 input rows x 3 columns, the middle row shared), no 64-bit index arithmetic (3-D grid), all 15
// 128-bit loads issued back to back.  Taps outside the image are replaced by a clamped in-window tap: max is
// idempotent, so re-reading a valid element changes nothing and no predication is needed.
template <bool UNSIGNED>
__global__ void __launch_bounds__(256) maxpool8_nhwc_3x3s2_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ y,
                                                                  PoolArgs a) {
  pdl_enter();
  const int cg = a.C >> 4;
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= a.OW * cg) return;
  const int ox = t / cg, g = t - ox * cg;
  const int oy = blockIdx.y * 2, n = blockIdx.z;
  const bool two = oy + 1 < a.OH;
  const uint8_t* img = x + (long long)n * a.H * a.W * a.C + g * 16;
  int cx[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) cx[j] = min(max(ox * 2 - a.pl + j, 0), a.W - 1) * a.C;
  const int ylo0 = max(oy * 2 - a.pt, 0), yhi0 = min(oy * 2 - a.pt + 2, a.H - 1);
  const int ylo1 = two ? max(oy * 2 + 2 - a.pt, 0) : ylo0, yhi1 = two ? min(oy * 2 + 4 - a.pt, a.H - 1) : yhi0;
  uint4 v[5][3];
#pragma unroll
  for (int r = 0; r < 5; ++r) {
    // rows 0..2 belong to output 0 (row 2 also to output 1), rows 3..4 to output 1; clamp each into its window
    int iy = oy * 2 - a.pt + r;
    iy = r < 2 ? min(max(iy, ylo0), yhi0) : (r == 2 ? min(max(iy, ylo0), yhi0) : min(max(iy, ylo1), yhi1));
    const uint8_t* row = img + (long long)iy * a.W * a.C;
#pragma unroll
    for (int j = 0; j < 3; ++j) v[r][j] = __ldg((const uint4*)(row + cx[j]));
  }
  auto vmax = [](uint4 p, uint4 q) {
    if (UNSIGNED) return make_uint4(__vmaxu4(p.x, q.x), __vmaxu4(p.y, q.y), __vmaxu4(p.z, q.z), __vmaxu4(p.w, q.w));
    return make_uint4(__vmaxs4(p.x, q.x), __vmaxs4(p.y, q.y), __vmaxs4(p.z, q.z), __vmaxs4(p.w, q.w));
  };
  uint4 rm[5];
#pragma unroll
  for (int r = 0; r < 5; ++r) rm[r] = vmax(vmax(v[r][0], v[r][1]), v[r][2]);
  uint8_t* out = y + (((long long)n * a.OH + oy) * a.OW + ox) * a.C + g * 16;
  *(uint4*)out = vmax(vmax(rm[0], rm[1]), rm[2]);
  // row 2 of the pair lies in output 1's window only when it is not clamped away (always true for pt <= 2)
  if (two) {
    const int iy2 = oy * 2 - a.pt + 2;
    uint4 m1 = vmax(rm[3], rm[4]);
    if (iy2 >= ylo1 && iy2 <= yhi1) m1 = vmax(m1, rm[2]);
    *(uint4*)(out + (long long)a.OW * a.C) = m1;
  }
}

// Global average pool of a dense channels-last int32 tensor: one thread per (n, c), the HW loads are independent
// (unrolled by 7) so several are in flight; sum wraps in int32 and the division truncates toward zero, as the
// reference's topi pool (include/tvm/topi/nn/pooling.h:389, :624-650).
// cast(8-bit -> int32) -> nn.global_avg_pool2d -> cast(int32 -> 8-bit), the frontend's apply_with_upcast form
// (python/tvm/relay/frontend/qnn_torch.py:587-590), in one launch: a thread owns 4 channels of one image, sums the
// H * W pixels in 32-bit wrap-around arithmetic, divides truncating (the int32 lowering, pooling.h:389) and writes the
// pooled int32 value and its truncating 8-bit cast.  The int32 copy of the input (the first cast's output) is written
// only when a trace wants it (cast_out != nullptr).  Same integers as the three separate kernels.
template <bool SIGNED>
__global__ void __launch_bounds__(128) upcast_gap_kernel(const uint32_t* __restrict__ x, int4* __restrict__ cast_out,
                                                         int4* __restrict__ pooled, uint32_t* __restrict__ out8, int HW, int C4) {
  pdl_enter();
  const int c = blockIdx.x * blockDim.x + threadIdx.x, n = blockIdx.y;
  if (c >= C4) return;
  const uint32_t* p = x + (long long)n * HW * C4 + c;
  int4* co = cast_out ? cast_out + (long long)n * HW * C4 + c : nullptr;
  uint32_t s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  auto acc = [&](uint32_t w, int i) {
    int4 v;
    if (SIGNED) v = make_int4((int8_t)(w & 0xFF), (int8_t)((w >> 8) & 0xFF), (int8_t)((w >> 16) & 0xFF), (int8_t)(w >> 24));
    else v = make_int4((int)(w & 0xFF), (int)((w >> 8) & 0xFF), (int)((w >> 16) & 0xFF), (int)(w >> 24));
    s0 += (uint32_t)v.x; s1 += (uint32_t)v.y; s2 += (uint32_t)v.z; s3 += (uint32_t)v.w;
    if (co) co[(long long)i * C4] = v;
  };
  int i = 0;
  for (; i + 7 <= HW; i += 7) {
    uint32_t w[7];
#pragma unroll
    for (int j = 0; j < 7; ++j) w[j] = __ldg(p + (long long)(i + j) * C4);
#pragma unroll
    for (int j = 0; j < 7; ++j) acc(w[j], i + j);
  }
  for (; i < HW; ++i) acc(__ldg(p + (long long)i * C4), i);
  const int4 q = make_int4((int32_t)s0 / HW, (int32_t)s1 / HW, (int32_t)s2 / HW, (int32_t)s3 / HW);
  pooled[(long long)n * C4 + c] = q;
  out8[(long long)n * C4 + c] = ((uint32_t)q.x & 0xFF) | (((uint32_t)q.y & 0xFF) << 8) | (((uint32_t)q.z & 0xFF) << 16) | ((uint32_t)q.w << 24);
}

void launch_upcast_gap(const void* x, int x_dtype, void* cast_out, void* pooled, void* out8, int N, int HW, int C, cudaStream_t s) {
  if (N <= 0 || HW <= 0 || C <= 0) return;
  const int C4 = C / 4;
  dim3 grid((C4 + 127) / 128, N);
  if (x_dtype == kI8) launch_pdl(upcast_gap_kernel<true>, grid, dim3(128), 0, s, (const uint32_t*)x, (int4*)cast_out, (int4*)pooled, (uint32_t*)out8, HW, C4);
  else launch_pdl(upcast_gap_kernel<false>, grid, dim3(128), 0, s, (const uint32_t*)x, (int4*)cast_out, (int4*)pooled, (uint32_t*)out8, HW, C4);
  ++g_launches;
}

__global__ void global_avgpool_nhwc_i32_kernel(const int32_t* __restrict__ x, int32_t* __restrict__ y, int HW, int C) {
  pdl_enter();
  const int c = blockIdx.x * blockDim.x + threadIdx.x, n = blockIdx.y;
  if (c >= C) return;
  const int32_t* p = x + (long long)n * HW * C + c;
  uint32_t sum = 0;
  int i = 0;
  for (; i + 7 <= HW; i += 7) {
    uint32_t v[7];
#pragma unroll
    for (int j = 0; j < 7; ++j) v[j] = (uint32_t)__ldg(p + (long long)(i + j) * C);
#pragma unroll
    for (int j = 0; j < 7; ++j) sum += v[j];
  }
  for (; i < HW; ++i) sum += (uint32_t)__ldg(p + (long long)i * C);
  y[(long long)n * C + c] = (int32_t)sum / HW;
}

void launch_pool2d(const void* x, void* y, int dtype, int is_avg, int count_include_pad, int N, int C, int H, int W,
                   int KH, int KW, int sh, int sw, int pt, int pl, int pb, int pr, const long long* xs,
                   const long long* ys, cudaStream_t s) {
  PoolArgs a;
  a.N = N; a.C = C; a.H = H; a.W = W; a.KH = KH; a.KW = KW; a.sh = sh; a.sw = sw;
  a.pt = pt; a.pl = pl; a.pb = pb; a.pr = pr; a.is_avg = is_avg; a.cip = count_include_pad; a.dtype = dtype;
  a.OH = (H + pt + pb - KH) / sh + 1;
  a.OW = (W + pl + pr - KW) / sw + 1;
  for (int i = 0; i < 4; ++i) { a.xs[i] = xs[i]; a.ys[i] = ys[i]; }
  const long long total = (long long)N * C * a.OH * a.OW;
  if (total == 0) return;
  const bool dense_nhwc = xs[1] == 1 && ys[1] == 1 && xs[3] == C && ys[3] == C && xs[2] == (long long)W * C &&
                          ys[2] == (long long)a.OW * C && xs[0] == (long long)H * W * C && ys[0] == (long long)a.OH * a.OW * C;
  if (!is_avg && (dtype == kU8 || dtype == kI8) && (C % 16) == 0 && dense_nhwc &&
      ((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0) {
    const long long work = total / 16;
    if (KH == 3 && KW == 3 && sh == 2 && sw == 2 && pt <= 2 && pl <= 2 && N <= 65535 && a.OH <= 131070 && H >= 2 && W >= 2) {
      dim3 grid((unsigned)((a.OW * (C / 16) + 255) / 256), (unsigned)((a.OH + 1) / 2), (unsigned)N);
      if (dtype == kU8) launch_pdl(maxpool8_nhwc_3x3s2_kernel<true>, grid, dim3(256), 0, s, (const uint8_t*)x, (uint8_t*)y, a);
      else launch_pdl(maxpool8_nhwc_3x3s2_kernel<false>, grid, dim3(256), 0, s, (const uint8_t*)x, (uint8_t*)y, a);
      ++g_launches;
      return;
    }
    if (dtype == kU8) launch_pdl(maxpool8_nhwc_kernel<true>, dim3(blocks_for(work, 256)), dim3(256), 0, s, (const uint8_t*)x, (uint8_t*)y, a);
    else launch_pdl(maxpool8_nhwc_kernel<false>, dim3(blocks_for(work, 256)), dim3(256), 0, s, (const uint8_t*)x, (uint8_t*)y, a);
    ++g_launches;
    return;
  }
  if (is_avg && dtype == kI32 && dense_nhwc && KH == H && KW == W && pt + pb + pl + pr == 0 && N <= 65535 && H * W > 0) {
    // int32: the adaptive (is_avg == 2) and the windowed divisor are both H * W
    dim3 grid((unsigned)((C + 127) / 128), (unsigned)N);
    launch_pdl(global_avgpool_nhwc_i32_kernel, grid, dim3(128), 0, s, (const int32_t*)x, (int32_t*)y, H * W, C);
    ++g_launches;
    return;
  }
  launch_pdl(pool2d_kernel, dim3(blocks_for(total, 256)), dim3(256), 0, s, x, y, a);
  ++g_launches;
}

// 8-bit [N][HW][C] (channels-last) -> [N][C][HW] transpose through a 64 x 64 shared-memory tile: 32-bit
// coalesced loads along C, 32-bit coalesced stores along HW.  Used for the logical-layout copies of the
// per-operator trace and for subgraph outputs.  Requires C % 4 == 0, HW % 4 == 0, 4-byte aligned bases.
__global__ void __launch_bounds__(256) nhwc_to_nchw8_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ y, int HW, int C) {
  __shared__ uint8_t tile[64][68];
  pdl_enter();
  const int n = blockIdx.z, hw0 = blockIdx.x * 64, c0 = blockIdx.y * 64;
  const uint8_t* xin = x + (long long)n * HW * C;
  uint8_t* yout = y + (long long)n * HW * C;
  const int t = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = t + i * 256;      // 0..1023: row = hw, 16 words along C
    const int r = idx >> 4, w = idx & 15;
    uint32_t v = 0;
    if (hw0 + r < HW && c0 + 4 * w < C) v = __ldg((const uint32_t*)(xin + (long long)(hw0 + r) * C + c0 + 4 * w));
    *(uint32_t*)&tile[r][4 * w] = v;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = t + i * 256;      // row = c, 16 words along HW
    const int c = idx >> 4, w = idx & 15;
    if (c0 + c < C && hw0 + 4 * w < HW) {
      const uint32_t v = (uint32_t)tile[4 * w][c] | ((uint32_t)tile[4 * w + 1][c] << 8) | ((uint32_t)tile[4 * w + 2][c] << 16) |
                         ((uint32_t)tile[4 * w + 3][c] << 24);
      *(uint32_t*)(yout + (long long)(c0 + c) * HW + hw0 + 4 * w) = v;
    }
  }
}

bool launch_nhwc_to_nchw8(const void* x, void* y, long long N, long long HW, long long C, cudaStream_t s) {
  if ((C & 3) || (HW & 3) || ((((uintptr_t)x) | ((uintptr_t)y)) & 3) || N > 65535 || (C + 63) / 64 > 65535) return false;
  if (N * HW * C == 0) return true;
  dim3 grid((unsigned)((HW + 63) / 64), (unsigned)((C + 63) / 64), (unsigned)N);
  launch_pdl(nhwc_to_nchw8_kernel, grid, dim3(256), 0, s, (const uint8_t*)x, (uint8_t*)y, (int)HW, (int)C);
  ++g_launches;
  return true;
}

// Generic [N][A][B] -> [N][B][A] transpose of 1/2/4-byte elements through a padded 32 x 32 tile (element-wise
// accesses, both sides coalesced).  Fallback of nhwc_to_nchw8 for odd spatial sizes and wider dtypes.
template <typename T>
__global__ void __launch_bounds__(256) transpose_tile_kernel(const T* __restrict__ x, T* __restrict__ y, int A, int B) {
  __shared__ T tile[32][33];
  pdl_enter();
  const long long base = (long long)blockIdx.z * A * B;
  const int a0 = blockIdx.x * 32, b0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int a = a0 + ty + 8 * i, b = b0 + tx;
    if (a < A && b < B) tile[ty + 8 * i][tx] = x[base + (long long)a * B + b];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int b = b0 + ty + 8 * i, a = a0 + tx;
    if (a < A && b < B) y[base + (long long)b * A + a] = tile[tx][ty + 8 * i];
  }
}

bool launch_transpose_tile(const void* x, void* y, int elem, long long N, long long A, long long B, cudaStream_t s) {
  if (N > 65535 || (B + 31) / 32 > 65535 || N * A * B == 0) return N * A * B == 0;
  dim3 grid((unsigned)((A + 31) / 32), (unsigned)((B + 31) / 32), (unsigned)N);
  if (elem == 1) launch_pdl(transpose_tile_kernel<uint8_t>, grid, dim3(256), 0, s, (const uint8_t*)x, (uint8_t*)y, (int)A, (int)B);
  else if (elem == 2) launch_pdl(transpose_tile_kernel<uint16_t>, grid, dim3(256), 0, s, (const uint16_t*)x, (uint16_t*)y, (int)A, (int)B);
  else if (elem == 4) launch_pdl(transpose_tile_kernel<uint32_t>, grid, dim3(256), 0, s, (const uint32_t*)x, (uint32_t*)y, (int)A, (int)B);
  else return false;
  ++g_launches;
  return true;
}

// Busy-wait on the device for about `us` microseconds (one thread).  bench.py queues it in front of a timed
// step so that the host enqueues the whole step while the GPU is still busy: the CUDA events recorded around
// every tensor-core launch then measure device time only, not host launch overhead.
__global__ void spin_kernel(long long ns) {
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1)); } while ((long long)(t1 - t0) < ns);
}
void launch_spin(double us, cudaStream_t s) {
  spin_kernel<<<1, 1, 0, s>>>((long long)(us * 1e3));
  ++g_launches;
}

struct Perm4 { long long shape[4], istr[4]; };

// Tiled transpose of the two dims that swap fastest-varying roles would be the speed-of-light
// version; graph-edge layout changes are small next to the convolutions, so this first version is a
// plain gather with coalesced stores.
template <typename T>
__global__ void permute4_kernel(const T* __restrict__ x, T* __restrict__ y, Perm4 p) {
  pdl_enter();
  const long long total = p.shape[0] * p.shape[1] * p.shape[2] * p.shape[3];
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    long long t = idx;
    const long long i3 = t % p.shape[3]; t /= p.shape[3];
    const long long i2 = t % p.shape[2]; t /= p.shape[2];
    const long long i1 = t % p.shape[1];
    const long long i0 = t / p.shape[1];
    y[idx] = x[i0 * p.istr[0] + i1 * p.istr[1] + i2 * p.istr[2] + i3 * p.istr[3]];
  }
}

void launch_permute4(const void* x, void* y, int elem, const long long* out_shape, const long long* in_strides,
                     cudaStream_t s) {
  Perm4 p;
  long long total = 1;
  for (int i = 0; i < 4; ++i) { p.shape[i] = out_shape[i]; p.istr[i] = in_strides[i]; total *= out_shape[i]; }
  if (total == 0) return;
  const int bl = blocks_for(total, 256);
  if (elem == 1) launch_pdl(permute4_kernel<uint8_t>, dim3(bl), dim3(256), 0, s, (const uint8_t*)x, (uint8_t*)y, p);
  else if (elem == 2) launch_pdl(permute4_kernel<uint16_t>, dim3(bl), dim3(256), 0, s, (const uint16_t*)x, (uint16_t*)y, p);
  else if (elem == 8) launch_pdl(permute4_kernel<uint64_t>, dim3(bl), dim3(256), 0, s, (const uint64_t*)x, (uint64_t*)y, p);
  else launch_pdl(permute4_kernel<uint32_t>, dim3(bl), dim3(256), 0, s, (const uint32_t*)x, (uint32_t*)y, p);
  ++g_launches;
}

struct Pad4 { long long ishape[4], before[4], oshape[4]; };

template <typename T>
__global__ void pad4_kernel(const T* __restrict__ x, T* __restrict__ y, Pad4 p, T value) {
  const long long total = p.oshape[0] * p.oshape[1] * p.oshape[2] * p.oshape[3];
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    long long t = idx, i[4];
    for (int d = 3; d >= 0; --d) { i[d] = t % p.oshape[d] - p.before[d]; t /= p.oshape[d]; }
    bool in = true;
    for (int d = 0; d < 4; ++d) in = in && i[d] >= 0 && i[d] < p.ishape[d];
    y[idx] = in ? x[((i[0] * p.ishape[1] + i[1]) * p.ishape[2] + i[2]) * p.ishape[3] + i[3]] : value;
  }
}

void launch_pad4(const void* x, void* y, int elem, const long long* in_shape, const long long* pad_before,
                 const long long* out_shape, int32_t value, cudaStream_t s) {
  Pad4 p;
  long long total = 1;
  for (int i = 0; i < 4; ++i) {
    p.ishape[i] = in_shape[i]; p.before[i] = pad_before[i]; p.oshape[i] = out_shape[i];
    total *= out_shape[i];
  }
  if (total == 0) return;
  const int bl = blocks_for(total, 256);
  if (elem == 1) pad4_kernel<uint8_t><<<bl, 256, 0, s>>>((const uint8_t*)x, (uint8_t*)y, p, (uint8_t)value);
  else if (elem == 2) pad4_kernel<uint16_t><<<bl, 256, 0, s>>>((const uint16_t*)x, (uint16_t*)y, p, (uint16_t)value);
  else pad4_kernel<uint32_t><<<bl, 256, 0, s>>>((const uint32_t*)x, (uint32_t*)y, p, (uint32_t)value);
  ++g_launches;
}

// Explicit im2col for the tensor-core GEMM path: A[m][(r*KW+s)*C + c] = x[n, iy, ix, c]; taps that fall
// outside the image read the INPUT ZERO POINT (convolution.cc:227-253 pads with zp_a), which makes the
// zero-point correction  acc = raw - zp_a * sum_k W[o,k]  uniform over the whole output.  Columns
// beyond KH*KW*C (K padded to a multiple of 16 for TMA) are zero and meet zero weights.
__global__ void im2col_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ a, ConvGeom g, int Kp) {
  const int K = g.KH * g.KW * g.C;
  const long long M = (long long)g.N * g.OH * g.OW;
  if ((g.C & 15) == 0) {
    const int chunks = Kp >> 4, cpc = g.C >> 4;
    const long long total = M * chunks;
    const uint32_t z = (uint32_t)(g.zp_a & 0xFF) * 0x01010101u;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
      const int ch = (int)(idx % chunks);
      long long m = idx / chunks;
      const int tap = ch / cpc, c16 = ch % cpc;
      const int r = tap / g.KW, s = tap % g.KW;
      const int ox = (int)(m % g.OW);
      long long t = m / g.OW;
      const int oy = (int)(t % g.OH);
      const int n = (int)(t / g.OH);
      const int iy = oy * g.sh + r * g.dh - g.pt, ix = ox * g.sw + s * g.dw - g.pl;
      uint4 v = make_uint4(z, z, z, z);
      if (iy >= 0 && iy < g.H && ix >= 0 && ix < g.W)
        v = *(const uint4*)(x + (((long long)n * g.H + iy) * g.W + ix) * g.C + c16 * 16);
      *(uint4*)(a + m * Kp + ch * 16) = v;
    }
  } else {
    const long long total = M * Kp;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
      const int k = (int)(idx % Kp);
      long long m = idx / Kp;
      uint8_t v = 0;
      if (k < K) {
        const int tap = k / g.C, c = k % g.C;
        const int r = tap / g.KW, s = tap % g.KW;
        const int ox = (int)(m % g.OW);
        long long t = m / g.OW;
        const int oy = (int)(t % g.OH);
        const int n = (int)(t / g.OH);
        const int iy = oy * g.sh + r * g.dh - g.pt, ix = ox * g.sw + s * g.dw - g.pl;
        v = (uint8_t)g.zp_a;
        if (iy >= 0 && iy < g.H && ix >= 0 && ix < g.W) v = x[(((long long)n * g.H + iy) * g.W + ix) * g.C + c];
      }
      a[idx] = v;
    }
  }
}

// im2col straight from a logical NCHW tensor (few input channels, e.g. the 3-channel network stem):
// A[m][(c*KH + r)*KW + s] = x[n, c, iy, ix], i.e. K is ordered like OIHW weights.  One thread produces 16
// consecutive bytes of a row, so a warp writes 512 contiguous bytes; the (c, r, s) decomposition of k comes
// from a small shared-memory table.  Out-of-image taps read the input zero point.
__global__ void im2col_nchw_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ a, ConvGeom g, int Kp) {
  extern __shared__ uint32_t ktab[];  // [Kp] c << 16 | r << 8 | s, 0xFFFFFFFF for the zero tail
  const int K = g.C * g.KH * g.KW;
  for (int k = threadIdx.x; k < Kp; k += blockDim.x) {
    const int c = k / (g.KH * g.KW), rs = k % (g.KH * g.KW);
    ktab[k] = k < K ? ((uint32_t)c << 16) | ((uint32_t)(rs / g.KW) << 8) | (uint32_t)(rs % g.KW) : 0xFFFFFFFFu;
  }
  __syncthreads();
  const int chunks = Kp >> 4;
  const long long M = (long long)g.N * g.OH * g.OW, total = M * chunks;
  const long long plane = (long long)g.H * g.W;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int ch = (int)(idx % chunks);
    const long long m = idx / chunks;
    const int ox = (int)(m % g.OW);
    const long long t = m / g.OW;
    const int oy = (int)(t % g.OH);
    const long long n = t / g.OH;
    const int iy0 = oy * g.sh - g.pt, ix0 = ox * g.sw - g.pl;
    const uint8_t* img = x + n * g.C * plane;
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t word = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const uint32_t e = ktab[ch * 16 + q * 4 + b];
        uint32_t v = 0;
        if (e != 0xFFFFFFFFu) {
          const int iy = iy0 + (int)((e >> 8) & 0xFF) * g.dh, ix = ix0 + (int)(e & 0xFF) * g.dw;
          v = (uint32_t)(g.zp_a & 0xFF);
          if (iy >= 0 && iy < g.H && ix >= 0 && ix < g.W) v = __ldg(img + (long long)(e >> 16) * plane + (long long)iy * g.W + ix);
        }
        word |= v << (8 * b);
      }
      w[q] = word;
    }
    *(uint4*)(a + m * Kp + ch * 16) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

void launch_im2col_nchw(const void* x, void* a, const ConvGeom& g, int Kp, cudaStream_t s) {
  const long long total = (long long)g.N * g.OH * g.OW * (Kp >> 4);
  if (total == 0) return;
  im2col_nchw_kernel<<<blocks_for(total, 256), 256, Kp * 4, s>>>((const uint8_t*)x, (uint8_t*)a, g, Kp);
  ++g_launches;
}

// "Vertical unfold" of a few-channel NCHW tensor (network stem):
//   X3[n][p][pl + w][r*C + c] = x[n, c, p*sh - pt + r*dh, w]      (CP = 32 bytes per pixel)
// for the KH vertical taps of output row p; rows are physically padded with pl / pr extra pixels.  Taps and
// pixels outside the image hold the input zero point (exactly the reference's pad-with-zp,
// convolution.cc:227-253), bytes beyond KH*C are zero.  The im2col row of output pixel (p, q) is then the
// CONTIGUOUS window of KW pixels starting at pixel q*sw of row (n, p): a tiled tensor map with an overlapping
// row stride of sw*CP bytes presents X3 as the [M][KW*CP] GEMM operand without materialising it.
// One thread per (n, p, padded w): loads coalesce along w, the CP bytes are stored as 16-byte vectors.
template <int C_, int KH_>
__global__ void unfold_rows_nchw_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ y, ConvGeom g) {
  constexpr int CP = 32;
  const int C = C_ ? C_ : g.C, KH = KH_ ? KH_ : g.KH;
  const int WP = g.W + g.pl + g.pr;
  const long long total = (long long)g.N * g.OH * WP;
  const long long plane = (long long)g.H * g.W;
  const uint32_t zp = (uint32_t)(g.zp_a & 0xFF);
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int wp = (int)(idx % WP);
    const long long t = idx / WP;
    const int p = (int)(t % g.OH);
    const long long n = t / g.OH;
    const int w = wp - g.pl;
    const bool win = w >= 0 && w < g.W;
    const uint8_t* img = x + n * g.C * plane + (win ? w : 0);
    const int iy0 = p * g.sh - g.pt;
    uint32_t words[CP / 4];
#pragma unroll
    for (int i = 0; i < CP / 4; ++i) words[i] = 0;
    if (C_ && KH_) {
#pragma unroll
      for (int r = 0; r < KH_; ++r) {
        const int iy = iy0 + r * g.dh;
        const bool in = win && iy >= 0 && iy < g.H;
#pragma unroll
        for (int c = 0; c < C_; ++c) {
          const int k = r * C_ + c;
          const uint32_t v = in ? (uint32_t)__ldg(img + c * plane + (long long)iy * g.W) : zp;
          words[k >> 2] |= v << (8 * (k & 3));
        }
      }
    } else {
      int k = 0;
      for (int r = 0; r < KH; ++r) {
        const int iy = iy0 + r * g.dh;
        const bool in = win && iy >= 0 && iy < g.H;
        for (int c = 0; c < C; ++c, ++k) {
          const uint32_t v = in ? (uint32_t)__ldg(img + c * plane + (long long)iy * g.W) : zp;
#pragma unroll
          for (int i = 0; i < CP / 4; ++i)
            if ((k >> 2) == i) words[i] |= v << (8 * (k & 3));
        }
      }
    }
    uint4* dst = (uint4*)(y + idx * CP);
    dst[0] = make_uint4(words[0], words[1], words[2], words[3]);
    dst[1] = make_uint4(words[4], words[5], words[6], words[7]);
  }
}

// Vectorised variant for W % 4 == 0: one warp produces 128 consecutive interior pixels of one (n, p) row.
// Lane l loads, for each of the KH*C (row, channel) lines, the 32-bit word holding input columns w0+4l .. +3
// (128 contiguous bytes per warp load), transposes bytes in registers (4x4 byte transposes with PRMT) into its
// four 32-byte pixels, and the warp then re-orders the 4 KB through a padded shared-memory slab so that every
// global store instruction writes 512 contiguous bytes.  The pl + pr padding pixels of the row are written by
// the first lanes of segment 0.
template <int C_, int KH_>
__global__ void __launch_bounds__(256) unfold_rows_nchw_vec_kernel(const uint8_t* __restrict__ x, uint8_t* __restrict__ y,
                                                                   ConvGeom g) {
  pdl_enter();
  constexpr int CP = 32, KC = KH_ * C_, NG = (KC + 3) / 4;  // NG groups of 4 source lines -> output word NG-1 max
  __shared__ uint4 slab[8][32 * 9];                          // per warp: 32 lanes x (128 + 16) bytes
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int segs = (g.W + 127) / 128;
  const int WP = g.W + g.pl + g.pr;
  const long long plane = (long long)g.H * g.W;
  const uint32_t zp = (uint32_t)(g.zp_a & 0xFF), zp4 = zp * 0x01010101u;
  const long long nwarps = (long long)g.N * g.OH * segs;
  for (long long wi = blockIdx.x * 8LL + warp; wi < nwarps; wi += gridDim.x * 8LL) {
    const int seg = (int)(wi % segs);
    const long long t = wi / segs;
    const int p = (int)(t % g.OH);
    const long long n = t / g.OH;
    const int w0 = seg * 128, w = w0 + 4 * lane;
    const bool act = w < g.W;
    const uint8_t* img = x + n * C_ * plane + w;
    const int iy0 = p * g.sh - g.pt;
    uint32_t L[NG * 4];
#pragma unroll
    for (int k = 0; k < NG * 4; ++k) {
      uint32_t v = 0;
      if (k < KC) {
        const int r = k / C_, c = k % C_;
        const int iy = iy0 + r * g.dh;
        v = zp4;
        if (act && iy >= 0 && iy < g.H) v = __ldg((const uint32_t*)(img + c * plane + (long long)iy * g.W));
      }
      L[k] = v;
    }
    // 4x4 byte transposes: out[j][i] = bytes j of L[4i .. 4i+3]
    uint32_t out[4][8];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int i = 0; i < 8; ++i) out[j][i] = 0;
#pragma unroll
    for (int i = 0; i < NG; ++i) {
      const uint32_t a = __byte_perm(L[4 * i], L[4 * i + 1], 0x5140), b = __byte_perm(L[4 * i], L[4 * i + 1], 0x7362);
      const uint32_t c2 = __byte_perm(L[4 * i + 2], L[4 * i + 3], 0x5140), d = __byte_perm(L[4 * i + 2], L[4 * i + 3], 0x7362);
      out[0][i] = __byte_perm(a, c2, 0x5410);
      out[1][i] = __byte_perm(a, c2, 0x7632);
      out[2][i] = __byte_perm(b, d, 0x5410);
      out[3][i] = __byte_perm(b, d, 0x7632);
    }
    uint4* mine = &slab[warp][lane * 9];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      mine[2 * j] = make_uint4(out[j][0], out[j][1], out[j][2], out[j][3]);
      mine[2 * j + 1] = make_uint4(out[j][4], out[j][5], out[j][6], out[j][7]);
    }
    __syncwarp();
    uint8_t* row = y + ((n * g.OH + p) * (long long)WP) * CP;
    uint4* dst = (uint4*)(row + (long long)(g.pl + w0) * CP);
    const int npx = min(128, g.W - w0);  // pixels of this segment
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int chunk = 32 * i + lane;   // 16-byte chunk of the segment's contiguous 4 KB
      if (chunk < npx * 2) dst[chunk] = slab[warp][(chunk >> 3) * 9 + (chunk & 7)];
    }
    if (seg == 0 && lane < g.pl + g.pr) {  // padding pixels: zero-point bytes in the first KH*C positions
      uint32_t wz[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        uint32_t v = 0;
#pragma unroll
        for (int bb = 0; bb < 4; ++bb)
          if (4 * i + bb < KC) v |= zp << (8 * bb);
        wz[i] = v;
      }
      const int px = lane < g.pl ? lane : g.pl + g.W + (lane - g.pl);
      uint4* pd = (uint4*)(row + (long long)px * CP);
      pd[0] = make_uint4(wz[0], wz[1], wz[2], wz[3]);
      pd[1] = make_uint4(wz[4], wz[5], wz[6], wz[7]);
    }
    __syncwarp();
  }
}

void launch_unfold_rows_nchw(const void* x, void* y, const ConvGeom& g, cudaStream_t s) {
  const long long total = (long long)g.N * g.OH * (g.W + g.pl + g.pr);
  if (total == 0) return;
  const int nb = blocks_for(total, 256);
  const uint8_t* xi = (const uint8_t*)x;
  uint8_t* yo = (uint8_t*)y;
  if ((g.W & 3) == 0 && (((uintptr_t)x) & 3) == 0 && (((uintptr_t)y) & 15) == 0 && g.pl + g.pr <= 32) {
    const long long nwarps = (long long)g.N * g.OH * ((g.W + 127) / 128);
    const int nbv = (int)std::min<long long>((nwarps + 7) / 8, 148LL * 64);
    bool done = true;
    if (g.C == 3 && g.KH == 7) launch_pdl(unfold_rows_nchw_vec_kernel<3, 7>, dim3(nbv), dim3(256), 0, s, xi, yo, g);
    else if (g.C == 3 && g.KH == 3) launch_pdl(unfold_rows_nchw_vec_kernel<3, 3>, dim3(nbv), dim3(256), 0, s, xi, yo, g);
    else if (g.C == 1 && g.KH == 5) launch_pdl(unfold_rows_nchw_vec_kernel<1, 5>, dim3(nbv), dim3(256), 0, s, xi, yo, g);
    else if (g.C == 1 && g.KH == 3) launch_pdl(unfold_rows_nchw_vec_kernel<1, 3>, dim3(nbv), dim3(256), 0, s, xi, yo, g);
    else done = false;
    if (done) { ++g_launches; return; }
  }
  if (g.C == 3 && g.KH == 7) unfold_rows_nchw_kernel<3, 7><<<nb, 256, 0, s>>>(xi, yo, g);
  else if (g.C == 3 && g.KH == 3) unfold_rows_nchw_kernel<3, 3><<<nb, 256, 0, s>>>(xi, yo, g);
  else if (g.C == 1 && g.KH == 5) unfold_rows_nchw_kernel<1, 5><<<nb, 256, 0, s>>>(xi, yo, g);
  else if (g.C == 1 && g.KH == 3) unfold_rows_nchw_kernel<1, 3><<<nb, 256, 0, s>>>(xi, yo, g);
  else unfold_rows_nchw_kernel<0, 0><<<nb, 256, 0, s>>>(xi, yo, g);
  ++g_launches;
}

// 1x1 / stride-1 / unpadded convolution whose channel count is not a multiple of 16 (e.g. 24): the im2col matrix is
// the activation matrix with every row zero-padded from C to Kp bytes; 8 bytes per thread, 32-bit indices.
__global__ void pad_rows8_kernel(const uint2* __restrict__ x, uint2* __restrict__ a, unsigned total, int c8, int k8) {
  pdl_enter();
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const unsigned m = i / (unsigned)k8, ch = i - m * (unsigned)k8;
  a[i] = ch < (unsigned)c8 ? __ldg(x + (size_t)m * c8 + ch) : make_uint2(0u, 0u);
}

void launch_im2col(const void* x, void* a, const ConvGeom& g, int Kp, cudaStream_t s) {
  const long long M = (long long)g.N * g.OH * g.OW;
  if (g.KH == 1 && g.KW == 1 && g.sh == 1 && g.sw == 1 && !(g.pt | g.pl | g.pb | g.pr) && (g.C & 15) != 0 && (g.C & 7) == 0 &&
      (Kp & 7) == 0 && M * (Kp >> 3) < (1LL << 31) && M > 0 && ((((uintptr_t)x) | ((uintptr_t)a)) & 7) == 0) {
    const unsigned total = (unsigned)(M * (Kp >> 3));
    launch_pdl(pad_rows8_kernel, dim3((total + 255) / 256), dim3(256), 0, s, (const uint2*)x, (uint2*)a, total, g.C >> 3, Kp >> 3);
    ++g_launches;
    return;
  }
  const long long total = (g.C & 15) == 0 ? M * (Kp >> 4) : M * Kp;
  if (total == 0) return;
  im2col_kernel<<<blocks_for(total, 256), 256, 0, s>>>((const uint8_t*)x, (uint8_t*)a, g, Kp);
  ++g_launches;
}

}  // namespace tkb
