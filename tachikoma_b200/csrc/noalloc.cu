// tkb_qnn_conv2d_noalloc: one qnn.conv2d without any allocation (include/tkb200.h).
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops).  It is the runtime's explicit-im2col convolution
// (runtime.cu plan_conv, "tc_explicit": column matrix in the workspace arena, then a GEMM) with the column matrix in
// a file-scope static __device__ array instead of the arena.  Arithmetic: src/relay/qnn/op/convolution.cc
// (out = sum (x - zp_x)(w - zp_w) in int32; out-of-image taps hold zp_x, i.e. contribute 0).
//
// The array's name is only a host-side stand-in in host code: the device address comes from cudaGetSymbolAddress,
// once per device with that device current, and is what the kernels receive.
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>
#include <string>

#include "../../include/tkb200.h"
#include "kernels.h"

namespace tkb {
void set_last_error(const std::string& msg);  // runtime.cu (thread-local message behind tkb_last_error)
}

// File scope.  Shared by every call on a device; holds the last call's column matrix until the next call overwrites it.
static __device__ __align__(256) uint8_t g_noalloc_scratch[TKB_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_mu;
uint8_t* g_scratch_ptr[kMaxDevices] = {};

// Device address of the scratch array on the CURRENT device (looked up once per device).
uint8_t* scratch_for_current_device(std::string* err) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    *err = std::string("tkb_qnn_conv2d_noalloc: cudaGetDevice failed: ") + cudaGetErrorString(e);
    return nullptr;
  }
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_scratch_ptr[dev]) {
    uint8_t* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      *err = std::string("tkb_qnn_conv2d_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e);
      return nullptr;
    }
    g_scratch_ptr[dev] = p;
  }
  return g_scratch_ptr[dev];
}

// out[n][o][p] (NCHW, p = oh*OW + ow) = sum_{k < K} (col[m][k] - zpx) * (w[o][k] - zpw), m = n*P + p.
// col rows have a pitch of Kp bytes (bytes K..Kp-1 are not read), weight rows a pitch of K bytes.
// 64 x 64 output tile per block of 256 threads, 32-deep k slices staged in shared memory as zero-point-corrected
// ints; thread (tx, ty) owns pixels tx + 16*i and channels ty*4 + j, so stores along p coalesce.
template <bool XS, bool WS>
__global__ void __launch_bounds__(256) conv_gemm_nchw_kernel(const uint8_t* __restrict__ col, const uint8_t* __restrict__ w,
                                                             int32_t* __restrict__ out, long long M, int P, int O, int K,
                                                             int Kp, int32_t zpx, int32_t zpw) {
  constexpr int TM = 64, TN = 64, TK = 32;
  __shared__ int sa[TK][TM + 1];
  __shared__ int sw[TK][TN + 1];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const long long m0 = (long long)blockIdx.x * TM;
  const int o0 = blockIdx.y * TN;
  int acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0;
  for (int k0 = 0; k0 < K; k0 += TK) {
    for (int t = threadIdx.x; t < TM * TK; t += 256) {
      const int kk = t % TK, r = t / TK;
      const long long m = m0 + r;
      const int k = k0 + kk;
      int va = 0, vw = 0;
      if (m < M && k < K) {
        const uint8_t b = col[m * Kp + k];
        va = (XS ? (int)(int8_t)b : (int)b) - zpx;
      }
      if (o0 + r < O && k < K) {
        const uint8_t b = w[(long long)(o0 + r) * K + k];
        vw = (WS ? (int)(int8_t)b : (int)b) - zpw;
      }
      sa[kk][r] = va;
      sw[kk][r] = vw;
    }
    __syncthreads();
#pragma unroll 8
    for (int kk = 0; kk < TK; ++kk) {
      int a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sa[kk][tx + 16 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sw[kk][ty * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * b[j];  // int32 wrap-around like the reference's int32 accumulation
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + tx + 16 * i;
    if (m >= M) continue;
    const long long n = m / P;
    const int p = (int)(m % P);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int o = o0 + ty * 4 + j;
      if (o < O) out[(n * O + o) * P + p] = acc[i][j];
    }
  }
}

}  // namespace

extern "C" int tkb_qnn_conv2d_noalloc(const tkb_conv2d_desc* d, const void* data, const void* weight, int32_t* out,
                                      void* cuda_stream) {
  using namespace tkb;
  char msg[320];
  auto bad = [&](const char* m) {
    set_last_error(std::string("tkb_qnn_conv2d_noalloc: ") + m);
    return -1;
  };
  if (!d || !data || !weight || !out) return bad("null argument");
  if (d->n <= 0 || d->c <= 0 || d->h <= 0 || d->w <= 0 || d->o <= 0 || d->kh <= 0 || d->kw <= 0)
    return bad("sizes must be positive");
  if (d->stride_h <= 0 || d->stride_w <= 0 || d->dilation_h <= 0 || d->dilation_w <= 0)
    return bad("strides and dilations must be positive");
  if (d->pad_top < 0 || d->pad_left < 0 || d->pad_bottom < 0 || d->pad_right < 0) return bad("negative padding");
  if (d->kh > 255 || d->kw > 255) return bad("kernel sides above 255 are not supported");
  if ((d->data_dtype_code != 0 && d->data_dtype_code != 1) || (d->weight_dtype_code != 0 && d->weight_dtype_code != 1))
    return bad("dtype codes must be 0 (int8) or 1 (uint8)");
  const long long eh = (long long)d->dilation_h * (d->kh - 1) + 1, ew = (long long)d->dilation_w * (d->kw - 1) + 1;
  const long long OH = ((long long)d->h + d->pad_top + d->pad_bottom - eh) / d->stride_h + 1;
  const long long OW = ((long long)d->w + d->pad_left + d->pad_right - ew) / d->stride_w + 1;
  if ((long long)d->h + d->pad_top + d->pad_bottom < eh || (long long)d->w + d->pad_left + d->pad_right < ew || OH <= 0 || OW <= 0)
    return bad("the kernel does not fit the padded input");
  const bool padded = d->pad_top || d->pad_left || d->pad_bottom || d->pad_right;
  const int32_t lo = d->data_dtype_code == 0 ? -128 : 0, hi = d->data_dtype_code == 0 ? 127 : 255;
  if (padded && (d->input_zero_point < lo || d->input_zero_point > hi))
    return bad("with padding the input zero point must be representable in the data dtype");
  const long long K = (long long)d->c * d->kh * d->kw, Kp = (K + 15) / 16 * 16;
  if (Kp > TKB_NOALLOC_MAX_K) {
    snprintf(msg, sizeof msg, "C*KH*KW = %lld exceeds the maximum of %d", K, TKB_NOALLOC_MAX_K);
    return bad(msg);
  }
  const long long M = (long long)d->n * OH * OW;
  if (M > (1ll << 31) - 1 || M * Kp > (long long)TKB_NOALLOC_SCRATCH_BYTES) {
    snprintf(msg, sizeof msg,
             "the im2col matrix needs %lld bytes (N*OH*OW = %lld rows of %lld), the static scratch holds %d: split the batch",
             M * Kp, M, Kp, (int)TKB_NOALLOC_SCRATCH_BYTES);
    return bad(msg);
  }
  std::string err;
  uint8_t* scratch = scratch_for_current_device(&err);
  if (!scratch) {
    set_last_error(err);
    return -1;
  }
  cudaStream_t s = (cudaStream_t)cuda_stream;
  ConvGeom g{};
  g.N = d->n; g.H = d->h; g.W = d->w; g.C = d->c;
  g.O = d->o; g.KH = d->kh; g.KW = d->kw; g.groups = 1;
  g.sh = d->stride_h; g.sw = d->stride_w; g.pt = d->pad_top; g.pl = d->pad_left; g.pb = d->pad_bottom; g.pr = d->pad_right;
  g.dh = d->dilation_h; g.dw = d->dilation_w;
  g.OH = (int)OH; g.OW = (int)OW;
  g.a_dtype = d->data_dtype_code == 0 ? kI8 : kU8;
  g.w_dtype = d->weight_dtype_code == 0 ? kI8 : kU8;
  g.zp_a = d->input_zero_point; g.zp_w = nullptr; g.zp_w0 = d->kernel_zero_point;
  // 1. column matrix -> static scratch (every byte of the [M][Kp] region is written: nothing has to start at zero)
  launch_im2col_nchw(data, scratch, g, (int)Kp, s);
  // 2. GEMM scratch x weights -> NCHW int32
  dim3 grid((unsigned)((M + 63) / 64), (unsigned)((d->o + 63) / 64));
  const int P = (int)(OH * OW);
  // The padding byte is (zp_x & 0xFF) and reads back as zp_x (checked above: representable in the data dtype).
  auto launch = [&](auto kern) {
    kern<<<grid, 256, 0, s>>>(scratch, (const uint8_t*)weight, out, M, P, d->o, (int)K, (int)Kp, d->input_zero_point,
                              d->kernel_zero_point);
  };
  if (g.a_dtype == kI8) {
    if (g.w_dtype == kI8) launch(conv_gemm_nchw_kernel<true, true>);
    else launch(conv_gemm_nchw_kernel<true, false>);
  } else {
    if (g.w_dtype == kI8) launch(conv_gemm_nchw_kernel<false, true>);
    else launch(conv_gemm_nchw_kernel<false, false>);
  }
  ++g_launches;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return bad(cudaGetErrorString(e));
  return 0;
}
