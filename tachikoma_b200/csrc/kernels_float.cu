// float32 operators of the reference's float tachikoma patterns: what the reference runs as
// oneDNN fp32 primitives with post-ops (src/runtime/contrib/tachikoma/tachikoma_json_runtime.cc:292-764).
//   convolution / deconvolution 1-D, 2-D, 3-D (+ bias, + sum, + activation)      :292-460
//   inner product, batch matmul (+ bias, + activation)                           :462-548
//   batch norm (inference), layer norm, softmax, pooling, eltwise, binary        :550-764
// These are plain fp32 FMA kernels: the reference's bar for this half is rtol = atol = 1e-3 against TVM's own fp32 build
// (tests/python/contrib/test_tachikoma.py:155), which rules out the tf32 tensor-core kinds (10-bit mantissa); the
// integer QNN chain is the tensor-core path of this library, this file only makes float subgraphs executable.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "kernels.h"

namespace tkb {

__device__ __forceinline__ float f_act(float v, int act, float a, float b) {
  switch (act) {
    case kFActRelu: return v > 0.f ? v : 0.f;
    case kFActTanh: return tanhf(v);
    case kFActSigmoid: return 1.f / (1.f + expf(-v));
    case kFActClip: return fminf(fmaxf(v, a), b);
    case kFActGelu: return 0.5f * v * (1.f + erff(v * 0.70710678118654752440f));
    case kFActSwish: return v / (1.f + expf(-v));                                   // x * sigmoid(x)
    case kFActMish: return v * tanhf(v > 20.f ? v : log1pf(expf(v)));                // x * tanh(softplus(x))
    case kFActAbs: return fabsf(v);
    case kFActExp: return expf(v);
    case kFActLog: return logf(v);
    case kFActSqrt: return sqrtf(v);
    case kFActRound: return nearbyintf(v);                                           // oneDNN eltwise_round: half to even
    case kFActLeakyRelu: return v > 0.f ? v : a * v;
    default: return v;
  }
}

// ---- direct convolution, NC[D]HW x OI[D]HW (transposed: IO[D]HW), one output element per thread ----------------------
__global__ void fconv_kernel(const float* __restrict__ x, const float* __restrict__ w, float* __restrict__ y, FConvGeom g,
                             FPost post) {
  const long long total = (long long)g.N * g.O * g.OD * g.OH * g.OW;
  const int cg = g.C / g.groups, og = g.O / g.groups;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    long long t = idx;
    const int ow = (int)(t % g.OW); t /= g.OW;
    const int oh = (int)(t % g.OH); t /= g.OH;
    const int od = (int)(t % g.OD); t /= g.OD;
    const int o = (int)(t % g.O);
    const int n = (int)(t / g.O);
    const int grp = o / og;
    float acc = 0.f;
    for (int ci = 0; ci < cg; ++ci) {
      const int c = grp * cg + ci;
      const float* xp = x + ((long long)n * g.C + c) * g.D * g.H * g.W;
      // weights: [O][C/g][KD][KH][KW], transposed convolution: [C][O/g][KD][KH][KW]
      const float* wp = g.transposed ? w + ((long long)c * og + (o - grp * og)) * g.KD * g.KH * g.KW
                                     : w + ((long long)o * cg + ci) * g.KD * g.KH * g.KW;
      for (int kd = 0; kd < g.KD; ++kd) {
        int id;
        if (g.transposed) {
          const int v = od + g.pd - kd * g.dd;
          if (v < 0 || v % g.sd) continue;
          id = v / g.sd;
        } else {
          id = od * g.sd - g.pd + kd * g.dd;
        }
        if (id < 0 || id >= g.D) continue;
        for (int kh = 0; kh < g.KH; ++kh) {
          int ih;
          if (g.transposed) {
            const int v = oh + g.ph - kh * g.dh;
            if (v < 0 || v % g.sh) continue;
            ih = v / g.sh;
          } else {
            ih = oh * g.sh - g.ph + kh * g.dh;
          }
          if (ih < 0 || ih >= g.H) continue;
          const float* xr = xp + ((long long)id * g.H + ih) * g.W;
          const float* wr = wp + (kd * g.KH + kh) * g.KW;
          for (int kw = 0; kw < g.KW; ++kw) {
            int iw;
            if (g.transposed) {
              const int v = ow + g.pw - kw * g.dw;
              if (v < 0 || v % g.sw) continue;
              iw = v / g.sw;
            } else {
              iw = ow * g.sw - g.pw + kw * g.dw;
            }
            if (iw < 0 || iw >= g.W) continue;
            acc = fmaf(__ldg(xr + iw), __ldg(wr + kw), acc);
          }
        }
      }
    }
    if (post.bias) acc += __ldg(post.bias + o);
    if (post.sum) acc += __ldg(post.sum + idx);
    y[idx] = f_act(acc, post.act, post.a, post.b);
  }
}

// ---- convolution as an implicit GEMM: per group, Y[o][pixel] = sum_k W[o][k] * X[pixel][k] with k = (ci, kd, kh, kw) gathered
// on the fly; 64 channels x 64 pixels x 16 k tiles, 4 x 4 outputs per thread, consecutive threads along the pixels (coalesced
// gathers along W and coalesced stores into the NC[D]HW output).  Used when a group has at least 16 output channels; depthwise
// and other thin groups keep the direct kernel.
__global__ void __launch_bounds__(256) fconv_gemm_kernel(const float* __restrict__ x, const float* __restrict__ w, float* __restrict__ y,
                                                         FConvGeom g, FPost post) {
  __shared__ float Ws[16][64 + 1], Xs[16][64 + 1];
  const int cg = g.C / g.groups, og = g.O / g.groups;
  const int P = g.OD * g.OH * g.OW, KK = g.KD * g.KH * g.KW, K = cg * KK;
  const long long pixels = (long long)g.N * P;
  const int grp = blockIdx.z, o0 = blockIdx.y * 64;
  const long long p0 = (long long)blockIdx.x * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  // the pixel this thread gathers for (fixed over the k loop)
  const int lr = threadIdx.x & 63, lk = threadIdx.x >> 6;  // lk: 0..3
  const long long gp = p0 + lr;
  const bool p_ok = gp < pixels;
  int img = 0, id0 = 0, ih0 = 0, iw0 = 0;
  if (p_ok) {
    img = (int)(gp / P);
    int r = (int)(gp - (long long)img * P);
    const int ow = r % g.OW; r /= g.OW;
    const int oh = r % g.OH;
    const int od = r / g.OH;
    id0 = od * g.sd - g.pd; ih0 = oh * g.sh - g.ph; iw0 = ow * g.sw - g.pw;
  }
  const float* xg = x + ((long long)img * g.C + (long long)grp * cg) * g.D * g.H * g.W;
  const float* wg = w + (long long)grp * og * K;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += 16) {
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {  // weights: consecutive threads along k (contiguous)
      const int kk = i & 15, r = i >> 4;
      Ws[kk][r] = (o0 + r < og && k0 + kk < K) ? __ldg(wg + (long long)(o0 + r) * K + k0 + kk) : 0.f;
    }
#pragma unroll
    for (int pass = 0; pass < 4; ++pass) {  // gathered activations: consecutive threads along the pixels
      const int kk = lk + 4 * pass, k = k0 + kk;
      float v = 0.f;
      if (p_ok && k < K) {
        const int ci = k / KK;
        int t = k - ci * KK;
        const int kw = t % g.KW; t /= g.KW;
        const int kh = t % g.KH;
        const int kd = t / g.KH;
        const int id = id0 + kd * g.dd, ih = ih0 + kh * g.dh, iw = iw0 + kw * g.dw;
        if (id >= 0 && id < g.D && ih >= 0 && ih < g.H && iw >= 0 && iw < g.W)
          v = __ldg(xg + (((long long)ci * g.D + id) * g.H + ih) * g.W + iw);
      }
      Xs[kk][lr] = v;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { av[i] = Ws[kk][ty * 4 + i]; bv[i] = Xs[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const long long q = p0 + tx * 4 + j;
    if (q >= pixels) continue;
    const int qi = (int)(q / P);
    const int qr = (int)(q - (long long)qi * P);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int o = o0 + ty * 4 + i;
      if (o >= og) continue;
      const int oc = grp * og + o;
      const long long idx = ((long long)qi * g.O + oc) * P + qr;
      float v = acc[i][j];
      if (post.bias) v += __ldg(post.bias + oc);
      if (post.sum) v += __ldg(post.sum + idx);
      y[idx] = f_act(v, post.act, post.a, post.b);
    }
  }
}

bool fconv_gemm_form(const FConvGeom& g) {
  static const bool direct_only = getenv("TKB_FCONV_DIRECT") != nullptr;  // A/B knob (tools/fconv_probe.py)
  const long long pixels = (long long)g.N * g.OD * g.OH * g.OW;
  return !direct_only && !g.transposed && g.O / g.groups >= 16 && g.groups <= 65535 && (pixels + 63) / 64 < 0x7fffffffLL;
}

void launch_fconv(const float* x, const float* w, float* y, const FConvGeom& g, const FPost& post, cudaStream_t s) {
  const long long total = (long long)g.N * g.O * g.OD * g.OH * g.OW;
  if (total <= 0) return;
  const int og = g.O / g.groups;
  const long long pixels = (long long)g.N * g.OD * g.OH * g.OW;
  if (fconv_gemm_form(g)) {
    dim3 grid((unsigned)((pixels + 63) / 64), (unsigned)((og + 63) / 64), (unsigned)g.groups);
    fconv_gemm_kernel<<<grid, 256, 0, s>>>(x, w, y, g, post);
    ++g_launches;
    return;
  }
  const int threads = 256;
  const long long blocks = (total + threads - 1) / threads;
  fconv_kernel<<<(unsigned)(blocks < 148 * 32 ? blocks : 148 * 32), threads, 0, s>>>(x, w, y, g, post);
  ++g_launches;
}

// ---- batched GEMM, C[b][m][n] = sum_k A[b](m,k) * B[b](n,k), arbitrary element strides, 64 x 64 x 16 tiles -------------
// (nn.dense: A = [M,K], B = [N,K]; nn.batch_matmul with either operand transposed through the strides)
__global__ void __launch_bounds__(256) fgemm_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ C,
                                                    int M, int N, int K, long long a_b, long long a_m, long long a_k,
                                                    long long b_b, long long b_n, long long b_k, FPost post) {
  __shared__ float As[16][64 + 1], Bs[16][64 + 1];
  const int bz = blockIdx.z, m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const float* Ab = A + (long long)bz * a_b;
  const float* Bb = B + (long long)bz * b_b;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += 16) {
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      const int kk = i & 15, r = i >> 4;  // consecutive threads along k (the contiguous axis of both operands by default)
      const int m = m0 + r, n = n0 + r, k = k0 + kk;
      As[kk][r] = (m < M && k < K) ? __ldg(Ab + (long long)m * a_m + (long long)k * a_k) : 0.f;
      Bs[kk][r] = (n < N && k < K) ? __ldg(Bb + (long long)n * b_n + (long long)k * b_k) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { av[i] = As[kk][ty * 4 + i]; bv[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      const long long idx = ((long long)bz * M + m) * N + n;
      float v = acc[i][j];
      if (post.bias) v += __ldg(post.bias + n);
      if (post.sum) v += __ldg(post.sum + idx);
      C[idx] = f_act(v, post.act, post.a, post.b);
    }
  }
}

void launch_fgemm(const float* A, const float* B, float* C, int batch, int M, int N, int K, long long a_b, long long a_m,
                  long long a_k, long long b_b, long long b_n, long long b_k, const FPost& post, cudaStream_t s) {
  if (batch <= 0 || M <= 0 || N <= 0) return;
  dim3 grid((N + 63) / 64, (M + 63) / 64, batch);
  fgemm_kernel<<<grid, 256, 0, s>>>(A, B, C, M, N, K, a_b, a_m, a_k, b_b, b_n, b_k, post);
  ++g_launches;
}

// ---- elementwise unary / binary with broadcasting ----------------------------------------------------------------------
__global__ void funary_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, int act, float a, float b) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = f_act(x[i], act, a, b);
}
void launch_funary(const float* x, float* y, long long n, int act, float a, float b, cudaStream_t s) {
  if (n <= 0) return;
  const long long blocks = (n + 255) / 256;
  funary_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, s>>>(x, y, n, act, a, b);
  ++g_launches;
}

// out[i0..i4] = lhs[...] op rhs[...]; a stride of 0 broadcasts that axis
__global__ void fbinary_kernel(const float* __restrict__ l, const float* __restrict__ r, float* __restrict__ y, FBcast5 bc, int mul) {
  const long long n = bc.shape[0] * bc.shape[1] * bc.shape[2] * bc.shape[3] * bc.shape[4];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    long long t = i, lo = 0, ro = 0;
#pragma unroll
    for (int d = 4; d >= 0; --d) {
      const long long c = t % bc.shape[d];
      t /= bc.shape[d];
      lo += c * bc.ls[d];
      ro += c * bc.rs[d];
    }
    y[i] = mul ? l[lo] * r[ro] : l[lo] + r[ro];
  }
}
void launch_fbinary(const float* l, const float* r, float* y, const FBcast5& bc, int mul, cudaStream_t s) {
  const long long n = bc.shape[0] * bc.shape[1] * bc.shape[2] * bc.shape[3] * bc.shape[4];
  if (n <= 0) return;
  const long long blocks = (n + 255) / 256;
  fbinary_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, s>>>(l, r, y, bc, mul);
  ++g_launches;
}

// ---- batch norm (inference, axis = 1): y = (x - mean[c]) / sqrt(var[c] + eps) * gamma[c] + beta[c] ------------------
__global__ void fbatchnorm_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, int C, long long inner,
                                  const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                                  const float* __restrict__ var, float eps) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)((i / inner) % C);
    y[i] = (x[i] - __ldg(mean + c)) / sqrtf(__ldg(var + c) + eps) * __ldg(gamma + c) + __ldg(beta + c);
  }
}
void launch_fbatchnorm(const float* x, float* y, long long outer, int C, long long inner, const float* gamma, const float* beta,
                       const float* mean, const float* var, float eps, cudaStream_t s) {
  const long long n = outer * C * inner;
  if (n <= 0) return;
  const long long blocks = (n + 255) / 256;
  fbatchnorm_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, s>>>(x, y, n, C, inner, gamma, beta, mean, var, eps);
  ++g_launches;
}

// ---- layer norm over the last axis: one warp per row -----------------------------------------------------------------------
__global__ void flayernorm_kernel(const float* __restrict__ x, float* __restrict__ y, long long rows, int C,
                                  const float* __restrict__ gamma, const float* __restrict__ beta, float eps) {
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5, nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = warp; r < rows; r += nwarps) {
    const float* xr = x + r * C;
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += xr[c];
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / (float)C;
    float v = 0.f;
    for (int c = lane; c < C; c += 32) { const float d = xr[c] - mean; v = fmaf(d, d, v); }
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const float inv = 1.f / sqrtf(v / (float)C + eps);
    for (int c = lane; c < C; c += 32) y[r * C + c] = (xr[c] - mean) * inv * __ldg(gamma + c) + __ldg(beta + c);
  }
}
void launch_flayernorm(const float* x, float* y, long long rows, int C, const float* gamma, const float* beta, float eps,
                       cudaStream_t s) {
  if (rows <= 0 || C <= 0) return;
  const long long blocks = (rows + 7) / 8;
  flayernorm_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, s>>>(x, y, rows, C, gamma, beta, eps);
  ++g_launches;
}

// ---- softmax along one axis of [outer][C][inner]: one warp per (outer, inner) pair ---------------------------------------
__global__ void fsoftmax_kernel(const float* __restrict__ x, float* __restrict__ y, long long outer, int C, long long inner) {
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5, nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long p = warp; p < outer * inner; p += nwarps) {
    const long long base = (p / inner) * C * inner + p % inner;
    float m = -INFINITY;
    for (int c = lane; c < C; c += 32) m = fmaxf(m, x[base + c * inner]);
#pragma unroll
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += expf(x[base + c * inner] - m);
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    for (int c = lane; c < C; c += 32) y[base + c * inner] = expf(x[base + c * inner] - m) / s;
  }
}
void launch_fsoftmax(const float* x, float* y, long long outer, int C, long long inner, cudaStream_t s) {
  if (outer * inner <= 0 || C <= 0) return;
  const long long blocks = (outer * inner + 7) / 8;
  fsoftmax_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, s>>>(x, y, outer, C, inner);
  ++g_launches;
}

// ---- pooling, NC[D]HW: max / average (count_include_pad or not) -------------------------------------------------------------
__global__ void fpool_kernel(const float* __restrict__ x, float* __restrict__ y, FConvGeom g, int is_avg, int count_include_pad) {
  const long long total = (long long)g.N * g.C * g.OD * g.OH * g.OW;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    long long t = idx;
    const int ow = (int)(t % g.OW); t /= g.OW;
    const int oh = (int)(t % g.OH); t /= g.OH;
    const int od = (int)(t % g.OD); t /= g.OD;
    const float* xp = x + t * g.D * g.H * g.W;  // t = n * C + c
    float acc = is_avg ? 0.f : -INFINITY;
    int cnt = 0;
    for (int kd = 0; kd < g.KD; ++kd) {
      const int id = od * g.sd - g.pd + kd * g.dd;
      if (id < 0 || id >= g.D) continue;
      for (int kh = 0; kh < g.KH; ++kh) {
        const int ih = oh * g.sh - g.ph + kh * g.dh;
        if (ih < 0 || ih >= g.H) continue;
        for (int kw = 0; kw < g.KW; ++kw) {
          const int iw = ow * g.sw - g.pw + kw * g.dw;
          if (iw < 0 || iw >= g.W) continue;
          const float v = xp[((long long)id * g.H + ih) * g.W + iw];
          acc = is_avg ? acc + v : fmaxf(acc, v);
          ++cnt;
        }
      }
    }
    if (is_avg) acc /= (float)(count_include_pad ? g.KD * g.KH * g.KW : (cnt > 0 ? cnt : 1));
    y[idx] = acc;
  }
}
void launch_fpool(const float* x, float* y, const FConvGeom& g, int is_avg, int count_include_pad, cudaStream_t s) {
  const long long total = (long long)g.N * g.C * g.OD * g.OH * g.OW;
  if (total <= 0) return;
  const long long blocks = (total + 255) / 256;
  fpool_kernel<<<(unsigned)(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, s>>>(x, y, g, is_avg, count_include_pad);
  ++g_launches;
}

}  // namespace tkb
