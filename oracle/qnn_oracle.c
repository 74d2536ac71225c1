/*
 * qnn_oracle.c -- CPU restatement of the reference's integer-only QNN lowering (path "A").
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under tachikoma_b200/ may import, link or call this
 * file; it is the checker for tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs.  The reference (a TVM fork) cannot be compiled in this image (no
 * LLVM, no oneDNN), so this file restates the algorithm op by op.  Parity is
 * PINNED against the reference's own literal golden vectors (tests/test_oracle_golden.py
 * replays tests/python/relay/test_op_qnn_{requantize,dense,conv2d,add,batch_matmul}.py).
 *
 * Every function cites the reference file:line (relative to /root/reference) it follows.
 * Plain C99 + OpenMP, no dependencies.  All tensors are dense row-major.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* dtype codes shared with oracle/interp.py */
enum { TKO_I8 = 0, TKO_U8 = 1, TKO_I32 = 2, TKO_I16 = 3, TKO_I64 = 4, TKO_F32 = 5 };
enum { TKO_UPWARD = 0, TKO_TONEAREST = 1 };
enum { TKO_CD_INT64 = 0, TKO_CD_FLOAT32 = 1, TKO_CD_FLOAT64 = 2 };

static inline int64_t load_int(const void* p, int dtype, int64_t i) {
  switch (dtype) {
    case TKO_I8: return ((const int8_t*)p)[i];
    case TKO_U8: return ((const uint8_t*)p)[i];
    case TKO_I32: return ((const int32_t*)p)[i];
    case TKO_I16: return ((const int16_t*)p)[i];
    case TKO_I64: return ((const int64_t*)p)[i];
    default: return 0;
  }
}
static inline void store_int(void* p, int dtype, int64_t i, int64_t v) {
  switch (dtype) {
    case TKO_I8: ((int8_t*)p)[i] = (int8_t)v; break;
    case TKO_U8: ((uint8_t*)p)[i] = (uint8_t)v; break;
    case TKO_I32: ((int32_t*)p)[i] = (int32_t)v; break;
    case TKO_I16: ((int16_t*)p)[i] = (int16_t)v; break;
    case TKO_I64: ((int64_t*)p)[i] = v; break;
    default: break;
  }
}
static inline void dtype_range(int dtype, int64_t* lo, int64_t* hi) {
  switch (dtype) {
    case TKO_I8: *lo = -128; *hi = 127; break;
    case TKO_U8: *lo = 0; *hi = 255; break;
    case TKO_I16: *lo = -32768; *hi = 32767; break;
    default: *lo = INT32_MIN; *hi = INT32_MAX; break;
  }
}
/* int32 arithmetic in the reference wraps (LLVM add/sub/mul without nsw on TVM int32). */
static inline int32_t wrap32(int64_t v) { return (int32_t)(uint32_t)(uint64_t)v; }

/* src/relay/qnn/utils.cc:33-57  GetFixedPointMultiplierShift */
void tko_fixed_point_multiplier_shift(double m, int32_t* significand, int32_t* exponent) {
  if (m == 0.) { *significand = 0; *exponent = 0; return; }
  int e;
  double sig = frexp(m, &e);
  sig = round(sig * (double)(1ll << 31));
  int64_t s64 = (int64_t)sig;
  if (s64 == (1ll << 31)) { s64 /= 2; ++e; }
  *significand = (int32_t)s64;
  *exponent = e;
}

/* src/target/intrin_rule.cc:165-195 QMultiplyShift (q = 31): UPWARD rounding, int64. */
static inline int32_t q_multiply_shift_generic(int32_t x, int32_t y, int32_t s) {
  int left = s > 0 ? s : 0, right = s > 0 ? 0 : -s;
  int64_t v = (int64_t)x;
  if (left) v = (int64_t)((uint64_t)v << left);
  v = v * (int64_t)y;
  int total = right + 31;
  v = v + ((int64_t)1 << (total - 1));
  v = v >> total; /* arithmetic */
  return (int32_t)v; /* truncating cast */
}

/* src/target/intrin_rule.cc:197-250 tir.q_multiply_shift default legalization (per-tensor
 * fixed_point_multiply, python/tvm/topi/math.py:643-674).  Power-of-two shortcut when the
 * multiplier is exactly 1<<30.  exp==0 would shift by -1 in the reference (undefined); it is
 * unreachable through qnn.requantize (equal scales skip the multiply, requantize.cc:226) and
 * is defined here as the identity, which is the mathematical value. */
int32_t tko_fixed_point_multiply(int32_t x, int32_t multiplier, int32_t shift) {
  if (multiplier == (1 << 30)) {
    int e = shift - 1;
    if (e > 0) return wrap32((int64_t)((uint64_t)(int64_t)x << e));
    if (e == 0) return x;
    e = -e;
    if (e >= 32) return x < 0 ? -1 : 0; /* x86 would mask the count; out of tested domain */
    int32_t r = wrap32((int64_t)x + ((int64_t)1 << (e - 1)));
    return r >> e;
  }
  return q_multiply_shift_generic(x, multiplier, shift);
}

/* src/relay/qnn/utils.cc:59-109 FixedPointMultiplyToNearest and :136-219 per-channel variant:
 * rounder is 2^(t-1) for v>=0 else 2^(t-1)-1; no shortcut. */
static inline int32_t fixed_point_multiply_tonearest(int32_t x, int32_t M, int32_t s) {
  int left = s > 0 ? s : 0, right = s > 0 ? 0 : -s;
  int64_t v = (int64_t)x;
  if (left) v = (int64_t)((uint64_t)v << left);
  v = v * (int64_t)M;
  int total = right + 31;
  int64_t pos = (int64_t)1 << (total - 1);
  v = v + (v >= 0 ? pos : pos - 1);
  v = v >> total;
  return (int32_t)v;
}

/* requantize.cc:131-173  Upward<Bits>/Tonearest<Bits> without SSE4.1 (plain "llvm" target). */
static inline double round_upward(double t) {
  double b = t + 0.5;
  double bi = (double)(int64_t)b;
  if (!isfinite(t)) return t;
  return (b == bi || b >= 0.0) ? bi : bi - 1.0;
}
static inline double round_tonearest(double t) {
  if (!isfinite(t)) return t;
  double m = t < 0 ? -1.0 : 1.0;
  double b = (t + 0.5 * m) * m;
  return (double)(int64_t)b * m;
}
static inline float round_upward_f(float t) {
  float b = t + 0.5f;
  float bi = (float)(int32_t)b;
  if (!isfinite(t)) return t;
  return (b == bi || b >= 0.0f) ? bi : bi - 1.0f;
}
static inline float round_tonearest_f(float t) {
  if (!isfinite(t)) return t;
  float m = t < 0 ? -1.0f : 1.0f;
  float b = (t + 0.5f * m) * m;
  return (float)(int32_t)b * m;
}

/*
 * qnn.requantize: src/relay/qnn/op/requantize.cc:195-273 (int64), :293-373 (fp32/fp64),
 * :383-406 dispatch.  Tensor viewed as [outer][C][inner] with the quantization axis = C.
 * n_scale is 1 (per-tensor) or C (per-channel); n_zp likewise for the input zero point.
 */
void tko_requantize(const void* in, int in_dtype, int64_t outer, int64_t C, int64_t inner,
                    const float* in_scale, int n_scale, const int32_t* in_zp, int n_zp,
                    float out_scale, int32_t out_zp, int rounding, int compute_dtype,
                    int out_dtype, void* out) {
  int32_t* M = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_scale > 0 ? n_scale : 1));
  int32_t* S = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_scale > 0 ? n_scale : 1));
  double* dm = (double*)malloc(sizeof(double) * (size_t)(n_scale > 0 ? n_scale : 1));
  for (int c = 0; c < n_scale; ++c) {
    dm[c] = (double)in_scale[c] / (double)out_scale; /* requantize.cc:222-224 */
    tko_fixed_point_multiplier_shift(dm[c], &M[c], &S[c]);
  }
  /* requantize.cc:226: per-tensor multiply is skipped when the two scale constants are equal */
  const int per_tensor = (n_scale == 1);
  const int skip = per_tensor && (in_scale[0] == out_scale);
  int64_t qmin, qmax;
  dtype_range(out_dtype, &qmin, &qmax);
  const int64_t total = outer * C * inner;
#pragma omp parallel for schedule(static) if (total > 65536)
  for (int64_t idx = 0; idx < total; ++idx) {
    int64_t c = (idx / inner) % C;
    int32_t zp = n_zp == 1 ? in_zp[0] : in_zp[c];
    int64_t y;
    if (compute_dtype == TKO_CD_INT64) {
      int32_t t = wrap32(load_int(in, in_dtype, idx) - (int64_t)zp);
      int32_t r;
      if (skip) r = t;
      else if (per_tensor)
        r = rounding == TKO_UPWARD ? tko_fixed_point_multiply(t, M[0], S[0])
                                   : fixed_point_multiply_tonearest(t, M[0], S[0]);
      else
        r = rounding == TKO_UPWARD ? q_multiply_shift_generic(t, M[c], S[c])
                                   : fixed_point_multiply_tonearest(t, M[c], S[c]);
      y = wrap32((int64_t)r + (int64_t)out_zp);
    } else if (compute_dtype == TKO_CD_FLOAT64) {
      double t = (double)load_int(in, in_dtype, idx) - (double)zp;
      if (!skip) t = per_tensor ? dm[0] * t : t * dm[c];
      t = t + (double)out_zp;
      t = rounding == TKO_UPWARD ? round_upward(t) : round_tonearest(t);
      y = (int32_t)t;
    } else {
      float t = (float)load_int(in, in_dtype, idx) - (float)zp;
      if (!skip) t = per_tensor ? (float)dm[0] * t : t * (float)dm[c];
      t = t + (float)out_zp;
      t = rounding == TKO_UPWARD ? round_upward_f(t) : round_tonearest_f(t);
      y = (int32_t)t;
    }
    if (out_dtype != TKO_I32) { if (y < qmin) y = qmin; if (y > qmax) y = qmax; }
    store_int(out, out_dtype, idx, y);
  }
  free(M); free(S); free(dm);
}

/*
 * qnn.conv2d: src/relay/qnn/op/convolution.cc:708-811 (QnnConv2DCanonicalize), padding with the
 * input zero point :227-253, terms :425-638, fallback :193-215.  All four expanded terms are
 * int32 ring arithmetic, so their sum equals the direct definition modulo 2^32:
 *   acc[n,k,y,x] = sum_{c,r,s} (QA[n, g*Cg+c, y*sh+r*dh-pt, x*sw+s*dw-pl] - zp_a) * (QW[k,c,r,s] - zp_w[k])
 * with out-of-image taps reading zp_a (contribute 0).  Layouts are expressed by element
 * strides: data (n,c,h,w), kernel (o,i,h,w), output (n,c,h,w).  zp_w is scalar (n_zpw==1) or
 * per-output-channel (convolution.cc:729-746).
 */
void tko_conv2d(const void* data, int d_dtype, const void* wgt, int w_dtype, int32_t* out,
                int64_t N, int64_t C, int64_t H, int64_t W, int64_t O, int64_t KH, int64_t KW,
                int64_t groups, int64_t sh, int64_t sw, int64_t pt, int64_t pl, int64_t pb,
                int64_t pr, int64_t dh, int64_t dw, const int64_t* ds, const int64_t* ws,
                const int64_t* os, int32_t zp_a, const int32_t* zp_w, int n_zpw) {
  const int64_t OH = (H + pt + pb - (KH - 1) * dh - 1) / sh + 1;
  const int64_t OW = (W + pl + pr - (KW - 1) * dw - 1) / sw + 1;
  const int64_t Cg = C / groups, Og = O / groups;
#pragma omp parallel for collapse(2) schedule(static)
  for (int64_t n = 0; n < N; ++n)
    for (int64_t k = 0; k < O; ++k) {
      const int64_t g = k / Og;
      const int32_t zw = n_zpw == 1 ? zp_w[0] : zp_w[k];
      for (int64_t y = 0; y < OH; ++y)
        for (int64_t x = 0; x < OW; ++x) {
          int32_t acc = 0;
          for (int64_t r = 0; r < KH; ++r) {
            const int64_t iy = y * sh + r * dh - pt;
            if (iy < 0 || iy >= H) continue; /* tap reads zp_a -> (zp_a - zp_a) = 0 */
            for (int64_t s = 0; s < KW; ++s) {
              const int64_t ix = x * sw + s * dw - pl;
              if (ix < 0 || ix >= W) continue;
              const int64_t dbase = n * ds[0] + (g * Cg) * ds[1] + iy * ds[2] + ix * ds[3];
              const int64_t wbase = k * ws[0] + r * ws[2] + s * ws[3];
              int32_t part = 0;
              for (int64_t c = 0; c < Cg; ++c) {
                int32_t a = (int32_t)load_int(data, d_dtype, dbase + c * ds[1]) - zp_a;
                int32_t b = (int32_t)load_int(wgt, w_dtype, wbase + c * ws[1]) - zw;
                part = wrap32((int64_t)part + (int64_t)a * (int64_t)b);
              }
              acc = wrap32((int64_t)acc + part);
            }
          }
          out[n * os[0] + k * os[1] + y * os[2] + x * os[3]] = acc;
        }
    }
}

/* qnn.dense: src/relay/qnn/op/dense.cc:160-205 (four-term expansion == direct definition
 * modulo 2^32).  X is [M,K], W is [N,K] (units x in), out [M,N] int32.  zp_w scalar. */
void tko_dense(const void* x, int x_dtype, const void* w, int w_dtype, int32_t* out, int64_t M,
               int64_t N, int64_t K, int32_t zp_x, int32_t zp_w) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int64_t m = 0; m < M; ++m)
    for (int64_t n = 0; n < N; ++n) {
      int32_t acc = 0;
      for (int64_t k = 0; k < K; ++k) {
        int32_t a = (int32_t)load_int(x, x_dtype, m * K + k) - zp_x;
        int32_t b = (int32_t)load_int(w, w_dtype, n * K + k) - zp_w;
        acc = wrap32((int64_t)acc + (int64_t)a * (int64_t)b);
      }
      out[m * N + n] = acc;
    }
}

/* qnn.batch_matmul: src/relay/qnn/op/batch_matmul.cc:161-200; only transpose_a=false,
 * transpose_b=true exists (:83-87).  X [B,M,K], Y [B,N,K] (B of Y may be 1: broadcast). */
void tko_batch_matmul(const void* x, int x_dtype, const void* y, int y_dtype, int32_t* out,
                      int64_t B, int64_t M, int64_t N, int64_t K, int64_t By, int32_t zp_x,
                      int32_t zp_y) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int64_t b = 0; b < B; ++b)
    for (int64_t m = 0; m < M; ++m)
      for (int64_t n = 0; n < N; ++n) {
        const int64_t yb = By == 1 ? 0 : b;
        int32_t acc = 0;
        for (int64_t k = 0; k < K; ++k) {
          int32_t a = (int32_t)load_int(x, x_dtype, (b * M + m) * K + k) - zp_x;
          int32_t c = (int32_t)load_int(y, y_dtype, (yb * N + n) * K + k) - zp_y;
          acc = wrap32((int64_t)acc + (int64_t)a * (int64_t)c);
        }
        out[(b * M + m) * N + n] = acc;
      }
}

/* One side of qnn.add: RequantizeOrUpcast, src/relay/qnn/op/op_common.h:196-210 -- requantize
 * to the output qparams (default config: UPWARD, int64; requantize_config.h:53-73) unless both
 * scale and zero point already match, in which case a plain cast to int32. */
static inline int32_t requantize_or_upcast(int32_t q, float s, int32_t zp, float so, int32_t zpo,
                                           int32_t M, int32_t S) {
  if (s == so && zp == zpo) return q;
  int32_t t = wrap32((int64_t)q - zp);
  if (s != so) t = tko_fixed_point_multiply(t, M, S);
  return wrap32((int64_t)t + zpo);
}

/* qnn.add: src/relay/qnn/op/add.cc:40-93.  Q_c = Q_a' + Q_b' - zp_c in int32, then
 * ConvertDtype (clip to the INPUT dtype range, cast; op_common.h:170-175).  Same-shape only. */
void tko_qnn_add(const void* a, const void* b, int dtype, int64_t n, float sa, int32_t zpa,
                 float sb, int32_t zpb, float so, int32_t zpo, void* out) {
  int32_t Ma, Sa, Mb, Sb;
  tko_fixed_point_multiplier_shift((double)sa / (double)so, &Ma, &Sa);
  tko_fixed_point_multiplier_shift((double)sb / (double)so, &Mb, &Sb);
  int64_t lo, hi;
  dtype_range(dtype, &lo, &hi);
#pragma omp parallel for schedule(static) if (n > 65536)
  for (int64_t i = 0; i < n; ++i) {
    int32_t qa = requantize_or_upcast((int32_t)load_int(a, dtype, i), sa, zpa, so, zpo, Ma, Sa);
    int32_t qb = requantize_or_upcast((int32_t)load_int(b, dtype, i), sb, zpb, so, zpo, Mb, Sb);
    int64_t y = wrap32((int64_t)wrap32((int64_t)qa + qb) - zpo);
    if (y < lo) y = lo;
    if (y > hi) y = hi;
    store_int(out, dtype, i, y);
  }
}

/* qnn.subtract, src/relay/qnn/op/subtract.cc:40-93: both sides through RequantizeOrUpcast to the output qparams,
 * Q_a' - Q_b' in int32, + zp_out (skipped when it is zero: adding zero changes nothing), saturate to the dtype. */
void tko_qnn_subtract(const void* a, const void* b, int dtype, int64_t n, float sa, int32_t zpa,
                      float sb, int32_t zpb, float so, int32_t zpo, void* out) {
  int32_t Ma, Sa, Mb, Sb;
  tko_fixed_point_multiplier_shift((double)sa / (double)so, &Ma, &Sa);
  tko_fixed_point_multiplier_shift((double)sb / (double)so, &Mb, &Sb);
  int64_t lo, hi;
  dtype_range(dtype, &lo, &hi);
#pragma omp parallel for schedule(static) if (n > 65536)
  for (int64_t i = 0; i < n; ++i) {
    int32_t qa = requantize_or_upcast((int32_t)load_int(a, dtype, i), sa, zpa, so, zpo, Ma, Sa);
    int32_t qb = requantize_or_upcast((int32_t)load_int(b, dtype, i), sb, zpb, so, zpo, Mb, Sb);
    int64_t y = wrap32((int64_t)wrap32((int64_t)qa - qb) + zpo);
    if (y < lo) y = lo;
    if (y > hi) y = hi;
    store_int(out, dtype, i, y);
  }
}

/* nn.max_pool2d / nn.avg_pool2d on integer tensors: include/tvm/topi/nn/pooling.h:384-389 (adaptive,
 * used by global pooling: is_avg == 2), :585-650 (windowed: is_avg == 1).  The window sum is reduced
 * IN THE TENSOR DTYPE (tvm::sum(temp(indices), daxis), :599; int32 after the frontend's upcast,
 * python/tvm/relay/frontend/qnn_torch.py:587-590), i.e. it wraps at 8 / 16 / 32 bits.  The windowed
 * form divides by an int32 element count (the narrow sum is promoted to int32 first, :612-648); the
 * adaptive form multiplies its divisor up in the tensor dtype (:384-387), so the divisor wraps too
 * (a zero divisor is a division by zero in the reference: reported here as 0, rejected by the product
 * at init).  TRUNCATING division.  Max-pool padding reads the dtype minimum.  Strides describe
 * (n,c,h,w) for in and out. */
static int64_t wrap_dtype(int64_t v, int dtype) {
  switch (dtype) {
    case 0: return (int8_t)v;
    case 1: return (uint8_t)v;
    case 3: return (int16_t)v;
    case 4: return v;
    default: return (int32_t)v;
  }
}
void tko_pool2d(const void* in, int dtype, void* out, int is_avg, int count_include_pad, int64_t N,
                int64_t C, int64_t H, int64_t W, int64_t KH, int64_t KW, int64_t sh, int64_t sw,
                int64_t pt, int64_t pl, int64_t pb, int64_t pr, const int64_t* is,
                const int64_t* os) {
  const int64_t OH = (H + pt + pb - KH) / sh + 1, OW = (W + pl + pr - KW) / sw + 1;
  int64_t lo, hi;
  dtype_range(dtype, &lo, &hi);
#pragma omp parallel for collapse(2) schedule(static)
  for (int64_t n = 0; n < N; ++n)
    for (int64_t c = 0; c < C; ++c)
      for (int64_t y = 0; y < OH; ++y)
        for (int64_t x = 0; x < OW; ++x) {
          int64_t y0 = y * sh - pt, x0 = x * sw - pl;
          int64_t y1 = y0 + KH - 1, x1 = x0 + KW - 1;
          int64_t best = lo, cnt = 0;
          int32_t sum = 0;
          for (int64_t iy = y0; iy <= y1; ++iy)
            for (int64_t ix = x0; ix <= x1; ++ix) {
              if (iy < 0 || iy >= H || ix < 0 || ix >= W) continue;
              int64_t v = load_int(in, dtype, n * is[0] + c * is[1] + iy * is[2] + ix * is[3]);
              if (v > best) best = v;
              sum = (int32_t)wrap_dtype((int64_t)sum + v, dtype);
              ++cnt;
            }
          int64_t r;
          if (!is_avg) r = best;
          else {
            int64_t div;
            if (is_avg == 2) {
              div = wrap_dtype(wrap_dtype(KH, dtype) * wrap_dtype(KW, dtype), dtype);
            } else if (count_include_pad) {
              int64_t ye = y1 < H + pb - 1 ? y1 : H + pb - 1, xe = x1 < W + pr - 1 ? x1 : W + pr - 1;
              div = (ye - y0 + 1) * (xe - x0 + 1);
            } else div = cnt > 1 ? cnt : 1;
            r = div == 0 ? 0 : (int32_t)sum / (int32_t)div; /* C division truncates toward zero */
          }
          store_int(out, dtype, n * os[0] + c * os[1] + y * os[2] + x * os[3], r);
        }
}

/* Tachikoma/DNNL float epilogue ("path B"), python/tvm/relay/op/contrib/tachikoma.py:1238-1277 and
 * src/runtime/contrib/tachikoma/tachikoma_json_runtime.cc:142-185: per element, in fp32,
 *   ((acc + bias_f[k]) * o_scl[k] -> clip[0,255]) * act_scl + sum_scl * sum + dst_zp -> saturating
 * round-to-nearest-even convert to the destination dtype (oneDNN's conversion; oneDNN itself is
 * not under /root/reference, so this mode is judged at the reference's own +-1 tolerance,
 * tests/python/contrib/test_tachikoma.py:1615-1616). Layout [outer][C][inner]. */
void tko_dnnl_float_epilogue(const int32_t* acc, int64_t outer, int64_t C, int64_t inner,
                             const float* bias_f, const float* o_scl, int n_oscl, float act_scl,
                             const void* sum, int sum_dtype, float sum_scl, float dst_zp,
                             int out_dtype, void* out) {
  int64_t lo, hi;
  dtype_range(out_dtype, &lo, &hi);
  const int64_t total = outer * C * inner;
#pragma omp parallel for schedule(static) if (total > 65536)
  for (int64_t idx = 0; idx < total; ++idx) {
    int64_t c = (idx / inner) % C;
    float v = (float)acc[idx];
    if (bias_f) v = v + bias_f[c];
    v = v * (n_oscl == 1 ? o_scl[0] : o_scl[c]);
    v = v < 0.f ? 0.f : (v > 255.f ? 255.f : v);
    v = v * act_scl;
    if (sum) v = v + sum_scl * (float)load_int(sum, sum_dtype, idx);
    v = v + dst_zp;
    float r = nearbyintf(v);
    int64_t y = r < (float)lo ? lo : (r > (float)hi ? hi : (int64_t)r);
    store_int(out, out_dtype, idx, y);
  }
}

/* qnn.quantize, src/relay/qnn/op/quantize.cc:113-150 (QuantizeLower): in float32,
 *   clip(round(x / scale) + float(zp), qmin, qmax) -> cast to out_dtype
 * round = tir.round = llvm.round (half away from zero: src/target/llvm/intrin_rule_llvm.cc:82-84).
 * Layout [outer][C][inner]; n_q = 1 (per tensor) or C (per channel on the axis). */
void tko_quantize(const float* in, int64_t outer, int64_t C, int64_t inner, const float* scale,
                  const int32_t* zp, int n_scale, int n_zp, int out_dtype, void* out) {
  int64_t lo, hi;
  dtype_range(out_dtype, &lo, &hi);
  const int64_t total = outer * C * inner;
#pragma omp parallel for schedule(static) if (total > 65536)
  for (int64_t idx = 0; idx < total; ++idx) {
    const int64_t c = (idx / inner) % C;
    float v = in[idx] / (n_scale == 1 ? scale[0] : scale[c]);
    v = roundf(v);
    v = v + (float)(n_zp == 1 ? zp[0] : zp[c]);
    v = v < (float)lo ? (float)lo : (v > (float)hi ? (float)hi : v);
    store_int(out, out_dtype, idx, (int64_t)v);
  }
}

/* qnn.dequantize, src/relay/qnn/op/dequantize.cc:95-127 (DequantizeLower):
 *   float32(int32(x) - zp) * scale   (int32 subtraction wraps, float32 multiply) */
void tko_dequantize(const void* in, int in_dtype, int64_t outer, int64_t C, int64_t inner,
                    const float* scale, const int32_t* zp, int n_scale, int n_zp, float* out) {
  const int64_t total = outer * C * inner;
#pragma omp parallel for schedule(static) if (total > 65536)
  for (int64_t idx = 0; idx < total; ++idx) {
    const int64_t c = (idx / inner) % C;
    const int32_t t = wrap32((int64_t)(int32_t)load_int(in, in_dtype, idx) - (int64_t)(n_zp == 1 ? zp[0] : zp[c]));
    out[idx] = (float)t * (n_scale == 1 ? scale[0] : scale[c]);
  }
}

/* Thread control for the timed CPU-baseline legs of bench.py: launchers such as torchrun export
 * OMP_NUM_THREADS=1, which would silently turn the "all host cores" baseline into a single-thread one. */
int tko_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}
