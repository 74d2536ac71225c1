// Fused integer epilogue shared by every conv/dense/bmm kernel (CUDA-core and tcgen05 paths) and by
// the stand-alone qnn.requantize / qnn.add kernels.  Bit-exact restatement, on the device, of
//   src/relay/qnn/op/requantize.cc:195-273 (RequantizeLowerInt), :293-373 (RequantizeLowerFP)
//   src/target/intrin_rule.cc:165-250 (q_multiply_shift, power-of-two shortcut)
//   src/relay/qnn/utils.cc:59-109, :136-219 (TONEAREST variants)
//   src/relay/qnn/op/add.cc:40-93 + op_common.h:170-210 (qnn.add, RequantizeOrUpcast)
//   python/tvm/relay/op/contrib/tachikoma.py:1238-1277 + tachikoma_json_runtime.cc:142-185 (float mode)
// The (multiplier, shift) pairs are derived on the HOST (frexp/round in double, utils.cc:33-57).
#pragma once
#include <cstdint>

namespace tkb {

enum DType : int { kI8 = 0, kU8 = 1, kI32 = 2, kI16 = 3, kI64 = 4, kF32 = 5 };
enum Rounding : int { kUpward = 0, kToNearest = 1 };
enum ComputeDT : int { kInt64 = 0, kFloat32 = 1, kFloat64 = 2 };

__host__ __device__ inline int dtype_size(int dt) {
  return dt == kI8 || dt == kU8 ? 1 : dt == kI16 ? 2 : dt == kI64 ? 8 : 4;
}

// One requantize step.  Per-channel tables live in device memory; scalars are used when
// per_channel == 0.
struct RqParams {
  const int32_t* mult;   // [C] fixed-point multipliers (per_channel)
  const int32_t* shift;  // [C]
  const int32_t* zp_in;  // [C] input zero points or nullptr
  const double* dmul;    // [C] double multipliers for the float compute dtypes
  int32_t mult0, shift0, zp_in0, zp_out;
  double dmul0;
  int32_t qmin, qmax;  // clamp range of the requantize out_dtype (skipped when no_clamp)
  int per_channel, skip_mul, rounding, compute, no_clamp;
};

enum EpiMode : int { kEpiRaw = 0, kEpiInt = 1, kEpiFloat = 2 };

struct EpiParams {
  int mode;             // EpiMode
  const int32_t* bias;  // [N] int32 or nullptr (kEpiInt/kEpiRaw); float bits for kEpiFloat
  RqParams rq;          // conv -> requantize
  int32_t clip_lo, clip_hi;  // relay clip after requantize (kEpiInt)
  int out_dtype;             // final store dtype
  // residual qnn.add (sum): lhs = this chain's value, rhs = sum tensor
  const void* sum;  // nullptr when absent; same shape/layout as the output
  int sum_dtype;
  RqParams lhs, rhs;  // RequantizeOrUpcast of each side; skip flags inside
  int lhs_identity, rhs_identity;
  int32_t sum_zp_out, sum_qmin, sum_qmax, sum_clip_lo, sum_clip_hi;
  int sum_sub;  // 1: qnn.subtract (lhs' - rhs' + zp_out, subtract.cc:40-93) instead of qnn.add (lhs' + rhs' - zp_out)
  // float (oneDNN-style) mode
  const float* o_scl;  // [N] or [1]
  int n_oscl;
  float act_scl, sum_scl, dst_zp;
  int32_t f_qmin, f_qmax;
};

__device__ __forceinline__ int32_t q_multiply_shift(int32_t x, int32_t y, int32_t s) {
  // intrin_rule.cc:165-195 with q = 31
  const int left = s > 0 ? s : 0, right = s > 0 ? 0 : -s;
  long long v = (long long)x;
  v = (long long)((unsigned long long)v << left);
  v = v * (long long)y;
  const int total = right + 31;
  v = v + (1ll << (total - 1));
  v = v >> total;
  return (int32_t)v;
}

__device__ __forceinline__ int32_t fixed_point_multiply_upward(int32_t x, int32_t m, int32_t s) {
  // intrin_rule.cc:197-250: power-of-two shortcut when m == 1<<30 (per-tensor only)
  if (m == (1 << 30)) {
    int e = s - 1;
    if (e > 0) return (int32_t)((uint32_t)x << e);
    if (e == 0) return x;
    e = -e;
    if (e >= 32) return x < 0 ? -1 : 0;
    int32_t r = (int32_t)((uint32_t)x + (1u << (e - 1)));
    return r >> e;
  }
  return q_multiply_shift(x, m, s);
}

__device__ __forceinline__ int32_t fixed_point_multiply_tonearest(int32_t x, int32_t m, int32_t s) {
  const int left = s > 0 ? s : 0, right = s > 0 ? 0 : -s;
  long long v = (long long)x;
  v = (long long)((unsigned long long)v << left);
  v = v * (long long)m;
  const int total = right + 31;
  const long long pos = 1ll << (total - 1);
  v = v + (v >= 0 ? pos : pos - 1);
  v = v >> total;
  return (int32_t)v;
}

__device__ __forceinline__ double round_upward_d(double t) {
  double b = t + 0.5;
  double bi = (double)(long long)b;
  if (!isfinite(t)) return t;
  return (b == bi || b >= 0.0) ? bi : bi - 1.0;
}
__device__ __forceinline__ double round_tonearest_d(double t) {
  if (!isfinite(t)) return t;
  double m = t < 0 ? -1.0 : 1.0;
  return (double)(long long)((t + 0.5 * m) * m) * m;
}
__device__ __forceinline__ float round_upward_f(float t) {
  float b = __fadd_rn(t, 0.5f);
  float bi = (float)(int)b;
  if (!isfinite(t)) return t;
  return (b == bi || b >= 0.0f) ? bi : __fsub_rn(bi, 1.0f);
}
__device__ __forceinline__ float round_tonearest_f(float t) {
  if (!isfinite(t)) return t;
  float m = t < 0 ? -1.0f : 1.0f;
  return __fmul_rn((float)(int)__fmul_rn(__fadd_rn(t, __fmul_rn(0.5f, m)), m), m);
}

// x: int32 input value, c: channel index along the quantization axis.
__device__ __forceinline__ int32_t requantize_one(int32_t x, int c, const RqParams& p) {
  const int32_t zp = p.zp_in ? p.zp_in[c] : p.zp_in0;
  int32_t y;
  if (p.compute == kInt64) {
    int32_t t = (int32_t)((uint32_t)x - (uint32_t)zp);
    if (!p.skip_mul) {
      if (!p.per_channel)
        t = p.rounding == kUpward ? fixed_point_multiply_upward(t, p.mult0, p.shift0)
                                  : fixed_point_multiply_tonearest(t, p.mult0, p.shift0);
      else
        t = p.rounding == kUpward ? q_multiply_shift(t, p.mult[c], p.shift[c])
                                  : fixed_point_multiply_tonearest(t, p.mult[c], p.shift[c]);
    }
    y = (int32_t)((uint32_t)t + (uint32_t)p.zp_out);
  } else if (p.compute == kFloat64) {
    double t = (double)x - (double)zp;
    if (!p.skip_mul) t = p.per_channel ? __dmul_rn(t, p.dmul[c]) : __dmul_rn(p.dmul0, t);
    t = t + (double)p.zp_out;
    t = p.rounding == kUpward ? round_upward_d(t) : round_tonearest_d(t);
    y = (int32_t)t;
  } else {
    float t = __fsub_rn((float)x, (float)zp);
    if (!p.skip_mul) t = p.per_channel ? __fmul_rn(t, (float)p.dmul[c]) : __fmul_rn((float)p.dmul0, t);
    t = __fadd_rn(t, (float)p.zp_out);
    t = p.rounding == kUpward ? round_upward_f(t) : round_tonearest_f(t);
    y = (int32_t)t;
  }
  if (!p.no_clamp) y = min(max(y, p.qmin), p.qmax);
  return y;
}

__device__ __forceinline__ int32_t cast_trunc(int32_t v, int dt) {
  // relay cast int32 -> narrower int is an LLVM trunc
  switch (dt) {
    case kI8: return (int32_t)(int8_t)v;
    case kU8: return (int32_t)(uint8_t)v;
    case kI16: return (int32_t)(int16_t)v;
    default: return v;
  }
}

__device__ __forceinline__ int32_t load_as_i32(const void* p, int dt, long long i) {
  switch (dt) {
    case kI8: return ((const int8_t*)p)[i];
    case kU8: return ((const uint8_t*)p)[i];
    case kI16: return ((const int16_t*)p)[i];
    default: return ((const int32_t*)p)[i];
  }
}

__device__ __forceinline__ void store_from_i32(void* p, int dt, long long i, int32_t v) {
  switch (dt) {
    case kI8: ((int8_t*)p)[i] = (int8_t)v; break;
    case kU8: ((uint8_t*)p)[i] = (uint8_t)v; break;
    case kI16: ((int16_t*)p)[i] = (int16_t)v; break;
    default: ((int32_t*)p)[i] = v; break;
  }
}

// qnn.add of two already-loaded values (add.cc:40-93).
__device__ __forceinline__ int32_t qnn_add_one(int32_t a, int32_t b, int c, const EpiParams& e) {
  int32_t qa = e.lhs_identity ? a : requantize_one(a, c, e.lhs);
  int32_t qb = e.rhs_identity ? b : requantize_one(b, c, e.rhs);
  int32_t y = e.sum_sub ? (int32_t)((uint32_t)qa - (uint32_t)qb + (uint32_t)e.sum_zp_out)
                        : (int32_t)((uint32_t)qa + (uint32_t)qb - (uint32_t)e.sum_zp_out);
  return min(max(y, e.sum_qmin), e.sum_qmax);
}

// Full fused chain on one int32 accumulator.  `acc` excludes the bias; `n` is the output channel,
// `idx` the linear output index (for the residual tensor).  Returns the value to store (already in
// the final dtype's range / truncated).
__device__ __forceinline__ int32_t epilogue_one(int32_t acc, int n, long long idx, const EpiParams& e) {
  if (e.mode == kEpiFloat) {
    float v = (float)acc;
    if (e.bias) v = __fadd_rn(v, __int_as_float(e.bias[n]));
    v = __fmul_rn(v, e.n_oscl == 1 ? e.o_scl[0] : e.o_scl[n]);
    v = fminf(fmaxf(v, 0.f), 255.f);
    v = __fmul_rn(v, e.act_scl);
    if (e.sum) v = __fadd_rn(v, __fmul_rn(e.sum_scl, (float)load_as_i32(e.sum, e.sum_dtype, idx)));
    v = __fadd_rn(v, e.dst_zp);
    float r = nearbyintf(v);
    return r < (float)e.f_qmin ? e.f_qmin : (r > (float)e.f_qmax ? e.f_qmax : (int32_t)r);
  }
  int32_t x = e.bias ? (int32_t)((uint32_t)acc + (uint32_t)e.bias[n]) : acc;
  if (e.mode == kEpiRaw) return x;
  int32_t y = requantize_one(x, n, e.rq);
  y = min(max(y, e.clip_lo), e.clip_hi);
  if (e.sum) {
    y = cast_trunc(y, e.sum_dtype);  // the cast between clip and qnn.add
    y = qnn_add_one(y, load_as_i32(e.sum, e.sum_dtype, idx), n, e);
    y = min(max(y, e.sum_clip_lo), e.sum_clip_hi);
  }
  return cast_trunc(y, e.out_dtype);
}

// ------------------------------------------------------------------------------------------------
// Fast fixed-point epilogue (tensor-core kernels).  Exactly the same integers as epilogue_one for the
// default configuration (UPWARD rounding, int64 compute, requantize multiplier < 0.5 so that the total
// right shift T = 31 - exponent is >= 32), with all per-channel work folded on the host into ONE 64-bit
// multiply-add and one 32-bit shift per element:
//
//   reference:  t = acc + bias - zp_in  (int32 ring)       v = int64(t) * M + 2^(T-1)
//               y = int32(v >> T) + zp_out                   y = clamp(y, qmin, qmax); y = clip(y, lo, hi)
//   here:       v = int64(acc) * M[c] + C[c],   C = (bias - zp_in) * M + 2^(T-1) + (zp_out << T)
//               y = hi32(v) >> (T[c] - 32);     y = clamp(y, max(qmin,lo), min(qmax,hi))
//
// floor((v + r)/2^T) + z == floor((v + r + z*2^T)/2^T); the host (runtime.cu build_fast_epi) proves from
// sum_k |w[c][k]| that neither the reference's int32 ring nor the 64-bit fold can overflow for any input,
// and only then enables this path -- otherwise the kernel runs epilogue_one.  The residual qnn.add is
// folded the same way: each side is one multiply-add and a 64-bit shift with per-tensor constants (its
// operands are 8-bit, so nothing can overflow); an identity side is M = 1, T = 0.
struct FastSide {        // q' = hi32(int64(q << ls) * M + C) >> s1   (host-normalised so that the shift is >= 32)
  int32_t M, ls, s1, pad;
  long long C;
};
struct FastEpi {
  int enabled;
  const int4* tabA;      // [N] {M, T - 32, C_lo, C_hi}
  const int4* tab_host;  // HOST copy of tabA (never dereferenced on the device): kernels with <= 64 output channels take the
                         // table through their parameter block (constant bank) instead of shared memory
  int32_t lo, hi;        // merged clamp of the requantize out_dtype and the relay clip
  int has_sum;
  uint32_t sum_xor;      // 0x80808080 for an int8 residual (b = (u ^ 0x80) - 128, the -128 is folded into rhs.C)
  FastSide lhs, rhs;     // qnn.add sides; lhs.C also carries "- zp_out" of the add
  int32_t sum_lo, sum_hi;
  // 1 / 2: the final clamp equals the uint8 / int8 range, so the saturating pack (cvt.pack.sat) performs it;
  // `sat` for the plain composite ([lo, hi]), `sum_sat` for the residual form ([sum_lo, sum_hi])
  int sat, sum_sat;
};

}  // namespace tkb
