// Internal host-side launch API of the CUDA kernels (kernels_simt.cu, kernels_tc.cu).
// All activations are channels-last ("NHWC") unless a stride array says otherwise.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "epilogue.cuh"

namespace tkb {

struct ConvGeom {
  int N, H, W, C;          // input, channels-last
  int O, KH, KW, groups;   // weights OHWI: [O][KH][KW][C/groups]
  int sh, sw, pt, pl, pb, pr, dh, dw;
  int OH, OW;
  int a_dtype, w_dtype;    // kI8 / kU8
  int32_t zp_a;            // input zero point (padding value)
  const int32_t* zp_w;     // device [O] or nullptr -> zp_w0
  int32_t zp_w0;
};

// ---- CUDA-core kernels (complete coverage; the fallback for every shape) ----------------------
void launch_conv2d_simt(const void* x, const void* w, void* y, const ConvGeom& g, const EpiParams& e,
                        cudaStream_t s);
// depthwise convolution on dense channels-last 8-bit tensors; w_tap: [KH*KW][C] (tap-major), corr[c] = -zp_a*sum_taps w
bool dwconv_nhwc_supported(const ConvGeom& g);
// w_row3 (3x3 kernels, may be null): [3][C] words {w(r,0), w(r,1), w(r,2), 0} for the dp4a kernel
// pointwise convolution with 8 / 16 / 24 / 32 input channels on CUDA cores (kernels_simt.cu pwconv_kernel); 0 on success
bool pwconv_supported(int K, int N);
int launch_pwconv(const void* x, const void* w, void* y, long long M, int N, int K, int a_dtype, int w_dtype, const FastEpi& f,
                  cudaStream_t s);
void launch_dwconv_nhwc(const void* x, const void* w_tap, const void* w_row3, const int32_t* corr, void* y,
                        const ConvGeom& g, const EpiParams& e, const FastEpi& f, cudaStream_t s);
// Y[b,m,n] = sum_k (X[b,m,k]-zpx)(W[b?,n,k]-zpw); w_batched: W has a batch dim of B, else shared.
void launch_gemm_simt(const void* x, const void* w, void* y, int B, int M, int N, int K, int w_batched,
                      int x_dtype, int w_dtype, int32_t zpx, int32_t zpw, const EpiParams& e, cudaStream_t s);
// elementwise over [outer][C][inner]
void launch_requantize(const void* x, int x_dtype, void* y, int y_dtype, long long outer, int C, long long inner,
                       const RqParams& p, cudaStream_t s);
void launch_qnn_add(const void* a, const void* b, void* y, int dtype, long long n, const EpiParams& e,
                    cudaStream_t s);
// qnn.mul (per tensor): y = requantize((a - zpa) * (b - zpb)) with rq built for scale_a * scale_b -> output qparams
void launch_qnn_mul(const void* a, const void* b, void* y, int dtype, long long n, int32_t zpa, int32_t zpb,
                    const RqParams& rq, cudaStream_t s);
// qnn.quantize / qnn.dequantize over [outer][C][inner]; parameters per tensor (n == 1: scale0 / zp0) or per channel
// (device arrays of C entries)
void launch_quantize(const float* x, void* y, int y_dtype, long long outer, int C, long long inner, const float* scale,
                     const int32_t* zp, float scale0, int32_t zp0, int n_scale, int n_zp, int32_t lo, int32_t hi, cudaStream_t s);
void launch_dequantize(const void* x, int x_dtype, float* y, long long outer, int C, long long inner, const float* scale,
                       const int32_t* zp, float scale0, int32_t zp0, int n_scale, int n_zp, cudaStream_t s);
// qnn.concatenate: copy (requant == 0) or requantize one input [outer][row_in] into rows of row_out elements at `off`
void launch_concat_piece(const void* x, void* y, int dtype, long long outer, long long row_in, long long row_out, long long off,
                         const RqParams& rq, int requant, cudaStream_t s);
// take(table, 8-bit indices): y[i] = table[x[i]] for a table of `entries` <= 256 elements of `elem` = 1 / 2 / 4 bytes
void launch_lut8(const void* x, const void* table, int entries, int elem, void* y, long long n, cudaStream_t s);
// qnn.leaky_relu, per tensor (leaky_relu.cc:70-139)
void launch_qnn_leaky_relu(const void* x, void* y, int dtype, long long n, const RqParams& rq, int identity, int32_t zp_in,
                           int32_t m_alpha, int32_t s_alpha, int32_t scaled_z, int32_t lo, int32_t hi, cudaStream_t s);
int32_t device_fixed_point_multiply(int32_t x, int32_t m, int32_t s);  // one value, synchronous (init time)
void launch_clip(const void* x, void* y, int dtype, long long n, double lo, double hi, cudaStream_t s);
void launch_cast(const void* x, int x_dtype, void* y, int y_dtype, long long n, cudaStream_t s);
// y = x + b[c] (int32 wrap / same dtype) over [outer][C][inner]; b has C elements of dtype b_dtype
void launch_bias_add(const void* x, const void* b, void* y, int dtype, long long outer, int C, long long inner,
                     cudaStream_t s);
void launch_add_same(const void* a, const void* b, void* y, int dtype, long long n, cudaStream_t s);
// pooling with explicit element strides (n,c,h,w) for input and output; is_avg: 0 max, 1 windowed average,
// 2 adaptive / global average (divisor cast to the tensor dtype, include/tvm/topi/nn/pooling.h:384-389)
// cast(8-bit -> int32) -> global average pool -> cast(-> 8-bit) in one launch over channels-last data [N][HW][C], C % 4 == 0;
// cast_out (the int32 copy of x, [N][HW][C]) may be nullptr
void launch_upcast_gap(const void* x, int x_dtype, void* cast_out, void* pooled, void* out8, int N, int HW, int C, cudaStream_t s);
void launch_pool2d(const void* x, void* y, int dtype, int is_avg, int count_include_pad, int N, int C, int H, int W,
                   int KH, int KW, int sh, int sw, int pt, int pl, int pb, int pr, const long long* xs,
                   const long long* ys, cudaStream_t s);
// generic 4-D permute copy: out[i0,i1,i2,i3] (dense) = in[... strides]; elem = 1/2/4 bytes
void launch_permute4(const void* x, void* y, int elem, const long long* out_shape, const long long* in_strides,
                     cudaStream_t s);
// tiled 8-bit channels-last -> NCHW transpose; returns false (nothing launched) if the shape needs launch_permute4
bool launch_nhwc_to_nchw8(const void* x, void* y, long long N, long long HW, long long C, cudaStream_t s);
// [N][A][B] -> [N][B][A] for 1/2/4-byte elements (32 x 32 tiles); false if nothing could be launched
bool launch_transpose_tile(const void* x, void* y, int elem, long long N, long long A, long long B, cudaStream_t s);
// constant pad of a dense tensor up to rank 4 (shape padded to 4 dims)
void launch_pad4(const void* x, void* y, int elem, const long long* in_shape, const long long* pad_before,
                 const long long* out_shape, int32_t value, cudaStream_t s);

// im2col into a [N*OH*OW][Kp] matrix (Kp >= KH*KW*C, multiple of 16) for the tensor-core GEMM path
void launch_im2col(const void* x, void* a, const ConvGeom& g, int Kp, cudaStream_t s);
// same matrix built from a logical NCHW tensor, K ordered (c, r, s) like OIHW weights; KH, KW <= 255
void launch_im2col_nchw(const void* x, void* a, const ConvGeom& g, int Kp, cudaStream_t s);

// vertical unfold of an NCHW tensor with C*KH <= 32 into [N][OH][W][32] (see kernels_simt.cu)
void launch_unfold_rows_nchw(const void* x, void* y, const ConvGeom& g, cudaStream_t s);

// ---- tcgen05 kernels ---------------------------------------------------------------------------
// Persistent GEMM / implicit-GEMM convolution on the 5th-gen tensor cores:
//   Y[b][M][N] = epilogue(A[b][M][K] . W[b?][N][K]^T), A/W int8 or uint8, K-major, zp_w == 0.
// implicit != 0: `a` is the NHWC activation tensor of geometry `g` and the A operand is gathered by an
// im2col-mode TMA tensor map (K = KH*KW*C ordered tap-major, i.e. OHWI weights); taps in the padding read
// 0, so the caller supplies the per-border-class zero-point correction `delta` (see runtime.cu).
struct TcArgs {
  const void* a;
  const void* w;
  void* y;
  int B, M, N, K, w_batched, a_dtype, w_dtype;
  int implicit;
  ConvGeom g;
  const int32_t* delta;   // [n_classes][N] int32, class 0 = interior (all zeros, never read); or nullptr
  const uint8_t* cls_h;   // [OH] class of each output row
  const uint8_t* cls_w;   // [OW] class of each output column
  int ncls_w;
  EpiParams e;
  FastEpi f;
  // Constant weights may be handed over pre-tiled ("packed"): for every (n-tile, k-block) the BN x bk tile in
  // its final swizzled shared-memory image, so that the kernel fetches it with one bulk copy instead of one
  // TMA request per row.  w_packed != 0: `w` points at tc_pack_weights() output made for (pack_bn, pack_bk).
  int w_packed, pack_bn, pack_bk;
};
// Tile configuration launch_tc will use for this problem (bn: N tile, bk: bytes of K per k-block).
void tc_tile_config(const TcArgs& a, int* bn, int* bk);
// [N][K] K-major weights -> packed tiles for (bn, bk); returns the byte size (dst may be nullptr to query).
size_t tc_pack_weights(const uint8_t* w, int N, int K, int bn, int bk, uint8_t* dst);
// Requirements of the plain-GEMM form: K % 16 == 0, pointers 16B aligned.
bool tc_gemm_supported(int M, int N, int K);
// Requirements of the implicit form: groups == 1, C % 32 == 0, TMA corner / offset ranges.
bool tc_implicit_conv_supported(const ConvGeom& g);
// halo-reuse convolution kernel (kernels_tc_halo.cu): TcArgs.implicit == 3
bool tc_halo_supported(const ConvGeom& g, int N);
void tc_halo_tile_config(const ConvGeom& g, int N, int* bn, int* bk);
int launch_tc_halo(const TcArgs& a, cudaStream_t s);
int launch_tc(const TcArgs& a, cudaStream_t s);
// cta_group::2 pair kernel (kernels_tc_pair.cu): 256 x BN tiles on two CTAs of a cluster, B halved per CTA; fast
// epilogues, unpacked K-major weights, plain GEMM or im2col-mode implicit convolution
bool tc_pair_supported(const TcArgs& a);
int tc_pair_bn(const TcArgs& a);
int launch_tc_pair(const TcArgs& a, cudaStream_t s);

// ---- chained launch (kernels_tc_chain.cu): several consecutive fast-epilogue layers in ONE persistent kernel, tiles of
// layer k+1 gated per image on layer k's completion counters.  A layer is a TcArgs of the implicit-conv (implicit == 1)
// or plain-GEMM (implicit == 0) form with weights packed for (bn, tc_chain_bk()).
#define TKB_CHAIN_MAX_SEG 16
struct TcChainLayer {
  TcArgs a;
  int rows_per_img;     // output rows (pixels) per image: the granularity of the completion counters
  int dep_a, dep_sum;   // index of the layer (in this chain) that produces the A operand / the residual, or -1
};
int tc_chain_bk(const TcArgs& a);
bool tc_chain_supported(const TcArgs& a, int bn);
int tc_chain_flag_ints(const TcChainLayer* layers, int n);   // ints of device memory for the claim counter + counters
int tc_chain_item_count(const TcChainLayer* layers, int n, int bn);  // entries of the work list
size_t tc_chain_state_bytes();
}  // namespace tkb
#include <vector>
namespace tkb {
// Encode the launch state (tensor maps, per-layer parameters) into `state`; `flags` / `tgt_dev`: device arrays of
// tc_chain_flag_ints() ints (flags zeroed before every launch; tgt_dev filled by the caller from *tgt_host);
// `items_dev`: device array of tc_chain_item_count() ints filled by the caller from *items_host (the work list).
int tc_chain_build(const TcChainLayer* layers, int n, int bn, int* flags, int* tgt_dev, std::vector<int>* tgt_host,
                   const int* items_dev, std::vector<int>* items_host, void* state);
int launch_tc_chain(const void* state, int bn, cudaStream_t s);
void tc_chain_stats(unsigned long long* out, bool reset);  // debug counters of the chained kernel (see g_chain_stats)
int probe_int8_peak(int iters, double* tops, cudaStream_t s);
int probe_int8_shape(int iters, int n, int commit_every, double* tops, cudaStream_t s);

void launch_spin(double us, cudaStream_t s);  // device busy-wait (measurement aid)

// ---- float32 operators of the float tachikoma patterns (kernels_float.cu) ------------------------------------------------
enum FAct : int { kFActNone = 0, kFActRelu, kFActTanh, kFActSigmoid, kFActClip, kFActGelu, kFActSwish, kFActMish, kFActAbs,
                  kFActExp, kFActLog, kFActSqrt, kFActRound, kFActLeakyRelu };
// post-ops of a float primitive in the reference's order (tachikoma_json_runtime.cc:185-233): + bias[channel], + sum, activation
struct FPost { int act; float a, b; const float* bias; const float* sum; };
// NC[D]HW convolution / pooling geometry (1-D and 2-D operators are the D = 1 (and H = 1) cases)
struct FConvGeom {
  int N, C, D, H, W, O, OD, OH, OW, KD, KH, KW, sd, sh, sw, dd, dh, dw, pd, ph, pw, groups, transposed;
};
struct FBcast5 { long long shape[5], ls[5], rs[5]; };
void launch_fconv(const float* x, const float* w, float* y, const FConvGeom& g, const FPost& post, cudaStream_t s);
bool fconv_gemm_form(const FConvGeom& g);  // launch_fconv takes the implicit-GEMM kernel (else the direct one)
void launch_fgemm(const float* A, const float* B, float* C, int batch, int M, int N, int K, long long a_b, long long a_m,
                  long long a_k, long long b_b, long long b_n, long long b_k, const FPost& post, cudaStream_t s);
void launch_funary(const float* x, float* y, long long n, int act, float a, float b, cudaStream_t s);
void launch_fbinary(const float* l, const float* r, float* y, const FBcast5& bc, int mul, cudaStream_t s);
void launch_fbatchnorm(const float* x, float* y, long long outer, int C, long long inner, const float* gamma, const float* beta,
                       const float* mean, const float* var, float eps, cudaStream_t s);
void launch_flayernorm(const float* x, float* y, long long rows, int C, const float* gamma, const float* beta, float eps,
                       cudaStream_t s);
void launch_fsoftmax(const float* x, float* y, long long outer, int C, long long inner, cudaStream_t s);
void launch_fpool(const float* x, float* y, const FConvGeom& g, int is_avg, int count_include_pad, cudaStream_t s);

extern thread_local long long g_launches;  // kernels launched by this thread (summed per handle)

}  // namespace tkb
