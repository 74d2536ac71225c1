/*
 * tkb200.h -- C ABI of libtkb200.so, the B200-native drop-in for the tachikoma BYOC JSON runtime.
 *
 * Every entry point replaces one interface of the reference (paths relative to the
 * CortexFoundation/tachikoma tree):
 *
 *   tkb_create            runtime.TachikomaJSONRuntimeCreate(symbol, graph_json, const_names)
 *                         src/runtime/contrib/tachikoma/tachikoma_json_runtime.cc:900-906
 *   tkb_load_binary       runtime.module.loadbinary_tachikoma_json          :908-909
 *   tkb_save_binary       JSONRuntimeBase::SaveToBinary   src/runtime/contrib/json/json_runtime.h:105-115
 *   tkb_get_symbol        PackedFunc "get_symbol"                           json_runtime.h:74-76
 *   tkb_get_const_vars    PackedFunc "get_const_vars"                       json_runtime.h:77-79
 *   tkb_init              PackedFunc "__init_<symbol>"(Array<NDArray>)      json_runtime.h:89-99,
 *                         TachikomaJSONRuntime::Init + BuildEngine  tachikoma_json_runtime.cc:64-71,:240-290
 *   tkb_run               PackedFunc "<symbol>"(in_0..in_k-1, out_0..out_m-1), thread-safe Run(args) const
 *                         tachikoma_json_runtime.cc:77-106
 *   tkb_last_error        TVMGetLastError()               include/tvm/runtime/c_runtime_api.h:238-249
 *   tkb_packed_run        TVMBackendPackedCFunc-shaped shim include/tvm/runtime/c_backend_api.h:49-51
 *   tkb_trace_*           the per-operator trace the reference describes (README.md:6) but does not
 *                         ship; container = TVM NDArray-list blob, src/runtime/file_utils.cc:210-229,
 *                         include/tvm/runtime/ndarray.h:449-494 (see DESIGN.md "Trace").
 *
 * Conventions (same as the reference's C API): every function returns 0 on success and -1 on
 * failure; the message is available from tkb_last_error() (thread local).  Unsupported operators
 * fail in tkb_init ("Unsupported op: <name>", tachikoma_json_runtime.cc:286) -- there is no CPU
 * fallback.  Tensors are DLPack DLTensor (3rdparty/dlpack/include/dlpack/dlpack.h:187-200), must be
 * compact (strides == NULL, tachikoma_json_runtime.cc:820) and may live on the host (kDLCPU,
 * kDLCUDAHost) or on the current CUDA device (kDLCUDA); host tensors are staged through pinned
 * buffers with cudaMemcpyAsync on the caller's stream.
 */
#ifndef TKB200_H_
#define TKB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* --- DLPack (ABI-compatible restatement so that no TVM/DLPack header is needed) ------------- */
typedef enum { tkbDLCPU = 1, tkbDLCUDA = 2, tkbDLCUDAHost = 3 } tkb_device_type;
typedef struct { int32_t device_type; int32_t device_id; } tkb_DLDevice;
typedef struct { uint8_t code; uint8_t bits; uint16_t lanes; } tkb_DLDataType; /* code: 0 int, 1 uint, 2 float */
typedef struct {
  void* data;
  tkb_DLDevice device;
  int32_t ndim;
  tkb_DLDataType dtype;
  int64_t* shape;
  int64_t* strides; /* must be NULL (compact row-major) */
  uint64_t byte_offset;
} tkb_DLTensor;

typedef struct tkb_runtime* tkb_handle;

/* --- module life cycle ---------------------------------------------------------------------- */
int tkb_create(const char* symbol, const char* graph_json, const char* const* const_names, int n_consts,
               tkb_handle* out);
int tkb_free(tkb_handle h);
/* blob = dmlc stream {string symbol; string graph_json; vector<string> const_names} (u64 length prefixes) */
int tkb_save_binary(tkb_handle h, const void** blob, size_t* len); /* blob owned by h */
int tkb_load_binary(const void* blob, size_t len, tkb_handle* out);
const char* tkb_type_key(void); /* "tachikoma_json" */

/* --- module functions ------------------------------------------------------------------------ */
const char* tkb_get_symbol(tkb_handle h);
int tkb_get_const_vars(tkb_handle h, const char* const** names, int* n);
int tkb_num_inputs(tkb_handle h);
int tkb_num_outputs(tkb_handle h);
/* Bind constants (in const_names order), build the execution plan, upload + repack weights.
 * Idempotent and mutex-guarded like "__init_<symbol>". */
int tkb_init(tkb_handle h, const tkb_DLTensor* const* consts, int n_consts);
/* args = inputs (graph input order) followed by caller-allocated outputs; n_args must equal
 * #inputs + #outputs.  cuda_stream is a cudaStream_t (NULL = default stream).  Re-entrant from
 * several host threads provided each uses its own stream.  With host outputs the call returns when the outputs are
 * valid; a trace being streamed to the host (mode 2) may still be in flight then: tkb_trace_sync_on / tkb_trace_blob_on
 * wait for it, and the next run on the same stream is ordered after it on the device.  Each stream has its own
 * workspace and pinned trace buffer, so one host thread may keep two batches in flight by alternating two streams. */
int tkb_run(tkb_handle h, const tkb_DLTensor* const* args, int n_args, void* cuda_stream);
/* TVMBackendPackedCFunc-shaped entry: args[0] = module handle (void*), args[1..] = DLTensor*.
 * type codes: 3 = kTVMOpaqueHandle, 7 = kTVMDLTensorHandle, 13 = kTVMNDArrayHandle. */
int tkb_packed_run(void* args, int* type_codes, int num_args, void* out_ret_value, int* out_ret_tcode,
                   void* resource_handle);

const char* tkb_last_error(void);

/* --- options / introspection ------------------------------------------------------------------ */
/* Known keys: "conv_impl" = "auto" (tensor cores where they pay) | "simt" (CUDA-core kernels only) | "tc" (tensor
 *                           cores wherever supported) | "tc_im2col" (as "tc" without the halo-reuse convolution
 *                           kernel: TMA-im2col implicit GEMM) | "tc_explicit" (as "tc" with an explicit im2col
 *                           matrix) | "tc_chain_all" / "tc_chain_siblings" / "tc_chain_split" (as "tc" with consecutive
 *                           layers merged into chained persistent launches, csrc/kernels_tc_chain.cu: dependent layers
 *                           too / layers reading the same activation / every layer, a launch per dependency) |
 *                           "tc_pair" (as "tc" with the cta_group::2 pair kernel, csrc/kernels_tc_pair.cu, for layers with
 *                           a multiple of 128 output channels) -- all
 *                           routes produce identical bits, everything after "tc" exists for A/B tests;
 *             "time_tc"   = "0"|"1" (bracket every run of consecutive tcgen05 launches with CUDA events on the compute
 *                           stream, see tkb_tc_time). */
int tkb_set_option(tkb_handle h, const char* key, const char* value);
/* Number of kernels this library launched on behalf of h since creation (bench.py "gpu_launches"). */
int64_t tkb_launch_count(tkb_handle h);
/* Human-readable plan: one line per kernel node with the kernel chosen and its algorithmic work. */
const char* tkb_plan_summary(tkb_handle h);
/* With option time_tc=1: summed device time (ms) of the tcgen05 launches of the last run and their count.  Launches that
 * follow one another with nothing else launched in between share one event pair (an event between two kernels costs
 * 2-4 us and breaks their programmatic-dependent-launch overlap, which is not the kernels' time). */
int tkb_tc_time(tkb_handle h, double* ms, int* launches);
/* Total int8 multiply-accumulates and algorithmic HBM bytes of one run (roofline numerators). */
int tkb_work(tkb_handle h, int64_t* macs, int64_t* hbm_bytes);
/* int8 multiply-accumulates of the plan's tcgen05 launches only (the launches tkb_tc_time times): depthwise and
 * few-channel pointwise convolutions run on CUDA cores and are counted by tkb_work but not here.  No reference counterpart
 * (measurement support for bench.py's roofline). */
int64_t tkb_tc_macs(tkb_handle h);

/* --- per-operator trace ------------------------------------------------------------------------
 * mode 0 = off;
 *      1 = device-resident capture: every kernel node's inputs / outputs exist in their LOGICAL layout on the
 *          device when the run ends (channels-last intermediates are permuted into a device ring on a side
 *          stream); the trace stays readable until the next run on the same stream and is copied out on demand
 *          by tkb_trace_blob / tkb_trace_save (deferred device -> host read-out);
 *      2 = additionally stream every distinct tensor to pinned host memory on the side stream during the run.
 * The pinned buffer IS the serialized NDArray-list blob (names and per-array headers are written once); after a
 * run, tkb_trace_blob returns it without re-serializing.  Trace state is per stream: tkb_run uses the handle's
 * default mode (tkb_trace_enable), tkb_run_traced passes the mode per call, so host threads on distinct streams
 * may trace differently at the same time (Run(args) const, tachikoma_json_runtime.cc:76-90). */
int tkb_trace_enable(tkb_handle h, int mode);
int tkb_run_traced(tkb_handle h, const tkb_DLTensor* const* args, int n_args, void* cuda_stream, int trace_mode);
int tkb_trace_sync(tkb_handle h);                         /* stream of the most recent run */
int tkb_trace_sync_on(tkb_handle h, void* cuda_stream);
int tkb_trace_save(tkb_handle h, const char* path);
/* Serialized trace of the last run on the stream; the pointer is owned by h and valid until the next traced run on
 * that stream.  tkb_trace_blob uses the stream of the most recent run. */
int tkb_trace_blob(tkb_handle h, const void** blob, size_t* len);
int tkb_trace_blob_on(tkb_handle h, void* cuda_stream, const void** blob, size_t* len);
/* Debug-executor-compatible dump (python/tvm/contrib/debugger/debug_result.py:25, :114-149, :191-202) of the last
 * run: <dir>/output_tensors.params (NDArray-list blob, one record per node OUTPUT, keys
 * "{name}____topo-index:{i}____output-num:{j}") and <dir>/_tvmdbg_graph_dump.json ({"nodes": [{op, name, inputs,
 * attrs, shape}], "arg_nodes", "node_row_ptr", "heads", "attrs"}). */
int tkb_trace_dump_debug(tkb_handle h, const char* dir);
int64_t tkb_trace_bytes(tkb_handle h); /* size of the serialized blob */
/* Bytes copied device -> host per traced run: every distinct activation once (a tensor that is one node's
 * output and another node's input is staged once and serialized twice); constants stay on the host. */
int64_t tkb_trace_d2h_bytes(tkb_handle h);

/* --- stand-alone helpers (used by tests and tools) -------------------------------------------- */
/* GetFixedPointMultiplierShift, src/relay/qnn/utils.cc:33-57 (host, double). */
void tkb_fixed_point_multiplier_shift(double multiplier, int32_t* significand, int32_t* exponent);
/* Dense int8 tcgen05 peak probe: issues back-to-back kind::i8 MMAs from resident smem tiles on all
 * SMs; returns achieved tera-ops/s (2*MAC) in *tops.  Roofline denominator for conv/dense. */
int tkb_probe_int8_peak(int iters, double* tops, void* cuda_stream);
/* Same probe with UMMA N = n (multiple of 16, <= 256) and a tcgen05.commit every `commit_every` groups of 4 MMAs
 * (0 = never): how the issue rate depends on the instruction shape and on per-stage commits. */
int tkb_probe_int8_shape(int iters, int n, int commit_every, double* tops, void* cuda_stream);
/* Measurement aid: enqueue a device busy-wait of about `microseconds` on the stream, so that a host can queue a
 * whole step behind it and per-kernel CUDA events then see device time only. */
int tkb_spin(double microseconds, void* cuda_stream);
/* Debug counters of the chained tcgen05 launches since the last reset: out8[0] = work items that had to wait for a
 * dependency, out8[1] = cycles waited (summed over CTAs); tools/chain_probe.py prints them. */
int tkb_debug_chain_stats(uint64_t* out8, int reset);
/* --- allocation-free entry point --------------------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * One qnn.conv2d (integer semantics of src/relay/qnn/op/convolution.cc, groups == 1):
 *   out[n,o,p,q] = sum_{c,r,s} (data[n,c,p*sh-pt+r*dh,q*sw-pl+s*dw] - input_zero_point) * (weight[o,c,r,s] - kernel_zero_point)
 * (int32, wrapping; taps outside the image contribute 0), computed as the runtime's explicit-im2col variant: the
 * [N*OH*OW][Kp] column matrix (Kp = C*KH*KW rounded up to 16) is staged in a file-scope static __device__ array of
 * TKB_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit, then multiplied with the weights.
 * `data` (NCHW, int8 or uint8), `weight` (OIHW, int8 or uint8) and `out` (NCHW int32) are DEVICE pointers of the
 * current device; dtypes are DLPack codes (0 = int, 1 = uint; 8 bits).  The call allocates nothing (no cudaMalloc /
 * cudaFree / allocating library routine) and only enqueues work on `cuda_stream`.
 * Errors (-1, tkb_last_error): N*OH*OW*Kp > TKB_NOALLOC_SCRATCH_BYTES, Kp > TKB_NOALLOC_MAX_K, kernel sides > 255,
 * non-positive sizes, bad dtype codes.
 * ALL calls on a device share the one scratch array, which keeps the previous call's contents (each call overwrites
 * the part it uses before reading it; nothing relies on zeros): calls that overlap on the GPU (two streams or two
 * host threads of one process) race and must be serialized by the caller.  Each process has its own copy. */
#define TKB_NOALLOC_SCRATCH_BYTES (16 << 20)
#define TKB_NOALLOC_MAX_K 12288
typedef struct {
  int n, c, h, w;                                 /* data [n][c][h][w] */
  int o, kh, kw;                                  /* weight [o][c][kh][kw] */
  int stride_h, stride_w, dilation_h, dilation_w;
  int pad_top, pad_left, pad_bottom, pad_right;
  int data_dtype_code, weight_dtype_code;         /* 0 = int8, 1 = uint8 */
  int32_t input_zero_point, kernel_zero_point;
} tkb_conv2d_desc;
int tkb_qnn_conv2d_noalloc(const tkb_conv2d_desc* desc, const void* data, const void* weight, int32_t* out,
                           void* cuda_stream);

/* Library build info: "sm_100a;<n kernels>;..." */
const char* tkb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* TKB200_H_ */
