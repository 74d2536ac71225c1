// libtkb200 runtime: the B200-native replacement of TachikomaJSONRuntime
// (src/runtime/contrib/tachikoma/tachikoma_json_runtime.cc:52-124 Init/Run/GetFunction, :240-290
// BuildEngine, :142-237 ParseAttrs) behind the C ABI declared in include/tkb200.h.
//
//   tkb_create : parse the JSON graph (graph.h)            <- TachikomaJSONRuntimeCreate
//   tkb_init   : bind constants, build the launch plan      <- Init + BuildEngine
//   tkb_run    : enqueue the plan on a CUDA stream          <- Run(args) const
//
// Internal tensor layout: 4-D activations of NCHW graphs are kept channels-last ("NHWC physical")
// between spatial operators and converted at the subgraph boundary or where an operator needs the
// logical layout (reshape / flatten / dense).  Every intermediate has its own slot in a per-stream
// arena (no reuse inside a run, so trace capture can read any tensor after the fact).
#include <cuda_runtime.h>

#include <algorithm>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <map>
#include <regex>
#include <set>
#include <sstream>
#include <thread>
#include <unordered_map>

#include "../../include/tkb200.h"
#include "graph.h"
#include "kernels.h"

namespace tkb {

static thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }  // for entry points in other files (noalloc.cu)

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

[[noreturn]] static void fail(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  throw Error(buf);
}

#define CUDA_OK(expr)                                                                      \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) fail("CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__, __LINE__, \
                                cudaGetErrorString(_e));                                   \
  } while (0)

static long long numel(const std::vector<int64_t>& s) {
  long long n = 1;
  for (auto d : s) n *= d;
  return n;
}

enum Phys : int { kLogical = 0, kNHWC = 1 };  // kNHWC: logical NCHW tensor stored as N,H,W,C

struct Entry {
  std::vector<int64_t> shape;
  int dtype = kI32;
  bool is_input = false, is_const = false;
  int input_index = -1;            // position among graph inputs
  std::vector<uint8_t> host;       // constants: host copy (planner reads scales / zero points from it)
  void* dev_const = nullptr;       // constants uploaded on demand (logical layout)
  long long off[2] = {-1, -1};     // arena offsets per physical layout (intermediates, staged inputs)
  int alias_of = -1;               // reshape / flatten: shares the buffer of another entry (logical)
  long long bytes() const { return numel(shape) * dtype_size(dtype); }
  bool nhwc_is_noop() const {      // NHWC and NCHW coincide in memory
    return shape.size() == 4 && (shape[1] == 1 || shape[2] * shape[3] == 1);
  }
};

struct RunCtx;
struct Step {
  std::function<void(RunCtx&)> fn;
  std::string desc;
  int nid = -1;   // JSON node this step belongs to (-1: boundary conversion)
  int cand = -1;  // index into tkb_runtime::chain_cands: a tcgen05 layer that may be merged into a chained launch
  bool is_tc = false;  // launches a tcgen05 layer (kernel timing: consecutive such steps form one timed run)
};

// A tcgen05 convolution / GEMM step in the form the chained launch (kernels_tc_chain.cu) takes.
struct ChainCand {
  TcArgs ta;
  int x_eid, y_eid, sum_eid, phys, rows_per_img, bn;
  std::string name;
};
struct ChainPlan {
  std::vector<int> cands;           // indices into chain_cands, execution order
  std::vector<int> dep_a, dep_sum;  // per layer: producing layer inside the chain or -1
  int bn = 128;
  int flag_off = 0, flag_ints = 0;  // slice of the workspace's counter array
  int item_off = 0, item_count = 0; // slice of the workspace's work-list array
};

struct TraceRec {
  std::string name;
  int eid;
  int nid;
  bool logical = false;  // a logical-layout copy of the tensor exists once node `nid` has run
};

// Everything one run touches lives here, one Workspace per caller stream: the plan is immutable after tkb_init and
// the handle carries no per-call state (tachikoma_json_runtime.cc:76-90: "Run(args) const", instance immutable), so
// host threads that use distinct streams never share mutable memory.
struct Workspace {
  uint8_t* arena = nullptr;
  long long arena_bytes = 0;
  uint8_t* trace_ring = nullptr;    // device: logical copies of NHWC-physical tensors
  uint8_t* trace_host = nullptr;    // pinned host image of the serialized NDArray-list blob (headers written once)
  long long trace_ring_bytes = 0, trace_host_bytes = 0;
  cudaStream_t side = nullptr;
  cudaEvent_t ev_join = nullptr;
  bool side_dirty = false;
  cudaEvent_t ev = nullptr;
  std::vector<cudaEvent_t> tc_events;  // begin/end pairs around tcgen05 GEMM launches (option time_tc)
  size_t tc_used = 0;
  long long tc_launches = 0;           // tcgen05 launches inside the recorded runs
  int trace_mode = 0;               // trace mode of the LAST run on this stream (0 = nothing captured)
  bool payload_on_host = false;     // first occurrence of every traced tensor has been copied into trace_host
  bool dups_filled = false;         // repeated records (a tensor that is also a consumer's input) have been filled in
  std::vector<void*> in_ptr;        // device pointers of the last run's graph inputs (deferred read-out, mode 1)
  int* chain_flags = nullptr;       // device: [counters of all chains | their targets]; the first half is zeroed per run
  std::vector<std::vector<uint8_t>> chain_state;  // per chain: encoded launch parameters (tensor maps point into arena)
};

}  // namespace tkb

struct tkb_runtime {
  std::string symbol, graph_json;
  std::vector<std::string> const_names;
  std::vector<const char*> const_name_ptrs;
  tkb::Graph g;
  std::vector<tkb::Entry> ent;
  std::vector<tkb::Step> steps;
  std::vector<void*> dev_allocs;  // freed in tkb_free
  std::deque<std::vector<int32_t>> host_tabs;  // host copies of the folded epilogue tables (FastEpi.tab_host points into these)
  long long arena_bytes = 0;
  bool initialized = false;
  std::mutex init_mu, ws_mu;
  std::unordered_map<cudaStream_t, std::unique_ptr<tkb::Workspace>> ws;
  std::string conv_impl = "auto";
  std::atomic<bool> time_tc{false};
  std::atomic<int> trace_mode{0};               // DEFAULT mode of tkb_run; tkb_run_traced passes its own per call
  std::vector<tkb::TraceRec> trace_recs;        // in emission order
  std::vector<long long> trace_host_off;        // per record: payload offset inside the serialized blob
  std::vector<long long> trace_ring_off;        // per record (-1: read the tensor in place)
  std::vector<int> trace_dup_of;                // per record: earlier record holding the same tensor, or -1
  long long trace_d2h_bytes = 0;                // bytes copied device -> host per run (each tensor once)
  long long trace_host_bytes = 0, trace_ring_bytes = 0;
  std::vector<std::pair<long long, std::vector<uint8_t>>> blob_headers;  // (offset, bytes): everything but payloads
  std::vector<uint8_t> blob;
  std::string plan_text;
  std::atomic<long long> launches{0};
  long long macs = 0, hbm_bytes = 0;
  long long tc_macs = 0;                       // MACs of the steps that run as tcgen05 launches (what tkb_tc_time times)
  std::vector<tkb::ChainCand> chain_cands;
  std::vector<tkb::ChainPlan> chains;
  int chain_flag_ints = 0;                      // counters of all chains (the workspace allocates twice this: + targets)
  int chain_item_ints = 0;                      // work-list entries of all chains (allocated behind the targets)
  bool chain_enabled = true;
  int chain_policy = 0;                         // 0 off, 1 sibling layers, 2 dependent layers merged, 3 every layer / split
  cudaStream_t last_stream = nullptr;           // guarded by ws_mu; only the stream-less convenience calls read it
};

namespace tkb {

struct RunCtx {
  tkb_runtime* rt;
  Workspace* ws;
  cudaStream_t s;
  int mode = 0;               // trace mode of THIS call
  bool time_tc = false;
  std::vector<void*> in_ptr;  // device pointer of each graph input (logical layout)
  // Kernel timing (option time_tc): CUDA events on the compute stream around every RUN of consecutive tcgen05 launches.
  // Back-to-back tensor-core layers (nothing else launched between them: ResNet's layer1 .. layer4) share one event pair,
  // so the ~2-4 us an event pair costs per launch and the loss of the programmatic-dependent-launch overlap an event
  // between two kernels causes are not billed to the kernels (they were a sixth of the sum with a pair per launch).
  bool tc_pending = false;      // a run is open: its end event is recorded when something else is about to be launched
  long long tc_launch_mark = 0;
  void tc_record() {
    if (ws->tc_used == ws->tc_events.size()) {
      cudaEvent_t ev;
      cudaEventCreate(&ev);
      ws->tc_events.push_back(ev);
    }
    cudaEventRecord(ws->tc_events[ws->tc_used++], s);
  }
  void tc_flush() {  // close the open run (call before launching anything that is not a tcgen05 layer)
    if (time_tc && tc_pending) { tc_record(); tc_pending = false; }
  }
  void tc_begin() {
    if (!time_tc) return;
    static const bool per_launch = getenv("TKB_TC_TIME_PER_LAUNCH") != nullptr;  // the old accounting, for comparison
    if (tc_pending && g_launches == tc_launch_mark && !per_launch) return;       // continues the open run
    tc_flush();
    tc_record();
  }
  void tc_end() {
    if (!time_tc) return;
    tc_pending = true;
    tc_launch_mark = g_launches;
    ++ws->tc_launches;
  }
  void* ptr(int eid, int phys) const {
    const Entry* e = &rt->ent[eid];
    while (e->alias_of >= 0) { e = &rt->ent[e->alias_of]; }
    if (e->is_const) return e->dev_const;
    if (e->is_input && phys == kLogical) return in_ptr[e->input_index];
    if (e->off[phys] < 0) fail("internal: entry %d has no buffer in layout %d", eid, phys);
    return ws->arena + e->off[phys];
  }
};

// ------------------------------------------------------------------------------------------------
// planner helpers
// ------------------------------------------------------------------------------------------------
struct Planner {
  tkb_runtime* rt;
  Graph& g;
  std::vector<Entry>& ent;
  std::ostringstream plan;
  int cur_nid = -1;

  explicit Planner(tkb_runtime* r) : rt(r), g(r->g), ent(r->ent) {}

  long long arena_alloc(long long bytes) {
    long long off = (rt->arena_bytes + 255) & ~255LL;
    rt->arena_bytes = off + std::max<long long>(bytes, 16);
    return off;
  }
  void* dev_upload(const void* host, size_t bytes) {
    void* d = nullptr;
    CUDA_OK(cudaMalloc(&d, std::max<size_t>(bytes, 16)));
    if (bytes) CUDA_OK(cudaMemcpy(d, host, bytes, cudaMemcpyHostToDevice));
    rt->dev_allocs.push_back(d);
    return d;
  }
  void add_step(const std::string& desc, std::function<void(RunCtx&)> fn, int cand = -1) {
    Step s;
    s.fn = std::move(fn);
    s.desc = desc;
    s.nid = cur_nid;
    s.cand = cand;
    s.is_tc = desc.find("tcgen05") != std::string::npos;
    rt->steps.push_back(std::move(s));
    plan << "  [" << cur_nid << "] " << desc << "\n";
  }
  int root(int eid) const {
    while (ent[eid].alias_of >= 0) eid = ent[eid].alias_of;
    return eid;
  }
  bool has(int eid, int phys) const {
    const Entry& e = ent[root(eid)];
    if (e.is_const) return phys == kLogical;
    if (e.is_input && phys == kLogical) return true;
    return e.off[phys] >= 0;
  }
  // constants ------------------------------------------------------------------------------------
  const Entry& cst(int eid, const char* what) const {
    if (eid < 0) fail("missing %s input", what);
    const Entry& e = ent[eid];
    if (!e.is_const) fail("%s must be a constant", what);
    return e;
  }
  std::vector<float> cfloats(int eid, const char* what) const {
    const Entry& e = cst(eid, what);
    if (e.dtype != kF32) fail("%s must be float32", what);
    std::vector<float> v(numel(e.shape));
    memcpy(v.data(), e.host.data(), v.size() * 4);
    return v;
  }
  std::vector<int32_t> cints(int eid, const char* what) const {
    const Entry& e = cst(eid, what);
    const long long n = numel(e.shape);
    std::vector<int32_t> v(n);
    for (long long i = 0; i < n; ++i) {
      switch (e.dtype) {
        case kI32: v[i] = ((const int32_t*)e.host.data())[i]; break;
        case kI8: v[i] = ((const int8_t*)e.host.data())[i]; break;
        case kU8: v[i] = ((const uint8_t*)e.host.data())[i]; break;
        case kI16: v[i] = ((const int16_t*)e.host.data())[i]; break;
        case kI64: v[i] = (int32_t)((const int64_t*)e.host.data())[i]; break;
        case kF32: v[i] = (int32_t)((const float*)e.host.data())[i]; break;
        default: fail("%s: bad dtype", what);
      }
    }
    return v;
  }
  void ensure_dev_const(int eid) {
    Entry& e = ent[eid];
    if (e.is_const && !e.dev_const) e.dev_const = dev_upload(e.host.data(), e.host.size());
  }
  // layout management ----------------------------------------------------------------------------
  // Make sure entry `eid` is materialised in physical layout `phys`; emits a conversion if needed.
  void need(int eid, int phys) {
    const int r = root(eid);
    Entry& e = ent[r];
    if (e.is_const) {
      if (phys != kLogical && !e.nhwc_is_noop()) fail("internal: constant activations must stay logical");
      ensure_dev_const(r);
      return;
    }
    if (has(r, phys)) return;
    const int other = 1 - phys;
    if (!has(r, other)) fail("internal: entry %d used before it is produced", r);
    if (e.shape.size() != 4) fail("internal: layout conversion of a rank-%d tensor", (int)e.shape.size());
    if (e.nhwc_is_noop() && !(e.is_input && other == kLogical)) {
      e.off[phys] = e.off[other];
      return;
    }
    e.off[phys] = arena_alloc(e.bytes());
    const long long N = e.shape[0], C = e.shape[1], H = e.shape[2], W = e.shape[3];
    const int elem = dtype_size(e.dtype);
    std::array<long long, 4> oshape, istr;
    if (phys == kNHWC) {  // out dims (N,H,W,C) reading logical NCHW
      oshape = {N, H, W, C};
      istr = {C * H * W, W, 1, H * W};
    } else {  // out dims (N,C,H,W) reading NHWC
      oshape = {N, C, H, W};
      istr = {H * W * C, 1, W * C, C};
    }
    rt->hbm_bytes += 2 * e.bytes();
    add_step(std::string("layout ") + (phys == kNHWC ? "NCHW->NHWC" : "NHWC->NCHW") + " e" + std::to_string(r),
             [=](RunCtx& c) {
               if (phys == kLogical && elem == 1 && launch_nhwc_to_nchw8(c.ptr(r, other), c.ptr(r, phys), N, H * W, C, c.s)) return;
               // NHWC -> NCHW is [N][HW][C] -> [N][C][HW]; NCHW -> NHWC is [N][C][HW] -> [N][HW][C]
               if (phys == kLogical ? launch_transpose_tile(c.ptr(r, other), c.ptr(r, phys), elem, N, H * W, C, c.s)
                                    : launch_transpose_tile(c.ptr(r, other), c.ptr(r, phys), elem, N, C, H * W, c.s))
                 return;
               launch_permute4(c.ptr(r, other), c.ptr(r, phys), elem, oshape.data(), istr.data(), c.s);
             });
  }
  void produce(int eid, int phys) {
    Entry& e = ent[eid];
    if (e.off[phys] < 0) e.off[phys] = arena_alloc(e.bytes());
  }
  int preferred(int eid) const {  // layout an elementwise consumer should work in
    const int r = root(eid);
    if (ent[r].is_const || ent[r].is_input) return has(r, kNHWC) ? kNHWC : kLogical;
    return ent[r].off[kNHWC] >= 0 ? kNHWC : kLogical;
  }
};

// (multiplier, shift) from double(float32 s_in) / double(float32 s_out): src/relay/qnn/utils.cc:33-57
static void fixed_point_multiplier_shift(double m, int32_t* sig, int32_t* exp) {
  if (m == 0.) { *sig = 0; *exp = 0; return; }
  int e;
  double s = std::frexp(m, &e);
  s = std::round(s * (double)(1ll << 31));
  long long s64 = (long long)s;
  if (s64 == (1ll << 31)) { s64 /= 2; ++e; }
  *sig = (int32_t)s64;
  *exp = e;
}

static void dtype_range(int dt, int32_t* lo, int32_t* hi) {
  switch (dt) {
    case kI8: *lo = -128; *hi = 127; break;
    case kU8: *lo = 0; *hi = 255; break;
    case kI16: *lo = -32768; *hi = 32767; break;
    default: *lo = INT32_MIN; *hi = INT32_MAX; break;
  }
}

static int parse_rounding(const std::string& s) {
  if (s == "UPWARD" || s == "None" || s.empty()) return kUpward;  // default config: requantize_config.h:53-73
  if (s == "TONEAREST") return kToNearest;
  fail("QNN requantize supports two rounding modes - UPWARD and TONEAREST (got %s)", s.c_str());
}
static int parse_compute(const std::string& s) {
  if (s == "int64" || s == "None" || s.empty()) return kInt64;
  if (s == "float32") return kFloat32;
  if (s == "float64") return kFloat64;
  fail("QNN requantize supports three compute_dtype variants (got %s)", s.c_str());
}

// Build a RqParams (RequantizeLower, requantize.cc:195-273).  C = channel count on the quantization axis.
struct RqHost {  // host copy of the fixed-point constants of one requantize (for build_fast_epi)
  std::vector<int32_t> M, S, zp_in;
};

static RqParams make_rq(Planner& P, const std::vector<float>& in_scale, const std::vector<int32_t>& in_zp,
                        float out_scale, int32_t out_zp, int rounding, int compute, int out_dtype, int C,
                        RqHost* host = nullptr) {
  RqParams p;
  memset(&p, 0, sizeof(p));
  if (in_scale.size() != 1 && (int)in_scale.size() != C) fail("requantize: %d scales for %d channels", (int)in_scale.size(), C);
  if (in_zp.size() != 1 && (int)in_zp.size() != C) fail("requantize: %d zero points for %d channels", (int)in_zp.size(), C);
  p.per_channel = in_scale.size() != 1;
  p.rounding = rounding;
  p.compute = compute;
  p.zp_out = out_zp;
  p.no_clamp = out_dtype == kI32;
  dtype_range(out_dtype, &p.qmin, &p.qmax);
  const size_t n = in_scale.size();
  std::vector<int32_t> M(n), S(n);
  std::vector<double> D(n);
  for (size_t i = 0; i < n; ++i) {
    D[i] = (double)in_scale[i] / (double)out_scale;
    fixed_point_multiplier_shift(D[i], &M[i], &S[i]);
  }
  if (!p.per_channel) {
    p.mult0 = M[0]; p.shift0 = S[0]; p.dmul0 = D[0];
    p.skip_mul = in_scale[0] == out_scale;  // IsEqualScalar(input_scale, output_scale), requantize.cc:226
  } else {
    p.mult = (const int32_t*)P.dev_upload(M.data(), n * 4);
    p.shift = (const int32_t*)P.dev_upload(S.data(), n * 4);
    p.dmul = (const double*)P.dev_upload(D.data(), n * 8);
  }
  if (in_zp.size() == 1) p.zp_in0 = in_zp[0];
  else p.zp_in = (const int32_t*)P.dev_upload(in_zp.data(), in_zp.size() * 4);
  if (host) { host->M = M; host->S = S; host->zp_in = in_zp; }
  return p;
}

static void init_epi(EpiParams& e) {
  memset(&e, 0, sizeof(e));
  e.mode = kEpiRaw;
  e.out_dtype = kI32;
}

// qnn.add parameters (add.cc:40-93, op_common.h:196-210) into e.lhs / e.rhs / sum_*
static void make_qadd(Planner& P, EpiParams& e, float sa, int32_t zpa, float sb, int32_t zpb, float so, int32_t zpo,
                      int dtype, RqHost* lhs_host = nullptr, RqHost* rhs_host = nullptr) {
  e.lhs_identity = sa == so && zpa == zpo;
  e.rhs_identity = sb == so && zpb == zpo;
  e.lhs = make_rq(P, {sa}, {zpa}, so, zpo, kUpward, kInt64, kI32, 1, lhs_host);
  e.rhs = make_rq(P, {sb}, {zpb}, so, zpo, kUpward, kInt64, kI32, 1, rhs_host);
  e.sum_zp_out = zpo;
  dtype_range(dtype, &e.sum_qmin, &e.sum_qmax);
  e.sum_clip_lo = e.sum_qmin;
  e.sum_clip_hi = e.sum_qmax;
  e.sum_dtype = dtype;
}

static void clip_bounds(double a_min, double a_max, int dtype, int32_t* lo, int32_t* hi) {
  int32_t qlo, qhi;
  dtype_range(dtype, &qlo, &qhi);
  const double l = std::max(a_min, (double)qlo), h = std::min(a_max, (double)qhi);
  *lo = (int32_t)l;  // relay clip casts the bounds to the tensor dtype
  *hi = (int32_t)h;
}

// ------------------------------------------------------------------------------------------------
// composite epilogue (ParseAttrs equivalent, tachikoma_json_runtime.cc:142-237)
// ------------------------------------------------------------------------------------------------
struct CompositeInfo {
  EpiParams epi;
  int bias_eid = -1;      // runtime bias tensor (non-constant), int32 [channels]
  int sum_eid = -1;       // residual input
  std::vector<int32_t> bias_host;  // constant int32 bias (int mode), empty if none
  bool has_const_bias = false;
  int32_t zp_a = 0;
  std::vector<int32_t> zp_w;
  bool float_mode = false;
  RqHost rq_host, lhs_host, rhs_host;
};

static int input_by_name(const Graph& g, const GNode& n, const char* key) {
  if (!n.has(key)) return -1;
  const int idx = (int)n.i64(key, -1);
  if (idx < 0 || idx >= (int)n.inputs.size()) fail("%s: index %d out of range", key, idx);
  return (int)g.eid(n.inputs[idx]);
}

// `channels`: size of the output channel axis; `axis_is_channel(axis)`: whether a requantize axis
// (logical, may be negative) designates that axis.
static CompositeInfo build_composite(Planner& P, const GNode& n, int nid, int channels, int out_rank, int channel_axis) {
  const Graph& g = P.g;
  CompositeInfo ci;
  init_epi(ci.epi);
  EpiParams& e = ci.epi;
  const int out_eid = (int)g.eid(nid, 0);
  const int out_dtype = P.ent[out_eid].dtype;
  e.out_dtype = out_dtype;
  ci.sum_eid = input_by_name(g, n, "sum_idx");
  const int bias_in = input_by_name(g, n, "bias_idx");
  if (n.has("rq_in_scl_idx")) {
    // ---- integer mode: TVM's exact lowering --------------------------------------------------
    e.mode = kEpiInt;
    ci.zp_a = P.cints(input_by_name(g, n, "src_zp_idx"), "src_zp")[0];
    ci.zp_w = P.cints(input_by_name(g, n, "wgh_zp_idx"), "wgh_zp");
    if (bias_in >= 0) {
      if (numel(P.ent[bias_in].shape) != channels) fail("bias has %lld elements, expected %d", numel(P.ent[bias_in].shape), channels);
      if (P.ent[bias_in].dtype != kI32) fail("bias must be int32");
      if (P.ent[bias_in].is_const) { ci.bias_host = P.cints(bias_in, "bias"); ci.has_const_bias = true; }
      else ci.bias_eid = bias_in;
    }
    auto in_scl = P.cfloats(input_by_name(g, n, "rq_in_scl_idx"), "rq_in_scl");
    auto in_zp = P.cints(input_by_name(g, n, "rq_in_zp_idx"), "rq_in_zp");
    const float out_scl = P.cfloats(input_by_name(g, n, "rq_out_scl_idx"), "rq_out_scl")[0];
    const int32_t out_zp = P.cints(input_by_name(g, n, "rq_out_zp_idx"), "rq_out_zp")[0];
    int axis = (int)n.i64("rq_axis", -1);
    if (axis < 0) axis += out_rank;
    if ((in_scl.size() > 1 || in_zp.size() > 1) && axis != channel_axis)
      fail("Unsupported op: %s with per-channel requantize on axis %d (channel axis is %d)", n.name.c_str(), axis, channel_axis);
    const int rq_dtype = dtype_from_string(n.str("rq_out_dtype", "int32"));
    e.rq = make_rq(P, in_scl, in_zp, out_scl, out_zp, parse_rounding(n.str("rounding", "None")),
                   parse_compute(n.str("compute_dtype", "None")), rq_dtype, channels, &ci.rq_host);
    clip_bounds(n.f64("clip_a_min", -2147483648.0), n.f64("clip_a_max", 2147483647.0), rq_dtype, &e.clip_lo, &e.clip_hi);
    const int cast_dtype = dtype_from_string(n.str("cast_dtype", n.str("rq_out_dtype", "int32")));
    if (ci.sum_eid >= 0) {
      const float sa = P.cfloats(input_by_name(g, n, "sum_lhs_scl_idx"), "sum_lhs_scl")[0];
      const int32_t za = P.cints(input_by_name(g, n, "sum_lhs_zp_idx"), "sum_lhs_zp")[0];
      const float sb = P.cfloats(input_by_name(g, n, "sum_rhs_scl_idx"), "sum_rhs_scl")[0];
      const int32_t zb = P.cints(input_by_name(g, n, "sum_rhs_zp_idx"), "sum_rhs_zp")[0];
      const float so = P.cfloats(input_by_name(g, n, "sum_out_scl_idx"), "sum_out_scl")[0];
      const int32_t zo = P.cints(input_by_name(g, n, "sum_out_zp_idx"), "sum_out_zp")[0];
      if (P.ent[ci.sum_eid].dtype != cast_dtype) fail("qnn.add operands must share a dtype");
      make_qadd(P, e, sa, za, sb, zb, so, zo, cast_dtype, &ci.lhs_host, &ci.rhs_host);
      clip_bounds(n.f64("sum_clip_a_min", -2147483648.0), n.f64("sum_clip_a_max", 2147483647.0), cast_dtype,
                  &e.sum_clip_lo, &e.sum_clip_hi);
      e.sum = (const void*)1;  // patched with the real pointer at run time
    } else {
      e.sum_dtype = cast_dtype;
    }
    if (ci.sum_eid < 0 && cast_dtype != out_dtype) fail("composite output dtype mismatch");
  } else if (n.has("o_scl_idx")) {
    // ---- oneDNN-style float post-ops (the reference's literal path B) -------------------------
    e.mode = kEpiFloat;
    ci.float_mode = true;
    ci.zp_w = {0};
    if (bias_in >= 0) {
      auto bf = P.cfloats(bias_in, "bias");
      std::vector<float> full(channels);
      if ((int)bf.size() == channels) full = bf;
      else if (bf.size() == 1) std::fill(full.begin(), full.end(), bf[0]);
      else fail("float bias has %d elements, expected %d", (int)bf.size(), channels);
      e.bias = (const int32_t*)P.dev_upload(full.data(), full.size() * 4);
    }
    auto oscl = P.cfloats(input_by_name(g, n, "o_scl_idx"), "o_scl");
    if (oscl.size() != 1 && (int)oscl.size() != channels) fail("o_scl has %d elements", (int)oscl.size());
    e.o_scl = (const float*)P.dev_upload(oscl.data(), oscl.size() * 4);
    e.n_oscl = (int)oscl.size();
    e.act_scl = 1.f;
    if (n.has("activation")) {
      auto& act = n.list("activation");
      if (act[0] != "clip") fail("Unsupported op: activation %s", act[0].c_str());
      const int ia = std::stoi(act[1]), il = std::stoi(act[2]), ih = std::stoi(act[3]);
      e.act_scl = P.cfloats((int)g.eid(n.inputs[ia]), "act_scale")[0];
      const float lo = P.cfloats((int)g.eid(n.inputs[il]), "alpha")[0], hi = P.cfloats((int)g.eid(n.inputs[ih]), "beta")[0];
      if (lo != 0.f || hi != 255.f) fail("Unsupported op: clip(%g, %g) activation", lo, hi);
    }
    const int ss = input_by_name(g, n, "sum_scl_idx");
    e.sum_scl = ss >= 0 ? P.cfloats(ss, "sum_scl")[0] : 0.f;
    const int dz = input_by_name(g, n, "dst_zp_idx");
    e.dst_zp = dz >= 0 ? P.cfloats(dz, "dst_zp")[0] : 0.f;
    dtype_range(out_dtype, &e.f_qmin, &e.f_qmax);
    if (ci.sum_eid >= 0) { e.sum = (const void*)1; e.sum_dtype = P.ent[ci.sum_eid].dtype; }
  } else {
    fail("Unsupported op: %s without quantization attributes", n.name.c_str());
  }
  return ci;
}

static std::vector<int32_t> add_vec(const std::vector<int32_t>& bias, const std::vector<int32_t>& corr, int n) {
  std::vector<int32_t> out(n);
  for (int i = 0; i < n; ++i)
    out[i] = (int32_t)((uint32_t)(bias.empty() ? 0 : bias[i]) + (uint32_t)corr[i]);
  return out;
}

static bool all_zero(const std::vector<int32_t>& v) {
  for (auto x : v) if (x != 0) return false;
  return true;
}

// ------------------------------------------------------------------------------------------------
// host-side folding for the tensor-core kernels
// ------------------------------------------------------------------------------------------------
static bool fits_i63(__int128 v) { return v < ((__int128)1 << 62) && v > -((__int128)1 << 62); }

// One side of the residual qnn.add (RequantizeOrUpcast, op_common.h:196-210) as q' = ((q << ls) * M + C) >> T.
// `p`/`h`: the requantize of this side towards the add's output qparams.  Returns false when the side
// cannot be expressed (absurd scale ratios); the caller then keeps the generic epilogue.
static bool fold_side(const RqParams& p, const RqHost& h, bool identity, FastSide* f) {
  memset(f, 0, sizeof(*f));
  const long long zin = h.zp_in[0], zo = p.zp_out;
  if (identity) { f->M = 1; return true; }
  if (p.skip_mul) { f->M = 1; f->C = zo - zin; return true; }
  const int32_t M = h.M[0], S = h.S[0];
  if (M <= 0) return false;
  if (M == (1 << 30)) {  // power-of-two shortcut of the per-tensor UPWARD path, intrin_rule.cc:221-239
    const int e = S - 1;
    if (e > 16 || e < -30) return false;
    f->M = 1;
    if (e > 0) { f->ls = e; f->C = -zin * (1ll << e) + zo; }
    else if (e == 0) { f->C = zo - zin; }
    else { f->s1 = -e; f->C = -zin + (1ll << (-e - 1)) + zo * (1ll << -e); }
    return true;
  }
  const int ls = S > 0 ? S : 0, rs = S > 0 ? 0 : -S, T = rs + 31;
  if (ls > 8 || T > 62) return false;
  const __int128 C = -(__int128)zin * ((__int128)M << ls) + ((__int128)1 << (T - 1)) + (__int128)zo * ((__int128)1 << T);
  if (!fits_i63(C)) return false;
  f->M = M; f->ls = ls; f->s1 = T; f->C = (long long)C;
  return true;
}

// Rewrite q' = ((q << ls) * M + C) >> T  (T in f->s1) as  hi32((q << ls') * M' + C') >> s  so that the kernel
// needs one IMAD.WIDE and one 32-bit shift: scale by 2^(32-T) when T < 32, split between the left shift
// of q (|q| <= 255, so ls' <= 22 keeps q << ls' inside int32) and the multiplier.
static bool normalize_side(FastSide* f) {
  int T = f->s1;
  __int128 C = f->C;
  long long M = f->M;
  int ls = f->ls;
  if (T < 32) {
    int k = 32 - T;
    const int k1 = std::min(k, 22 - ls);
    if (k1 < 0) return false;
    ls += k1;
    k -= k1;
    M <<= k;
    if (M > INT32_MAX) return false;
    C <<= (32 - T);
    T = 32;
  }
  if (ls > 22 || !fits_i63(C) || !fits_i63(((__int128)256 << ls) * M + (C < 0 ? -C : C))) return false;
  f->M = (int32_t)M; f->ls = ls; f->s1 = T - 32; f->C = (long long)C;
  return true;
}

// Fold bias, zero points, rounding constant and the two clamps of the int-mode composite into FastEpi
// (see epilogue.cuh).  bias2[c]: constant int32 bias plus the uniform zero-point correction; wabs[c]:
// sum_k |w[c][k]| (bounds every accumulator the kernel or the reference can form for channel c).
static bool build_fast_epi(Planner& P, const CompositeInfo& ci, const std::vector<int32_t>& bias2,
                           const std::vector<long long>& wabs, int channels, int out_dtype, FastEpi* out) {
  memset(out, 0, sizeof(*out));
  const EpiParams& e = ci.epi;
  if (e.mode != kEpiInt || ci.bias_eid >= 0) return false;
  const RqParams& rq = e.rq;
  if (rq.compute != kInt64 || rq.rounding != kUpward || rq.skip_mul) return false;
  const RqHost& h = ci.rq_host;
  if (h.M.empty()) return false;
  const bool has_sum = ci.sum_eid >= 0;
  const int mid_dtype = has_sum ? e.sum_dtype : out_dtype;  // dtype of the cast that follows the clip
  int32_t dlo, dhi;
  dtype_range(mid_dtype, &dlo, &dhi);
  const int32_t lo = std::max(rq.no_clamp ? INT32_MIN : rq.qmin, e.clip_lo);
  const int32_t hi = std::min(rq.no_clamp ? INT32_MAX : rq.qmax, e.clip_hi);
  if (lo > hi || lo < dlo || hi > dhi) return false;  // the cast must be the identity on [lo, hi]
  std::vector<int32_t> tab((size_t)channels * 4);
  for (int c = 0; c < channels; ++c) {
    const int32_t M = h.M[h.M.size() == 1 ? 0 : c], S = h.S[h.S.size() == 1 ? 0 : c];
    const long long zin = h.zp_in[h.zp_in.size() == 1 ? 0 : c];
    if (M <= 0 || S > -1) return false;  // T = 31 - S must be >= 32 (multiplier < 0.5)
    const int T = 31 - S;
    if (T > 62) return false;
    const long long b = bias2.empty() ? 0 : bias2[c];
    // every |acc|, |acc + delta|, |acc + bias|, |acc + bias - zp_in| stays below 2^31: the reference's int32
    // ring arithmetic equals plain integer arithmetic
    const __int128 amax = (__int128)2 * 255 * wabs[c] + (b < 0 ? -b : b) + (zin < 0 ? -zin : zin) + 1;
    if (amax >= ((__int128)1 << 31)) return false;
    if (h.M.size() == 1 && M == (1 << 30) && (-S > 29 || amax >= ((__int128)1 << 30))) return false;  // shortcut branch
    const __int128 C = (__int128)(b - zin) * M + ((__int128)1 << (T - 1)) + (__int128)rq.zp_out * ((__int128)1 << T);
    if (!fits_i63(C) || !fits_i63(amax * M + (C < 0 ? -C : C))) return false;
    const long long C64 = (long long)C;
    tab[4 * c + 0] = M;
    tab[4 * c + 1] = T - 32;
    tab[4 * c + 2] = (int32_t)(uint32_t)(C64 & 0xFFFFFFFFll);
    tab[4 * c + 3] = (int32_t)(C64 >> 32);
  }
  out->lo = lo;
  out->hi = hi;
  if (has_sum) {
    if (e.sum_dtype != kU8 && e.sum_dtype != kI8) return false;
    if (!fold_side(e.lhs, ci.lhs_host, e.lhs_identity != 0, &out->lhs)) return false;
    if (!fold_side(e.rhs, ci.rhs_host, e.rhs_identity != 0, &out->rhs)) return false;
    const __int128 Cl = (__int128)out->lhs.C - (__int128)e.sum_zp_out * ((__int128)1 << out->lhs.s1);
    if (!fits_i63(Cl)) return false;
    out->lhs.C = (long long)Cl;
    out->has_sum = 1;
    if (e.sum_dtype == kI8) {  // the kernel reads the residual bytes as u' = u ^ 0x80, i.e. b = u' - 128
      out->sum_xor = 0x80808080u;
      const __int128 Cr = (__int128)out->rhs.C - ((__int128)128 << out->rhs.ls) * out->rhs.M;
      if (!fits_i63(Cr)) return false;
      out->rhs.C = (long long)Cr;
    }
    if (!normalize_side(&out->lhs) || !normalize_side(&out->rhs)) return false;
    out->sum_lo = std::max(e.sum_qmin, e.sum_clip_lo);
    out->sum_hi = std::min(e.sum_qmax, e.sum_clip_hi);
    int32_t olo, ohi;
    dtype_range(out_dtype, &olo, &ohi);
    if (out->sum_lo > out->sum_hi || out->sum_lo < olo || out->sum_hi > ohi) return false;
  }
  auto sat_code = [](int dtype, int32_t l, int32_t h) {
    return dtype == kU8 && l == 0 && h == 255 ? 1 : (dtype == kI8 && l == -128 && h == 127 ? 2 : 0);
  };
  out->sat = has_sum ? 0 : sat_code(out_dtype, lo, hi);
  out->sum_sat = has_sum ? sat_code(out_dtype, out->sum_lo, out->sum_hi) : 0;
  out->tabA = (const int4*)P.dev_upload(tab.data(), tab.size() * 4);
  P.rt->host_tabs.emplace_back(tab.begin(), tab.end());
  out->tab_host = (const int4*)P.rt->host_tabs.back().data();
  out->enabled = 1;
  return true;
}

// sum_k |w[o][k]| of an [O][K] int8/uint8 matrix
static std::vector<long long> abs_row_sums(const uint8_t* w, int O, int K, int w_dtype) {
  std::vector<long long> out(O, 0);
  for (int o = 0; o < O; ++o) {
    long long s = 0;
    for (int k = 0; k < K; ++k) {
      const int v = w_dtype == kU8 ? (int)w[(size_t)o * K + k] : (int)(int8_t)w[(size_t)o * K + k];
      s += v < 0 ? -v : v;
    }
    out[o] = s;
  }
  return out;
}

// Implicit-GEMM convolution reads 0 (TMA out-of-bounds fill) where the reference pads with the input zero
// point (convolution.cc:227-253), so  acc_ref = raw - zp_a * sum_{valid taps} W.  The uniform part
// (-zp_a * sum_{all taps} W) goes into the bias; an output pixel whose window overlaps the padding gets
// +zp_a * sum_{taps in the padding} W back.  Pixels are classified per axis by their set of valid taps;
// class 0 is "all taps valid".  Returns false if the geometry has too many classes for uint8 tables.
struct BorderPlan {
  std::vector<uint8_t> cls_h, cls_w;
  std::vector<int32_t> delta;  // [nh*nw][O]
  int nh = 1, nw = 1;
};
static bool axis_classes(int in, int out, int k, int stride, int dil, int pad, std::vector<uint8_t>* cls,
                         std::vector<uint64_t>* masks) {
  if (k > 63) return false;
  masks->assign(1, (1ull << k) - 1);
  cls->resize(out);
  for (int p = 0; p < out; ++p) {
    uint64_t m = 0;
    for (int r = 0; r < k; ++r) {
      const int i = p * stride + r * dil - pad;
      if (i >= 0 && i < in) m |= 1ull << r;
    }
    size_t id = 0;
    for (; id < masks->size(); ++id) if ((*masks)[id] == m) break;
    if (id == masks->size()) masks->push_back(m);
    if (id > 255) return false;
    (*cls)[p] = (uint8_t)id;
  }
  return true;
}
static bool border_plan(const ConvGeom& g, const std::vector<uint8_t>& w_ohwi, int Cg, BorderPlan* bp) {
  std::vector<uint64_t> mh, mw;
  if (!axis_classes(g.H, g.OH, g.KH, g.sh, g.dh, g.pt, &bp->cls_h, &mh)) return false;
  if (!axis_classes(g.W, g.OW, g.KW, g.sw, g.dw, g.pl, &bp->cls_w, &mw)) return false;
  bp->nh = (int)mh.size();
  bp->nw = (int)mw.size();
  if (bp->nh * bp->nw == 1 || g.zp_a == 0) { bp->delta.clear(); return true; }
  std::vector<long long> tap((size_t)g.O * g.KH * g.KW, 0);  // sum over channels of each tap
  for (int o = 0; o < g.O; ++o)
    for (int t = 0; t < g.KH * g.KW; ++t) {
      long long s = 0;
      const uint8_t* src = &w_ohwi[((size_t)o * g.KH * g.KW + t) * Cg];
      for (int c = 0; c < Cg; ++c) s += g.w_dtype == kU8 ? (int)src[c] : (int)(int8_t)src[c];
      tap[(size_t)o * g.KH * g.KW + t] = s;
    }
  bp->delta.assign((size_t)bp->nh * bp->nw * g.O, 0);
  for (int ch = 0; ch < bp->nh; ++ch)
    for (int cw = 0; cw < bp->nw; ++cw)
      for (int o = 0; o < g.O; ++o) {
        long long s = 0;
        for (int r = 0; r < g.KH; ++r)
          for (int x = 0; x < g.KW; ++x)
            if (!((mh[ch] >> r) & 1) || !((mw[cw] >> x) & 1)) s += tap[((size_t)o * g.KH + r) * g.KW + x];
        bp->delta[((size_t)ch * bp->nw + cw) * g.O + o] = (int32_t)((uint32_t)g.zp_a * (uint32_t)(int32_t)s);
      }
  return true;
}

// ------------------------------------------------------------------------------------------------
// convolution (Convolution builder, tachikoma_json_runtime.cc:292-388; arithmetic convolution.cc:708-811)
// ------------------------------------------------------------------------------------------------
static std::vector<uint8_t> repack_ohwi(const Entry& w, const std::string& layout, int O, int Cg, int KH, int KW) {
  std::vector<uint8_t> out((size_t)O * KH * KW * Cg);
  const uint8_t* src = w.host.data();
  auto at = [&](int o, int i, int h, int x) -> uint8_t {
    if (layout == "OIHW") return src[(((size_t)o * Cg + i) * KH + h) * KW + x];
    if (layout == "HWIO") return src[(((size_t)h * KW + x) * Cg + i) * O + o];
    if (layout == "OHWI") return src[(((size_t)o * KH + h) * KW + x) * Cg + i];
    // HWOI depthwise: [H][W][C][mult], output channel = c*mult + m, one input channel per group
    return src[((size_t)h * KW + x) * O + o];
  };
  for (int o = 0; o < O; ++o)
    for (int h = 0; h < KH; ++h)
      for (int x = 0; x < KW; ++x)
        for (int i = 0; i < Cg; ++i) out[(((size_t)o * KH + h) * KW + x) * Cg + i] = at(o, i, h, x);
  return out;
}

// Another tensor-core composite (convolution / dense) reads the same activation as node `nid`: the two are independent
// and can share one chained launch (a residual block's 3x3 convolution and its 1x1 projection; Q / K / V projections).
static bool has_sibling_composite(const Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  if (n.inputs.empty()) return false;
  const uint32_t x = P.g.eid(n.inputs[0]);
  for (size_t j = 0; j < P.g.nodes.size(); ++j) {
    const GNode& m = P.g.nodes[j];
    if ((int)j == nid || m.op_type != "kernel" || m.inputs.empty()) continue;
    if (m.name != "tachikoma.qnn.conv2d" && m.name != "tachikoma.qnn.dense") continue;
    if (P.g.eid(m.inputs[0]) == x) return true;
  }
  return false;
}

static void plan_conv(Planner& P, int nid, bool composite) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), w_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  const Entry& Wt = P.cst(w_eid, "convolution weight");
  Entry& Y = P.ent[y_eid];
  const std::string dl = n.str("data_layout", "NCHW"), kl = n.str("kernel_layout", "OIHW");
  std::string ol = n.str("out_layout", "");
  if (ol.empty()) ol = dl;
  if ((dl != "NCHW" && dl != "NHWC") || ol != dl) fail("Unsupported op: %s with data_layout %s / out_layout %s", n.name.c_str(), dl.c_str(), ol.c_str());
  if (kl != "OIHW" && kl != "HWIO" && kl != "OHWI" && kl != "HWOI") fail("Unsupported op: %s with kernel_layout %s", n.name.c_str(), kl.c_str());
  if (X.shape.size() != 4 || Wt.shape.size() != 4) fail("conv2d expects 4-D data and kernel");
  ConvGeom g;
  memset(&g, 0, sizeof(g));
  auto dim = [&](const std::vector<int64_t>& shp, const std::string& lay, char c) { return (int)shp[lay.find(c)]; };
  g.N = dim(X.shape, dl, 'N'); g.C = dim(X.shape, dl, 'C'); g.H = dim(X.shape, dl, 'H'); g.W = dim(X.shape, dl, 'W');
  g.groups = (int)n.i64("groups", 1);
  g.KH = dim(Wt.shape, kl, 'H'); g.KW = dim(Wt.shape, kl, 'W');
  int Cg;
  if (kl == "HWOI") {  // depthwise, channels-last convention
    g.O = (int)(Wt.shape[2] * Wt.shape[3]);
    Cg = 1;
    if (g.groups != g.C || Wt.shape[2] != g.C) fail("Unsupported op: HWOI kernel layout is only valid for depthwise convolution");
  } else {
    g.O = dim(Wt.shape, kl, 'O');
    Cg = dim(Wt.shape, kl, 'I');
  }
  if (Cg * g.groups != g.C || g.O % g.groups) fail("conv2d: channels %d, groups %d, kernel I %d do not agree", g.C, g.groups, Cg);
  auto st = n.ints("strides"), pd = n.ints("padding"), di = n.ints("dilation");
  if (st.size() != 2 || di.size() != 2) fail("conv2d: strides / dilation must have 2 elements");
  g.sh = (int)st[0]; g.sw = (int)st[1]; g.dh = (int)di[0]; g.dw = (int)di[1];
  if (pd.size() == 4) { g.pt = (int)pd[0]; g.pl = (int)pd[1]; g.pb = (int)pd[2]; g.pr = (int)pd[3]; }
  else if (pd.size() == 2) { g.pt = g.pb = (int)pd[0]; g.pl = g.pr = (int)pd[1]; }
  else if (pd.size() == 1) { g.pt = g.pb = g.pl = g.pr = (int)pd[0]; }
  else fail("conv2d: bad padding");
  g.OH = (g.H + g.pt + g.pb - (g.KH - 1) * g.dh - 1) / g.sh + 1;
  g.OW = (g.W + g.pl + g.pr - (g.KW - 1) * g.dw - 1) / g.sw + 1;
  if (dim(Y.shape, ol, 'N') != g.N || dim(Y.shape, ol, 'C') != g.O || dim(Y.shape, ol, 'H') != g.OH || dim(Y.shape, ol, 'W') != g.OW)
    fail("conv2d: output shape in the graph does not match the attributes");
  g.a_dtype = X.dtype; g.w_dtype = Wt.dtype;
  if ((g.a_dtype != kI8 && g.a_dtype != kU8) || (g.w_dtype != kI8 && g.w_dtype != kU8))
    fail("Unsupported op: %s with data/kernel dtypes other than int8/uint8", n.name.c_str());

  CompositeInfo ci;
  const int ch_axis = (int)ol.find('C');
  if (composite) {
    ci = build_composite(P, n, nid, g.O, 4, ch_axis);
  } else {
    init_epi(ci.epi);
    if (Y.dtype != kI32) fail("qnn.conv2d out_dtype must be int32");
    ci.zp_a = P.cints((int)P.g.eid(n.inputs[2]), "input_zero_point")[0];
    ci.zp_w = P.cints((int)P.g.eid(n.inputs[3]), "kernel_zero_point");
  }
  if (ci.zp_w.size() != 1 && (int)ci.zp_w.size() != g.O) fail("kernel zero point must be a scalar or have one value per output channel");
  g.zp_a = ci.float_mode ? 0 : ci.zp_a;
  g.zp_w0 = ci.zp_w[0];
  if (ci.zp_w.size() > 1) g.zp_w = (const int32_t*)P.dev_upload(ci.zp_w.data(), ci.zp_w.size() * 4);

  const int phys = dl == "NCHW" ? kNHWC : kLogical;
  std::vector<uint8_t> w_ohwi = repack_ohwi(Wt, kl, g.O, Cg, g.KH, g.KW);
  const int K = g.KH * g.KW * Cg;
  const long long M = (long long)g.N * g.OH * g.OW;
  const bool direct = g.KH == 1 && g.KW == 1 && g.sh == 1 && g.sw == 1 && !(g.pt | g.pl | g.pb | g.pr) && (g.C % 16) == 0;
  const bool tc_ok = P.rt->conv_impl != "simt" && g.groups == 1 && all_zero(ci.zp_w) &&
                     (g.zp_a == 0 || ci.bias_eid < 0) && M >= 1;
  const bool force_tc = P.rt->conv_impl == "tc" || P.rt->conv_impl == "tc_explicit" || P.rt->conv_impl == "tc_im2col" ||
                        P.rt->conv_impl == "tc_nochain" || P.rt->conv_impl == "tc_pair" || P.rt->conv_impl.rfind("tc_chain", 0) == 0;
  const bool use_tc = tc_ok && (force_tc || M * g.O >= 4096);
  // "Kernel view" of the convolution.  A few-channel convolution whose input still lives in logical NCHW
  // form (the network stem) is rewritten as a 1 x KW convolution over the vertically unfolded tensor
  // X3[n][p][w][r*C + c] (32 bytes per pixel, launch_unfold_rows_nchw): same integers, but addressable by the
  // TMA im2col path and 2.4x fewer bytes than an explicit im2col matrix.
  std::vector<uint8_t> wk;
  int Cgk = Cg, Kk = K;
  const bool unfold = use_tc && !direct && P.rt->conv_impl != "tc_explicit" && g.groups == 1 && phys == kNHWC &&
                      P.has(x_eid, kLogical) && !P.has(x_eid, kNHWC) && !X.nhwc_is_noop() && (g.C % 32) != 0 &&
                      g.C * g.KH <= 32 && g.KW <= 63 && g.dw == 1;
  if (unfold) {
    Cgk = 32; Kk = g.KW * 32;
    wk.assign((size_t)g.O * Kk, 0);
    for (int o = 0; o < g.O; ++o)
      for (int r = 0; r < g.KH; ++r)
        for (int x = 0; x < g.KW; ++x)
          for (int c = 0; c < Cg; ++c)
            wk[((size_t)o * g.KW + x) * 32 + r * Cg + c] = w_ohwi[(((size_t)o * g.KH + r) * g.KW + x) * Cg + c];
  } else {
    wk = w_ohwi;
  }
  const bool want_implicit = use_tc && !direct && !unfold && P.rt->conv_impl != "tc_explicit" &&
                             tc_implicit_conv_supported(g) && g.KH <= 63 && g.KW <= 63;
  // Explicit im2col of a tensor that still lives in its logical NCHW form: gather from NCHW directly instead
  // of permuting to channels-last first.
  const bool from_nchw = use_tc && !direct && !want_implicit && !unfold && phys == kNHWC && P.has(x_eid, kLogical) &&
                         !P.has(x_eid, kNHWC) && !X.nhwc_is_noop() && g.KH <= 255 && g.KW <= 255;
  P.need(x_eid, (from_nchw || unfold) ? kLogical : phys);
  if (ci.sum_eid >= 0) P.need(ci.sum_eid, phys);
  if (ci.bias_eid >= 0) P.need(ci.bias_eid, kLogical);
  P.produce(y_eid, phys);
  P.rt->macs += M * g.O * K;
  P.rt->hbm_bytes += X.bytes() + Y.bytes() + (long long)w_ohwi.size() + (ci.sum_eid >= 0 ? Y.bytes() : 0);

  EpiParams epi = ci.epi;
  const int bias_eid = ci.bias_eid, sum_eid = ci.sum_eid;
  char desc[256];
  if (use_tc) {
    // A-operand source: (1) the activation tensor itself for 1x1/stride-1/unpadded convolutions, (2) the
    // activation tensor through an im2col-mode TMA map (implicit GEMM), (3) an explicit im2col matrix for
    // channel counts TMA cannot address (C % 32 != 0, e.g. the 3-channel stem).
    BorderPlan bp;
    const bool implicit = want_implicit && border_plan(g, wk, Cgk, &bp);
    const int K = Kk;  // from here on: the kernel view
    const int Kp = (direct || implicit || unfold) ? K : (K + 15) / 16 * 16;
    std::vector<uint8_t> wp((size_t)g.O * Kp, 0);
    std::vector<int32_t> corr(g.O, 0);
    for (int o = 0; o < g.O; ++o) {
      if (from_nchw) {  // K ordered (c, r, s) to match launch_im2col_nchw
        for (int c = 0; c < Cg; ++c)
          for (int t = 0; t < g.KH * g.KW; ++t) wp[(size_t)o * Kp + (size_t)c * g.KH * g.KW + t] = wk[((size_t)o * g.KH * g.KW + t) * Cg + c];
      } else {
        memcpy(&wp[(size_t)o * Kp], &wk[(size_t)o * K], K);
      }
      long long sum = 0;
      for (int k = 0; k < K; ++k) sum += g.w_dtype == kU8 ? (int)wk[(size_t)o * K + k] : (int)(int8_t)wk[(size_t)o * K + k];
      corr[o] = (int32_t)(0u - (uint32_t)g.zp_a * (uint32_t)(int32_t)sum);  // -zp_a * sum_k W[o,k]  (int32 ring)
    }
    std::vector<int32_t> bias2;
    if (!ci.float_mode) {
      if (ci.has_const_bias || g.zp_a != 0) {
        bias2 = add_vec(ci.bias_host, corr, g.O);
        epi.bias = (const int32_t*)P.dev_upload(bias2.data(), bias2.size() * 4);
      }
    }
    TcArgs ta;
    memset(&ta, 0, sizeof(ta));
    ta.B = 1; ta.M = (int)M; ta.N = g.O; ta.K = Kp;
    ta.a_dtype = g.a_dtype; ta.w_dtype = g.w_dtype;
    ta.implicit = unfold ? 2 : (implicit ? 1 : 0);
    ta.g = g;
    if (composite && !ci.float_mode)
      build_fast_epi(P, ci, bias2, abs_row_sums(wk.data(), g.O, K, g.w_dtype), g.O, Y.dtype, &ta.f);
    // 1x1 convolution with 8 .. 32 input channels: one or two MMAs per 128-pixel tile make the tensor-core kernel
    // barrier-bound (profiles/r02_notes.md sections 13, 20); the default plan runs these on the CUDA-core pointwise
    // kernel.  conv_impl "tc" keeps the tcgen05 form (A/B, tests).
    // (measured, MobileNetV2 batch 128: 32 -> 16 at 112x112 48.4 -> 26.0 us, 16 -> 96 91 -> 79.5 us, 24 -> 144 at 56x56
    // 59 -> 42.5 us incl. the K-padding pass it no longer needs; at 28x28 the tensor-core kernel wins, 17.4 vs 21.3 us, hence
    // the pixel-count threshold; TKB_PW_MIN_M overrides it, read at plan time so that tests can exercise small shapes)
    static const bool pw_off = getenv("TKB_NO_PW") != nullptr;
    const char* pw_min_env = getenv("TKB_PW_MIN_M");
    const long long pw_min_m = pw_min_env ? atoll(pw_min_env) : (1LL << 18);
    if (P.rt->conv_impl == "auto" && !pw_off && M >= pw_min_m && g.KH == 1 && g.KW == 1 && g.sh == 1 && g.sw == 1 && !(g.pt | g.pl | g.pb | g.pr) &&
        !unfold && !from_nchw && pwconv_supported(Cg, g.O) && ta.f.enabled && !ta.f.has_sum && bias_eid < 0 && sum_eid < 0 &&
        dtype_size(Y.dtype) == 1 && (int)wp.size() == g.O * ((Cg + 15) / 16 * 16)) {
      void* wdev = P.dev_upload(wp.data(), wp.size());
      const FastEpi fe = ta.f;
      const int Kc = Cg, O = g.O, adt = g.a_dtype, wdt = g.w_dtype;
      snprintf(desc, sizeof(desc), "%s -> CUDA-core pointwise conv (dp4a, 4 pixels x %d channels per thread) M=%lld N=%d K=%d epi=fast",
               n.name.c_str(), O % 16 == 0 ? 16 : 8, M, O, Kc);
      P.add_step(desc, [=](RunCtx& c) {
        if (launch_pwconv(c.ptr(x_eid, phys), wdev, c.ptr(y_eid, phys), M, O, Kc, adt, wdt, fe, c.s) != 0)
          fail("pointwise convolution launch failed: %s", cudaGetErrorString(cudaGetLastError()));
      });
      return;
    }
    // Chained launch (kernels_tc_chain.cu): layers with >= 96 output channels in the implicit-GEMM / plain-GEMM form
    // become candidates; form_chains() merges consecutive candidates into one persistent launch.
    const int chain_bn = 128;
    bool chain = P.rt->chain_enabled && composite && !ci.float_mode && ta.f.enabled && !unfold && !from_nchw &&
                 (P.rt->chain_policy != 1 || has_sibling_composite(P, nid)) &&
                 (direct || implicit) && ci.bias_eid < 0 && g.O >= 96 && !P.ent[P.root(x_eid)].is_input &&
                 (ci.sum_eid < 0 || !P.ent[P.root(ci.sum_eid)].is_input);
    if (chain) {
      TcArgs probe = ta;
      probe.e = epi;
      chain = tc_chain_supported(probe, chain_bn);
    }
    // cta_group::2 pair kernel (kernels_tc_pair.cu): opt-in (conv_impl = tc_pair) for GEMM-shaped layers with >= 128
    // output channels; takes precedence over the halo kernel where both apply
    bool pair = false;
    if (!chain && P.rt->conv_impl == "tc_pair" && composite && !ci.float_mode && ta.f.enabled && !unfold && !from_nchw &&
        (direct || implicit) && ci.bias_eid < 0 && g.O % 128 == 0) {
      TcArgs probe = ta;
      probe.e = epi;
      probe.w_packed = 0;
      pair = tc_pair_supported(probe);
    }
    bool halo = !chain && !pair && implicit && ta.f.enabled && P.rt->conv_impl != "tc_im2col" && tc_halo_supported(g, g.O);
    if (halo) ta.implicit = 3;
    // Stride-2 stem over the unfolded tensor: two horizontally adjacent 32-byte pixels of X3 form one 64-byte
    // "pair pixel", output column q reads pair pixels q .. q + ceil(KW/2) - 1: a stride-1 1 x ceil(KW/2)
    // convolution with 64 channels, which the halo-reuse kernel runs with one window load per output row
    // (each X3 byte enters shared memory once instead of KW/2 times) and the weights resident.
    if (unfold && g.sw == 2 && (g.W + g.pl + g.pr) % 2 == 0 && ta.f.enabled && P.rt->conv_impl != "tc_im2col") {
      ConvGeom g2 = g;
      g2.C = 64; g2.H = g.OH; g2.W = (g.W + g.pl + g.pr) / 2; g2.KH = 1; g2.KW = (g.KW + 1) / 2;
      g2.sh = g2.sw = g2.dh = g2.dw = 1;
      g2.pt = g2.pb = g2.pl = g2.pr = 0;
      if (g2.W - g2.KW + 1 == g.OW && tc_halo_supported(g2, g.O)) {
        const int K2 = g2.KW * 64;
        std::vector<uint8_t> w2((size_t)g.O * K2, 0);
        for (int o = 0; o < g.O; ++o) memcpy(&w2[(size_t)o * K2], &wp[(size_t)o * Kp], Kp);
        wp.swap(w2);
        ta.g = g2;
        ta.K = K2;
        ta.implicit = 3;
        halo = true;
      }
    }
    const int Kw = ta.K;  // row length of wp
    if (pair) {  // the pair kernel reads the plain K-major weight matrix through a tensor map (half of the rows per CTA)
      ta.w = P.dev_upload(wp.data(), wp.size());
      ta.w_packed = 0;
    } else {  // constant weights: pre-tile into the swizzled shared-memory image (one bulk copy per tile at run time)
      if (chain) { ta.pack_bn = chain_bn; ta.pack_bk = tc_chain_bk(ta); }
      else tc_tile_config(ta, &ta.pack_bn, &ta.pack_bk);
      std::vector<uint8_t> packed(tc_pack_weights(wp.data(), g.O, Kw, ta.pack_bn, ta.pack_bk, nullptr));
      tc_pack_weights(wp.data(), g.O, Kw, ta.pack_bn, ta.pack_bk, packed.data());
      ta.w = P.dev_upload(packed.data(), packed.size());
      ta.w_packed = 1;
    }
    if (implicit && !bp.delta.empty()) {
      ta.delta = (const int32_t*)P.dev_upload(bp.delta.data(), bp.delta.size() * 4);
      ta.cls_h = (const uint8_t*)P.dev_upload(bp.cls_h.data(), bp.cls_h.size());
      ta.cls_w = (const uint8_t*)P.dev_upload(bp.cls_w.data(), bp.cls_w.size());
      ta.ncls_w = bp.nw;
    }
    long long scratch_off = -1;
    if (unfold) scratch_off = P.arena_alloc((long long)g.N * g.OH * (g.W + g.pl + g.pr) * 32 + 256);
    else if (!direct && !implicit) scratch_off = P.arena_alloc(M * Kp);
    snprintf(desc, sizeof(desc), "%s -> tcgen05 i8 %s M=%lld N=%d K=%d (%dx%d s%d p%d) epi=%s", n.name.c_str(),
             pair ? (direct ? "cta_group::2 pair GEMM" : "cta_group::2 pair implicit-GEMM conv (TMA im2col)") : direct ? "GEMM" : (unfold ? (halo ? "halo-reuse conv over pixel pairs of the vertically unfolded NCHW input" : "conv as windowed-row GEMM over vertically unfolded NCHW input") : halo ? "halo-reuse conv (TMA patch, resident weights)" : implicit ? "implicit-GEMM conv (TMA im2col)" : (from_nchw ? "GEMM + explicit im2col from NCHW" : "GEMM + explicit im2col")), M, g.O, Kp, g.KH,
             g.KW, g.sh, g.pt, ta.f.enabled ? "fast" : (epi.mode == kEpiInt ? "int" : (epi.mode == kEpiFloat ? "float" : "raw")));
    P.rt->tc_macs += M * g.O * (long long)(g.KH * g.KW * Cg);
    if (chain) {
      ChainCand cc;
      cc.ta = ta;
      cc.ta.e = epi;
      cc.x_eid = x_eid; cc.y_eid = y_eid; cc.sum_eid = sum_eid; cc.phys = phys;
      cc.rows_per_img = g.OH * g.OW;
      cc.bn = chain_bn;
      cc.name = n.name;
      P.rt->chain_cands.push_back(cc);
      P.add_step(std::string(desc) + " [chain candidate]", [](RunCtx&) { fail("internal: chain candidate executed outside its chain"); },
                 (int)P.rt->chain_cands.size() - 1);
      return;
    }
    P.add_step(desc, [=](RunCtx& c) {
      TcArgs t = ta;
      t.e = epi;
      if (bias_eid >= 0) t.e.bias = (const int32_t*)c.ptr(bias_eid, kLogical);
      if (sum_eid >= 0) t.e.sum = c.ptr(sum_eid, phys);
      t.a = c.ptr(x_eid, (from_nchw || unfold) ? kLogical : phys);
      if (unfold) {
        void* x3 = c.ws->arena + scratch_off;
        c.tc_flush();
        launch_unfold_rows_nchw(t.a, x3, g, c.s);
        t.a = x3;
      }
      if (!direct && !implicit && !unfold) {
        void* col = c.ws->arena + scratch_off;
        c.tc_flush();
        if (from_nchw) launch_im2col_nchw(t.a, col, g, Kp, c.s);
        else launch_im2col(t.a, col, g, Kp, c.s);
        t.a = col;
      }
      t.y = c.ptr(y_eid, phys);
      c.tc_begin();
      if ((pair ? launch_tc_pair(t, c.s) : launch_tc(t, c.s)) != 0) fail("tcgen05 launch failed: %s", cudaGetErrorString(cudaGetLastError()));
      c.tc_end();
    });
  } else if (P.rt->conv_impl != "simt" && phys == kNHWC && Cg == 1 && all_zero(ci.zp_w) && dwconv_nhwc_supported(g)) {
    // depthwise: dedicated channels-last kernel (16 channels per thread), tap-major weights
    const int KK = g.KH * g.KW;
    std::vector<uint8_t> w_tap((size_t)KK * g.C);
    std::vector<int32_t> corr(g.C, 0);
    for (int c = 0; c < g.C; ++c) {
      long long sum = 0;
      for (int t = 0; t < KK; ++t) {
        const uint8_t b = w_ohwi[(size_t)c * KK + t];
        w_tap[(size_t)t * g.C + c] = b;
        sum += g.w_dtype == kU8 ? (int)b : (int)(int8_t)b;
      }
      corr[c] = (int32_t)(0u - (uint32_t)g.zp_a * (uint32_t)(int32_t)sum);
    }
    void* wdev = P.dev_upload(w_tap.data(), w_tap.size());
    void* wrow_dev = nullptr;
    if (g.KH == 3 && g.KW == 3) {  // row-packed weights for the dp4a kernel: word (r, c) = {w(r,0), w(r,1), w(r,2), 0}
      std::vector<uint8_t> w_row((size_t)3 * g.C * 4, 0);
      for (int c = 0; c < g.C; ++c)
        for (int rr = 0; rr < 3; ++rr)
          for (int q = 0; q < 3; ++q) w_row[((size_t)rr * g.C + c) * 4 + q] = w_ohwi[(size_t)c * KK + rr * 3 + q];
      wrow_dev = P.dev_upload(w_row.data(), w_row.size());
    }
    const int32_t* corr_dev = g.zp_a != 0 ? (const int32_t*)P.dev_upload(corr.data(), corr.size() * 4) : nullptr;
    FastEpi fe;
    memset(&fe, 0, sizeof(fe));
    if (composite && !ci.float_mode) {
      std::vector<int32_t> bias2;
      if (ci.has_const_bias || g.zp_a != 0) bias2 = add_vec(ci.bias_host, corr, g.O);
      build_fast_epi(P, ci, bias2, abs_row_sums(w_ohwi.data(), g.O, KK, g.w_dtype), g.O, Y.dtype, &fe);
    }
    if (!ci.float_mode && ci.has_const_bias) epi.bias = (const int32_t*)P.dev_upload(ci.bias_host.data(), ci.bias_host.size() * 4);
    snprintf(desc, sizeof(desc), "%s -> CUDA-core depthwise conv (NHWC, 16 ch/thread) N=%d %dx%dx%d -> %dx%d k%dx%d s%d epi=%s",
             n.name.c_str(), g.N, g.H, g.W, g.C, g.OH, g.OW, g.KH, g.KW, g.sh, fe.enabled && !fe.has_sum ? "fast" : "generic");
    P.add_step(desc, [=](RunCtx& c) {
      EpiParams e = epi;
      if (bias_eid >= 0) e.bias = (const int32_t*)c.ptr(bias_eid, kLogical);
      if (sum_eid >= 0) e.sum = c.ptr(sum_eid, phys);
      launch_dwconv_nhwc(c.ptr(x_eid, phys), wdev, wrow_dev, corr_dev, c.ptr(y_eid, phys), g, e, fe, c.s);
    });
  } else {
    void* wdev = P.dev_upload(w_ohwi.data(), w_ohwi.size());
    if (!ci.float_mode && ci.has_const_bias) epi.bias = (const int32_t*)P.dev_upload(ci.bias_host.data(), ci.bias_host.size() * 4);
    snprintf(desc, sizeof(desc), "%s -> CUDA-core direct conv N=%d %dx%dx%d -> %dx%dx%d k%dx%d g%d epi=%d", n.name.c_str(),
             g.N, g.H, g.W, g.C, g.OH, g.OW, g.O, g.KH, g.KW, g.groups, epi.mode);
    P.add_step(desc, [=](RunCtx& c) {
      EpiParams e = epi;
      if (bias_eid >= 0) e.bias = (const int32_t*)c.ptr(bias_eid, kLogical);
      if (sum_eid >= 0) e.sum = c.ptr(sum_eid, phys);
      launch_conv2d_simt(c.ptr(x_eid, phys), wdev, c.ptr(y_eid, phys), g, e, c.s);
    });
  }
}

// ------------------------------------------------------------------------------------------------
// dense / batch_matmul (Dense :462-502, BatchMatMul :504-548; arithmetic dense.cc:160-205,
// batch_matmul.cc:161-200)
// ------------------------------------------------------------------------------------------------
static void plan_dense(Planner& P, int nid, bool composite) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), w_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[x_eid];
  const Entry& Wt = P.ent[w_eid];
  Entry& Y = P.ent[y_eid];
  if (Wt.shape.size() != 2 || X.shape.empty()) fail("dense expects a 2-D weight");
  const int K = (int)Wt.shape[1], N = (int)Wt.shape[0];
  if (X.shape.back() != K) fail("dense: input has %lld features, weight expects %d", (long long)X.shape.back(), K);
  const long long M = numel(X.shape) / K;
  if ((X.dtype != kI8 && X.dtype != kU8) || (Wt.dtype != kI8 && Wt.dtype != kU8))
    fail("Unsupported op: %s with data/kernel dtypes other than int8/uint8", n.name.c_str());
  CompositeInfo ci;
  if (composite) {
    ci = build_composite(P, n, nid, N, (int)Y.shape.size(), (int)Y.shape.size() - 1);
  } else {
    init_epi(ci.epi);
    if (Y.dtype != kI32) fail("qnn.dense out_dtype must be int32");
    ci.zp_a = P.cints((int)P.g.eid(n.inputs[2]), "input_zero_point")[0];
    ci.zp_w = P.cints((int)P.g.eid(n.inputs[3]), "kernel_zero_point");
  }
  if (ci.zp_w.size() != 1) fail("Unsupported op: qnn.dense with a per-channel kernel zero point");
  const int32_t zpx = ci.float_mode ? 0 : ci.zp_a, zpw = ci.zp_w[0];
  P.need(x_eid, kLogical);
  P.need(w_eid, kLogical);
  if (ci.sum_eid >= 0) P.need(ci.sum_eid, kLogical);
  if (ci.bias_eid >= 0) P.need(ci.bias_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->macs += M * N * K;
  P.rt->hbm_bytes += X.bytes() + Wt.bytes() + Y.bytes() + (ci.sum_eid >= 0 ? Y.bytes() : 0);
  EpiParams epi = ci.epi;
  const int bias_eid = ci.bias_eid, sum_eid = ci.sum_eid, xd = X.dtype, wd = Wt.dtype;
  const bool foldable = zpx == 0 || (Wt.is_const && bias_eid < 0);
  const bool use_tc = P.rt->conv_impl != "simt" && tc_gemm_supported((int)M, N, K) && zpw == 0 && foldable &&
                      (P.rt->conv_impl == "tc" || P.rt->conv_impl == "tc_nochain" || P.rt->conv_impl.rfind("tc_chain", 0) == 0 || P.rt->conv_impl == "tc_explicit" || P.rt->conv_impl == "tc_im2col" || P.rt->conv_impl == "tc_pair" || M * N >= 4096);
  char desc[256];
  if (use_tc) {
    std::vector<int32_t> bias2;
    if (!ci.float_mode && (ci.has_const_bias || zpx != 0)) {
      std::vector<int32_t> corr(N, 0);
      if (zpx != 0)
        for (int o = 0; o < N; ++o) {
          long long sum = 0;
          for (int k = 0; k < K; ++k) sum += wd == kU8 ? (int)Wt.host[(size_t)o * K + k] : (int)(int8_t)Wt.host[(size_t)o * K + k];
          corr[o] = (int32_t)(0u - (uint32_t)zpx * (uint32_t)(int32_t)sum);
        }
      bias2 = add_vec(ci.bias_host, corr, N);
      epi.bias = (const int32_t*)P.dev_upload(bias2.data(), bias2.size() * 4);
    }
    TcArgs ta;
    memset(&ta, 0, sizeof(ta));
    ta.B = 1; ta.M = (int)M; ta.N = N; ta.K = K;
    ta.a_dtype = xd; ta.w_dtype = wd;
    if (composite && !ci.float_mode && Wt.is_const)
      build_fast_epi(P, ci, bias2, abs_row_sums(Wt.host.data(), N, K, wd), N, Y.dtype, &ta.f);
    const void* wpacked = nullptr;
    P.rt->tc_macs += (long long)M * N * K;
    const int chain_bn = 128;
    bool chain = P.rt->chain_enabled && composite && !ci.float_mode && ta.f.enabled && Wt.is_const && bias_eid < 0 && N >= 96 &&
                 (P.rt->chain_policy != 1 || has_sibling_composite(P, nid)) &&
                 M >= 128 && !P.ent[P.root(x_eid)].is_input && (sum_eid < 0 || !P.ent[P.root(sum_eid)].is_input);
    if (chain) {
      TcArgs probe = ta;
      probe.e = epi;
      chain = tc_chain_supported(probe, chain_bn);
    }
    if (chain) {
      ta.pack_bn = chain_bn; ta.pack_bk = tc_chain_bk(ta);
      std::vector<uint8_t> packed(tc_pack_weights(Wt.host.data(), N, K, ta.pack_bn, ta.pack_bk, nullptr));
      tc_pack_weights(Wt.host.data(), N, K, ta.pack_bn, ta.pack_bk, packed.data());
      ta.w = P.dev_upload(packed.data(), packed.size());
      ta.w_packed = 1;
      ChainCand cc;
      cc.ta = ta;
      cc.ta.e = epi;
      cc.x_eid = x_eid; cc.y_eid = y_eid; cc.sum_eid = sum_eid; cc.phys = kLogical;
      cc.rows_per_img = 128;  // a dense row depends on the same row of its input only: 128-row blocks are the "images"
      cc.bn = chain_bn;
      cc.name = n.name;
      P.rt->chain_cands.push_back(cc);
      snprintf(desc, sizeof(desc), "%s -> tcgen05 i8 GEMM M=%lld N=%d K=%d epi=fast [chain candidate]", n.name.c_str(), M, N, K);
      P.add_step(desc, [](RunCtx&) { fail("internal: chain candidate executed outside its chain"); }, (int)P.rt->chain_cands.size() - 1);
      return;
    }
    if (Wt.is_const) {
      tc_tile_config(ta, &ta.pack_bn, &ta.pack_bk);
      std::vector<uint8_t> packed(tc_pack_weights(Wt.host.data(), N, K, ta.pack_bn, ta.pack_bk, nullptr));
      tc_pack_weights(Wt.host.data(), N, K, ta.pack_bn, ta.pack_bk, packed.data());
      wpacked = P.dev_upload(packed.data(), packed.size());
      ta.w_packed = 1;
    }
    snprintf(desc, sizeof(desc), "%s -> tcgen05 i8 GEMM M=%lld N=%d K=%d epi=%s", n.name.c_str(), M, N, K,
             ta.f.enabled ? "fast" : (epi.mode == kEpiInt ? "int" : (epi.mode == kEpiFloat ? "float" : "raw")));
    P.add_step(desc, [=](RunCtx& c) {
      TcArgs t = ta;
      t.e = epi;
      if (bias_eid >= 0) t.e.bias = (const int32_t*)c.ptr(bias_eid, kLogical);
      if (sum_eid >= 0) t.e.sum = c.ptr(sum_eid, kLogical);
      t.a = c.ptr(x_eid, kLogical);
      t.w = wpacked ? wpacked : c.ptr(w_eid, kLogical);
      t.y = c.ptr(y_eid, kLogical);
      c.tc_begin();
      if (launch_tc(t, c.s) != 0) fail("tcgen05 launch failed: %s", cudaGetErrorString(cudaGetLastError()));
      c.tc_end();
    });
  } else {
    if (!ci.float_mode && ci.has_const_bias) epi.bias = (const int32_t*)P.dev_upload(ci.bias_host.data(), ci.bias_host.size() * 4);
    snprintf(desc, sizeof(desc), "%s -> CUDA-core GEMM M=%lld N=%d K=%d epi=%d", n.name.c_str(), M, N, K, epi.mode);
    P.add_step(desc, [=](RunCtx& c) {
      EpiParams e = epi;
      if (bias_eid >= 0) e.bias = (const int32_t*)c.ptr(bias_eid, kLogical);
      if (sum_eid >= 0) e.sum = c.ptr(sum_eid, kLogical);
      launch_gemm_simt(c.ptr(x_eid, kLogical), c.ptr(w_eid, kLogical), c.ptr(y_eid, kLogical), 1, (int)M, N, K, 0, xd, wd,
                       zpx, zpw, e, c.s);
    });
  }
}

static void plan_batch_matmul(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), w_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[x_eid];
  const Entry& Wt = P.ent[w_eid];
  if (X.shape.size() != 3 || Wt.shape.size() != 3 || X.shape[2] != Wt.shape[2])
    fail("qnn.batch_matmul expects x [B,M,K] and y [B,N,K] (transpose_b only, batch_matmul.cc:83-87)");
  if (n.i64("transpose_a", 0) != 0 || n.i64("transpose_b", 1) != 1) fail("Unsupported op: qnn.batch_matmul transpose flags");
  const int B = (int)std::max(X.shape[0], Wt.shape[0]), M = (int)X.shape[1], N = (int)Wt.shape[1], K = (int)X.shape[2];
  if (X.shape[0] != B) fail("Unsupported op: qnn.batch_matmul broadcasting the first operand");
  const int wb = Wt.shape[0] == B && B > 1 ? 1 : 0;
  if ((X.dtype != kI8 && X.dtype != kU8) || (Wt.dtype != kI8 && Wt.dtype != kU8)) fail("Unsupported op: batch_matmul dtypes");
  const int32_t zpx = P.cints((int)P.g.eid(n.inputs[2]), "x_zero_point")[0];
  const int32_t zpw = P.cints((int)P.g.eid(n.inputs[3]), "y_zero_point")[0];
  P.need(x_eid, kLogical);
  P.need(w_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->macs += (long long)B * M * N * K;
  P.rt->hbm_bytes += X.bytes() + Wt.bytes() + P.ent[y_eid].bytes();
  EpiParams epi;
  init_epi(epi);
  const int xd = X.dtype, wd = Wt.dtype;
  const bool use_tc = P.rt->conv_impl != "simt" && tc_gemm_supported(M, N, K) && zpx == 0 && zpw == 0 &&
                      (P.rt->conv_impl == "tc" || P.rt->conv_impl == "tc_nochain" || P.rt->conv_impl.rfind("tc_chain", 0) == 0 || P.rt->conv_impl == "tc_explicit" || P.rt->conv_impl == "tc_im2col" || P.rt->conv_impl == "tc_pair" || (long long)B * M * N >= 4096);
  char desc[256];
  if (use_tc) P.rt->tc_macs += (long long)B * M * N * K;
  snprintf(desc, sizeof(desc), "qnn.batch_matmul -> %s B=%d M=%d N=%d K=%d", use_tc ? "tcgen05 i8 GEMM" : "CUDA-core GEMM", B, M, N, K);
  P.add_step(desc, [=](RunCtx& c) {
    if (use_tc) {
      TcArgs t;
      memset(&t, 0, sizeof(t));
      t.a = c.ptr(x_eid, kLogical); t.w = c.ptr(w_eid, kLogical); t.y = c.ptr(y_eid, kLogical);
      t.B = B; t.M = M; t.N = N; t.K = K; t.w_batched = wb; t.a_dtype = xd; t.w_dtype = wd;
      t.e = epi;
      c.tc_begin();
      if (launch_tc(t, c.s) != 0) fail("tcgen05 launch failed: %s", cudaGetErrorString(cudaGetLastError()));
      c.tc_end();
    } else {
      launch_gemm_simt(c.ptr(x_eid, kLogical), c.ptr(w_eid, kLogical), c.ptr(y_eid, kLogical), B, M, N, K, wb, xd, wd, zpx, zpw, epi, c.s);
    }
  });
}

// ------------------------------------------------------------------------------------------------
// elementwise / pooling / data-movement operators
// ------------------------------------------------------------------------------------------------
// (outer, C, inner) decomposition of a tensor around logical axis `axis` in physical layout `phys`.
static void axis_split(const Entry& e, int axis, int phys, long long* outer, int* C, long long* inner) {
  const int rank = (int)e.shape.size();
  if (rank == 0) { *outer = 1; *C = 1; *inner = 1; return; }
  if (axis < 0) axis += rank;
  if (axis < 0 || axis >= rank) fail("axis %d out of range for rank %d", axis, rank);
  std::vector<int64_t> shp = e.shape;
  int ax = axis;
  if (phys == kNHWC && rank == 4) {
    shp = {e.shape[0], e.shape[2], e.shape[3], e.shape[1]};
    static const int map[4] = {0, 3, 1, 2};  // logical axis -> physical position
    ax = map[axis];
  }
  *outer = 1; *inner = 1;
  for (int i = 0; i < ax; ++i) *outer *= shp[i];
  *C = (int)shp[ax];
  for (int i = ax + 1; i < rank; ++i) *inner *= shp[i];
}

static void plan_requantize(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[x_eid];
  const int phys = P.preferred(x_eid);
  auto in_scl = P.cfloats((int)P.g.eid(n.inputs[1]), "input_scale");
  auto in_zp = P.cints((int)P.g.eid(n.inputs[2]), "input_zero_point");
  const float out_scl = P.cfloats((int)P.g.eid(n.inputs[3]), "output_scale")[0];
  const int32_t out_zp = P.cints((int)P.g.eid(n.inputs[4]), "output_zero_point")[0];
  long long outer, inner;
  int C;
  axis_split(X, (int)n.i64("axis", -1), phys, &outer, &C, &inner);
  if (in_scl.size() == 1 && in_zp.size() == 1) { outer *= (long long)C * inner; C = 1; inner = 1; }
  const int ydt = P.ent[y_eid].dtype, xdt = X.dtype;
  RqParams rq = make_rq(P, in_scl, in_zp, out_scl, out_zp, parse_rounding(n.str("rounding", "None")),
                        parse_compute(n.str("compute_dtype", "None")), ydt, C);
  P.need(x_eid, phys);
  P.produce(y_eid, phys);
  P.rt->hbm_bytes += X.bytes() + P.ent[y_eid].bytes();
  P.add_step("qnn.requantize -> elementwise kernel", [=](RunCtx& c) {
    launch_requantize(c.ptr(x_eid, phys), xdt, c.ptr(y_eid, phys), ydt, outer, C, inner, rq, c.s);
  });
}

static void plan_qnn_add(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int a_eid = (int)P.g.eid(n.inputs[0]), b_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& A = P.ent[a_eid];
  const bool sub = n.name == "qnn.subtract";  // subtract.cc:40-93: the same two requantizes, lhs' - rhs' + zp_out
  if (A.shape != P.ent[b_eid].shape || A.dtype != P.ent[b_eid].dtype) fail("Unsupported op: %s with broadcasting operands", n.name.c_str());
  const int phys = (P.preferred(a_eid) == kNHWC || P.preferred(b_eid) == kNHWC) && A.shape.size() == 4 &&
                           !P.ent[P.root(a_eid)].is_const && !P.ent[P.root(b_eid)].is_const ? kNHWC : kLogical;
  EpiParams e;
  init_epi(e);
  make_qadd(P, e, P.cfloats((int)P.g.eid(n.inputs[2]), "lhs_scale")[0], P.cints((int)P.g.eid(n.inputs[3]), "lhs_zero_point")[0],
            P.cfloats((int)P.g.eid(n.inputs[4]), "rhs_scale")[0], P.cints((int)P.g.eid(n.inputs[5]), "rhs_zero_point")[0],
            P.cfloats((int)P.g.eid(n.inputs[6]), "output_scale")[0], P.cints((int)P.g.eid(n.inputs[7]), "output_zero_point")[0],
            A.dtype);
  e.sum_sub = sub ? 1 : 0;
  P.need(a_eid, phys);
  P.need(b_eid, phys);
  P.produce(y_eid, phys);
  const long long cnt = numel(A.shape);
  const int dt = A.dtype;
  P.rt->hbm_bytes += 3 * A.bytes();
  P.add_step(sub ? "qnn.subtract -> vectorised elementwise kernel" : "qnn.add -> vectorised elementwise kernel", [=](RunCtx& c) {
    launch_qnn_add(c.ptr(a_eid, phys), c.ptr(b_eid, phys), c.ptr(y_eid, phys), dt, cnt, e, c.s);
  });
}

// qnn.mul, per-tensor form (src/relay/qnn/op/mul.cc:66-101)
static void plan_qnn_mul(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int a_eid = (int)P.g.eid(n.inputs[0]), b_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& A = P.ent[a_eid];
  if (A.shape != P.ent[b_eid].shape || A.dtype != P.ent[b_eid].dtype) fail("Unsupported op: qnn.mul with broadcasting operands");
  if (A.dtype == kF32 || A.dtype == kI64) fail("Unsupported op: qnn.mul on non-integer tensors");
  auto sa = P.cfloats((int)P.g.eid(n.inputs[2]), "lhs_scale"), sb = P.cfloats((int)P.g.eid(n.inputs[4]), "rhs_scale");
  if (sa.size() != 1 || sb.size() != 1) fail("Unsupported op: qnn.mul with per-channel scales");
  const int32_t zpa = P.cints((int)P.g.eid(n.inputs[3]), "lhs_zero_point")[0], zpb = P.cints((int)P.g.eid(n.inputs[5]), "rhs_zero_point")[0];
  const float so = P.cfloats((int)P.g.eid(n.inputs[6]), "output_scale")[0];
  const int32_t zpo = P.cints((int)P.g.eid(n.inputs[7]), "output_zero_point")[0];
  const float new_scale = sa[0] * sb[0];  // float32 product, mul.cc:91-93
  const int dt = A.dtype;
  RqParams rq = make_rq(P, {new_scale}, {0}, so, zpo, kUpward, kInt64, dt, 1);
  const int phys = (P.preferred(a_eid) == kNHWC || P.preferred(b_eid) == kNHWC) && A.shape.size() == 4 &&
                           !P.ent[P.root(a_eid)].is_const && !P.ent[P.root(b_eid)].is_const ? kNHWC : kLogical;
  P.need(a_eid, phys);
  P.need(b_eid, phys);
  P.produce(y_eid, phys);
  const long long cnt = numel(A.shape);
  P.rt->hbm_bytes += 3 * A.bytes();
  P.add_step("qnn.mul -> vectorised elementwise kernel", [=](RunCtx& c) {
    launch_qnn_mul(c.ptr(a_eid, phys), c.ptr(b_eid, phys), c.ptr(y_eid, phys), dt, cnt, zpa, zpb, rq, c.s);
  });
}

// qnn.quantize / qnn.dequantize at the float edges of a graph (quantize.cc:113-150, dequantize.cc:95-127)
static void plan_quantize(Planner& P, int nid, bool dequant) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[x_eid];
  const Entry& Y = P.ent[y_eid];
  if (dequant ? (X.dtype == kF32 || X.dtype == kI64 || Y.dtype != kF32) : (X.dtype != kF32 || Y.dtype == kF32 || Y.dtype == kI64))
    fail("Unsupported op: %s from dtype %d to dtype %d", n.name.c_str(), X.dtype, Y.dtype);
  auto scl = P.cfloats((int)P.g.eid(n.inputs[1]), "scale");
  auto zps = P.cints((int)P.g.eid(n.inputs[2]), "zero_point");
  long long outer, inner;
  int C;
  axis_split(X, (int)n.i64("axis", -1), kLogical, &outer, &C, &inner);
  if (scl.size() == 1 && zps.size() == 1) { outer *= (long long)C * inner; C = 1; inner = 1; }
  if ((scl.size() != 1 && (int)scl.size() != C) || (zps.size() != 1 && (int)zps.size() != C))
    fail("%s: %d scales / %d zero points for an axis of %d", n.name.c_str(), (int)scl.size(), (int)zps.size(), C);
  const float* scl_dev = scl.size() > 1 ? (const float*)P.dev_upload(scl.data(), scl.size() * 4) : nullptr;
  const int32_t* zp_dev = zps.size() > 1 ? (const int32_t*)P.dev_upload(zps.data(), zps.size() * 4) : nullptr;
  const float s0 = scl[0];
  const int32_t z0 = zps[0];
  const int ns = (int)scl.size(), nz = (int)zps.size(), xdt = X.dtype, ydt = Y.dtype;
  int32_t lo, hi;
  dtype_range(ydt, &lo, &hi);
  P.need(x_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += X.bytes() + Y.bytes();
  if (dequant)
    P.add_step("qnn.dequantize -> vectorised elementwise kernel", [=](RunCtx& c) {
      launch_dequantize(c.ptr(x_eid, kLogical), xdt, (float*)c.ptr(y_eid, kLogical), outer, C, inner, scl_dev, zp_dev, s0, z0, ns, nz, c.s);
    });
  else
    P.add_step("qnn.quantize -> vectorised elementwise kernel", [=](RunCtx& c) {
      launch_quantize((const float*)c.ptr(x_eid, kLogical), c.ptr(y_eid, kLogical), ydt, outer, C, inner, scl_dev, zp_dev, s0, z0, ns, nz, lo, hi, c.s);
    });
}

// qnn.concatenate (concatenate.cc:138-213): inputs = n tensors, n scales, n zero points, output scale, output zero point
static void plan_qnn_concatenate(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int y_eid = (int)P.g.eid(nid, 0);
  const Entry& Y = P.ent[y_eid];
  if (n.inputs.size() < 5 || (n.inputs.size() - 2) % 3 != 0) fail("qnn.concatenate: expected n tensors, n scales, n zero points, 2 output parameters");
  const int cnt = (int)(n.inputs.size() - 2) / 3;
  const int rank = (int)Y.shape.size();
  int axis = (int)n.i64("axis", 0);
  if (axis < 0) axis += rank;
  if (axis < 0 || axis >= rank) fail("qnn.concatenate: axis out of range");
  if (Y.dtype == kF32 || Y.dtype == kI64) fail("Unsupported op: qnn.concatenate on non-integer tensors");
  const float so = P.cfloats((int)P.g.eid(n.inputs[3 * cnt]), "output_scale")[0];
  const int32_t zpo = P.cints((int)P.g.eid(n.inputs[3 * cnt + 1]), "output_zero_point")[0];
  long long outer = 1, inner = 1;
  for (int i = 0; i < axis; ++i) outer *= Y.shape[i];
  for (int i = axis + 1; i < rank; ++i) inner *= Y.shape[i];
  const long long row_out = Y.shape[axis] * inner;
  P.produce(y_eid, kLogical);
  long long off = 0;
  const int dt = Y.dtype;
  for (int i = 0; i < cnt; ++i) {
    const int x_eid = (int)P.g.eid(n.inputs[i]);
    const Entry& X = P.ent[x_eid];
    if (X.dtype != dt || (int)X.shape.size() != rank) fail("qnn.concatenate: inputs must share dtype and rank");
    for (int d = 0; d < rank; ++d)
      if (d != axis && X.shape[d] != Y.shape[d]) fail("qnn.concatenate: input %d does not match the output shape", i);
    const float si = P.cfloats((int)P.g.eid(n.inputs[cnt + i]), "input_scale")[0];
    const int32_t zi = P.cints((int)P.g.eid(n.inputs[2 * cnt + i]), "input_zero_point")[0];
    const int requant = !(si == so && zi == zpo);  // IsEqualScalar on both, concatenate.cc:186-187
    RqParams rq = make_rq(P, {si}, {zi}, so, zpo, kUpward, kInt64, dt, 1);
    const long long row_in = X.shape[axis] * inner, o = off;
    P.need(x_eid, kLogical);
    P.rt->hbm_bytes += 2 * X.bytes();
    P.add_step("qnn.concatenate piece " + std::to_string(i) + (requant ? " -> requantize into the output slice" : " -> copy into the output slice"),
               [=](RunCtx& c) { launch_concat_piece(c.ptr(x_eid, kLogical), c.ptr(y_eid, kLogical), dt, outer, row_in, row_out, o, rq, requant, c.s); });
    off += row_in;
  }
  if (off != row_out) fail("qnn.concatenate: the inputs do not fill the output along axis %d", axis);
}

static void plan_unary(Planner& P, int nid) {  // clip, cast
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[x_eid];
  const int phys = P.preferred(x_eid);
  P.need(x_eid, phys);
  P.produce(y_eid, phys);
  const long long cnt = numel(X.shape);
  const int xdt = X.dtype, ydt = P.ent[y_eid].dtype;
  P.rt->hbm_bytes += X.bytes() + P.ent[y_eid].bytes();
  if (n.name == "clip") {
    double lo = n.f64("a_min"), hi = n.f64("a_max");
    if (xdt != kF32) {
      int32_t l, h;
      clip_bounds(lo, hi, xdt, &l, &h);
      lo = l; hi = h;
    }
    P.add_step("clip -> elementwise kernel", [=](RunCtx& c) { launch_clip(c.ptr(x_eid, phys), c.ptr(y_eid, phys), xdt, cnt, lo, hi, c.s); });
  } else {
    P.add_step("cast -> elementwise kernel", [=](RunCtx& c) { launch_cast(c.ptr(x_eid, phys), xdt, c.ptr(y_eid, phys), ydt, cnt, c.s); });
  }
}

static void plan_add(Planner& P, int nid) {  // nn.bias_add, add
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), b_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[x_eid];
  const Entry& Bv = P.ent[b_eid];
  if (X.dtype != Bv.dtype || X.dtype == kF32 || X.dtype == kI64) fail("Unsupported op: %s on dtypes other than matching integers", n.name.c_str());
  if (P.ent[y_eid].shape != X.shape) fail("Unsupported op: %s broadcasting the first operand", n.name.c_str());
  const int dt = X.dtype;
  const int rank = (int)X.shape.size();
  P.rt->hbm_bytes += 2 * X.bytes();
  if (n.name == "add" && Bv.shape == X.shape) {
    const int phys = P.preferred(x_eid);
    P.need(x_eid, phys);
    P.need(b_eid, phys);
    P.produce(y_eid, phys);
    const long long cnt = numel(X.shape);
    P.add_step("add -> elementwise kernel", [=](RunCtx& c) { launch_add_same(c.ptr(x_eid, phys), c.ptr(b_eid, phys), c.ptr(y_eid, phys), dt, cnt, c.s); });
    return;
  }
  // broadcast of a vector along one axis: nn.bias_add(axis) or add(x, const[C,1,1] / [C] / scalar)
  int axis;
  if (n.name == "nn.bias_add") {
    axis = (int)n.i64("axis", 1);
    if (axis < 0) axis += rank;
  } else {
    const int br = (int)Bv.shape.size();
    axis = -1;
    for (int i = 0; i < br; ++i) {
      if (Bv.shape[i] == 1) continue;
      if (axis >= 0) fail("Unsupported op: add with a multi-axis broadcast operand");
      axis = rank - br + i;
    }
    if (axis < 0) axis = rank > 0 ? rank - 1 : 0;  // scalar: broadcast over everything
  }
  const int phys = P.preferred(x_eid);
  long long outer, inner;
  int C;
  axis_split(X, axis, phys, &outer, &C, &inner);
  const long long bn = numel(Bv.shape);
  if (bn == 1) { outer *= (long long)C * inner; C = 1; inner = 1; }
  else if (bn != C) fail("%s: operand has %lld elements, axis has %d", n.name.c_str(), bn, C);
  P.need(x_eid, phys);
  P.need(b_eid, kLogical);
  P.produce(y_eid, phys);
  P.add_step(n.name + " -> elementwise kernel", [=](RunCtx& c) {
    launch_bias_add(c.ptr(x_eid, phys), c.ptr(b_eid, kLogical), c.ptr(y_eid, phys), dt, outer, C, inner, c.s);
  });
}

static void plan_pool(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  const Entry& Y = P.ent[y_eid];
  const std::string lay = n.str("layout", "NCHW");
  if ((lay != "NCHW" && lay != "NHWC") || X.shape.size() != 4) fail("Unsupported op: %s with layout %s", n.name.c_str(), lay.c_str());
  if (X.dtype == kF32 || X.dtype == kI64) fail("Unsupported op: %s on non-integer tensors", n.name.c_str());
  const int N = (int)X.shape[lay.find('N')], C = (int)X.shape[lay.find('C')], H = (int)X.shape[lay.find('H')], W = (int)X.shape[lay.find('W')];
  int KH, KW, sh = 1, sw = 1, pt = 0, pl = 0, pb = 0, pr = 0, cip = 0;
  const bool global = n.name == "nn.global_avg_pool2d";
  const bool is_avg = global || n.name == "nn.avg_pool2d";
  if (global) { KH = H; KW = W; }
  else {
    auto ps = n.ints("pool_size"), st = n.ints("strides"), pd = n.ints("padding"), di = n.ints("dilation");
    if (ps.size() != 2 || st.size() != 2) fail("%s: pool_size / strides must have 2 elements", n.name.c_str());
    for (auto d : di) if (d != 1) fail("Unsupported op: dilated pooling");
    if (n.i64("ceil_mode", 0)) fail("Unsupported op: pooling with ceil_mode");
    KH = (int)ps[0]; KW = (int)ps[1]; sh = (int)st[0]; sw = (int)st[1];
    if (pd.size() == 4) { pt = (int)pd[0]; pl = (int)pd[1]; pb = (int)pd[2]; pr = (int)pd[3]; }
    else if (pd.size() == 2) { pt = pb = (int)pd[0]; pl = pr = (int)pd[1]; }
    else if (pd.size() == 1) { pt = pb = pl = pr = (int)pd[0]; }
    cip = (int)n.i64("count_include_pad", 0);
  }
  const int OH = (H + pt + pb - KH) / sh + 1, OW = (W + pl + pr - KW) / sw + 1;
  if (Y.shape[lay.find('H')] != OH || Y.shape[lay.find('W')] != OW) fail("%s: output shape mismatch", n.name.c_str());
  const int phys = lay == "NCHW" ? kNHWC : kLogical;  // channels-last in memory either way
  P.need(x_eid, phys);
  P.produce(y_eid, phys);
  const int dt = X.dtype;
  // The reference reduces in the tensor dtype and, for the adaptive (global) form, multiplies the divisor up in that
  // dtype too (include/tvm/topi/nn/pooling.h:384-389, :599): a divisor that wraps to zero is a division by zero there.
  const int avg_kind = global ? 2 : (is_avg ? 1 : 0);
  if (global) {
    auto wrap = [&](long long v) -> long long {
      return dt == kI8 ? (int8_t)v : dt == kU8 ? (uint8_t)v : dt == kI16 ? (int16_t)v : (int32_t)v;
    };
    if (wrap(wrap(KH) * wrap(KW)) == 0)
      fail("Unsupported op: %s over %dx%d on a %d-bit tensor (the divisor wraps to zero in the tensor dtype)", n.name.c_str(), KH, KW, 8 * dtype_size(dt));
  }
  std::array<long long, 4> xs = {(long long)H * W * C, 1, (long long)W * C, C}, ys = {(long long)OH * OW * C, 1, (long long)OW * C, C};
  P.rt->hbm_bytes += X.bytes() + Y.bytes();
  P.add_step(n.name + " -> pooling kernel (channels-last)", [=](RunCtx& c) {
    launch_pool2d(c.ptr(x_eid, phys), c.ptr(y_eid, phys), dt, avg_kind, cip, N, C, H, W, KH, KW, sh, sw, pt, pl, pb, pr, xs.data(), ys.data(), c.s);
  });
}

static void plan_pad(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[x_eid];
  const int rank = (int)X.shape.size();
  if (rank > 4 || n.str("pad_mode", "constant") != "constant") fail("Unsupported op: nn.pad (rank %d, mode %s)", rank, n.str("pad_mode").c_str());
  auto pw = n.ints("pad_width");
  if ((int)pw.size() != 2 * rank) fail("nn.pad: pad_width must have 2*rank entries");
  const Entry& V = P.cst((int)P.g.eid(n.inputs[1]), "pad_value");
  int32_t value = V.dtype == kF32 ? (int32_t)((const float*)V.host.data())[0] : P.cints((int)P.g.eid(n.inputs[1]), "pad_value")[0];
  if (X.dtype == kF32) {  // float tensors: the kernel writes the pad value's bit pattern
    const float fv = V.dtype == kF32 ? ((const float*)V.host.data())[0] : (float)value;
    memcpy(&value, &fv, 4);
  }
  std::array<long long, 4> ishape = {1, 1, 1, 1}, before = {0, 0, 0, 0}, oshape = {1, 1, 1, 1};
  for (int i = 0; i < rank; ++i) {
    const int d = 4 - rank + i;
    ishape[d] = X.shape[i]; before[d] = pw[2 * i]; oshape[d] = X.shape[i] + pw[2 * i] + pw[2 * i + 1];
  }
  P.need(x_eid, kLogical);
  P.produce(y_eid, kLogical);
  const int elem = dtype_size(X.dtype);
  P.rt->hbm_bytes += X.bytes() + P.ent[y_eid].bytes();
  P.add_step("nn.pad -> pad kernel", [=](RunCtx& c) {
    launch_pad4(c.ptr(x_eid, kLogical), c.ptr(y_eid, kLogical), elem, ishape.data(), before.data(), oshape.data(), value, c.s);
  });
}

static void plan_reshape(Planner& P, int nid) {  // reshape, nn.batch_flatten: alias, no kernel
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  if (numel(P.ent[x_eid].shape) != numel(P.ent[y_eid].shape)) fail("%s: element count changes", n.name.c_str());
  P.need(x_eid, kLogical);
  P.ent[y_eid].alias_of = P.root(x_eid);
  P.plan << "  [" << nid << "] " << n.name << " -> alias of e" << P.root(x_eid) << " (no kernel)\n";
}

// reinterpret: same bytes under another dtype of the same width -> alias, no kernel
static void plan_reinterpret(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  if (P.ent[x_eid].shape != P.ent[y_eid].shape || dtype_size(P.ent[x_eid].dtype) != dtype_size(P.ent[y_eid].dtype))
    fail("Unsupported op: reinterpret between dtypes of different width");
  P.need(x_eid, kLogical);
  P.ent[y_eid].alias_of = P.root(x_eid);
  P.plan << "  [" << nid << "] reinterpret -> alias of e" << P.root(x_eid) << " (no kernel)\n";
}

// take(table, indices, axis = 0, mode = "fast") with a constant 1-D table and 8-bit indices: what the table-lookup
// unary QNN operators legalize to (python/tvm/relay/qnn/op/canonicalizations.py:106-160)
static void plan_take(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int t_eid = (int)P.g.eid(n.inputs[0]), i_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& Tb = P.ent[P.root(t_eid)];
  const Entry& I = P.ent[i_eid];
  if (!Tb.is_const || Tb.shape.size() != 1 || Tb.shape[0] > 256 || n.i64("axis", 0) != 0 || n.i64("batch_dims", 0) != 0)
    fail("Unsupported op: take other than a constant table of <= 256 entries along axis 0");
  if (I.dtype != kU8) fail("Unsupported op: take with indices that are not uint8 (reinterpret first)");
  if (Tb.shape[0] < 256 && n.str("mode", "fast") != "fast") fail("Unsupported op: take with a short table outside mode=fast");
  const int elem = dtype_size(Tb.dtype), entries = (int)Tb.shape[0];
  if (elem > 4 || Tb.dtype != P.ent[y_eid].dtype || P.ent[y_eid].shape != I.shape) fail("Unsupported op: take with this table dtype / shape");
  const int phys = P.preferred(i_eid);
  P.need(i_eid, phys);
  P.need(t_eid, kLogical);
  P.produce(y_eid, phys);
  const long long cnt = numel(I.shape);
  P.rt->hbm_bytes += I.bytes() + P.ent[y_eid].bytes();
  P.add_step("take -> 256-entry table lookup kernel", [=](RunCtx& c) {
    launch_lut8(c.ptr(i_eid, phys), c.ptr(t_eid, kLogical), entries, elem, c.ptr(y_eid, phys), cnt, c.s);
  });
}

// qnn.leaky_relu (src/relay/qnn/op/leaky_relu.cc:70-139), per-tensor parameters
static void plan_qnn_leaky_relu(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[x_eid];
  if (X.dtype == kF32 || X.dtype == kI64 || X.dtype != P.ent[y_eid].dtype) fail("Unsupported op: qnn.leaky_relu on this dtype");
  auto si = P.cfloats((int)P.g.eid(n.inputs[1]), "input_scale"), so = P.cfloats((int)P.g.eid(n.inputs[3]), "output_scale");
  auto zi = P.cints((int)P.g.eid(n.inputs[2]), "input_zero_point"), zo = P.cints((int)P.g.eid(n.inputs[4]), "output_zero_point");
  if (si.size() != 1 || so.size() != 1 || zi.size() != 1 || zo.size() != 1) fail("Unsupported op: qnn.leaky_relu with per-channel parameters");
  const double alpha = n.f64("alpha", 0.0);
  int32_t ma, sa, mz, sz;
  fixed_point_multiplier_shift(alpha, &ma, &sa);        // GetFixedPointMultiplierShift(alpha)
  fixed_point_multiplier_shift(1.0 - alpha, &mz, &sz);  // ... (1 - alpha)
  const int32_t scaled_z = device_fixed_point_multiply(zo[0], mz, sz);
  const int identity = si[0] == so[0] && zi[0] == zo[0];  // RequantizeOrUpcast, op_common.h:196-210
  RqParams rq = make_rq(P, si, zi, so[0], zo[0], kUpward, kInt64, kI32, 1);
  int32_t lo, hi;
  dtype_range(X.dtype, &lo, &hi);
  const int phys = P.preferred(x_eid);
  P.need(x_eid, phys);
  P.produce(y_eid, phys);
  const long long cnt = numel(X.shape);
  const int dt = X.dtype;
  const int32_t zp_in = zi[0];
  P.rt->hbm_bytes += 2 * X.bytes();
  P.add_step("qnn.leaky_relu -> elementwise kernel", [=](RunCtx& c) {
    launch_qnn_leaky_relu(c.ptr(x_eid, phys), c.ptr(y_eid, phys), dt, cnt, rq, identity, zp_in, ma, sa, scaled_z, lo, hi, c.s);
  });
}

// cast(8-bit -> int32) -> nn.global_avg_pool2d -> cast(-> 8-bit), each with a single consumer (the frontend's
// apply_with_upcast tail of ResNet / MobileNet): one launch instead of three.  Returns the two nodes whose work the step of
// `nid` already does (the main loop only emits their trace records), or an empty list if the pattern is not there.
static std::vector<int> plan_upcast_gap(Planner& P, int nid) {
  static const bool off = getenv("TKB_NO_GAP_FUSION") != nullptr;
  Graph& g = P.g;
  auto sole_consumer = [&](int producer) -> int {
    int found = -1, count = 0;
    for (size_t k = 0; k < g.nodes.size(); ++k)
      for (auto& in : g.nodes[k].inputs)
        if ((int)in[0] == producer && in[1] == 0) { found = (int)k; ++count; }
    for (auto& h : g.heads)
      if ((int)h[0] == producer) ++count;
    return count == 1 ? found : -1;
  };
  const GNode& c1 = g.nodes[nid];
  if (off || c1.name != "cast" || c1.inputs.size() != 1) return {};
  const int pool_nid = sole_consumer(nid);
  if (pool_nid < 0 || g.nodes[pool_nid].name != "nn.global_avg_pool2d") return {};
  const int c2_nid = sole_consumer(pool_nid);
  if (c2_nid < 0 || g.nodes[c2_nid].name != "cast") return {};
  const int x_eid = (int)g.eid(c1.inputs[0]), e1 = (int)g.eid((uint32_t)nid, 0), e2 = (int)g.eid((uint32_t)pool_nid, 0), e3 = (int)g.eid((uint32_t)c2_nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  const std::string lay = g.nodes[pool_nid].str("layout", "NCHW");
  if (X.shape.size() != 4 || (X.dtype != kU8 && X.dtype != kI8) || P.ent[e1].dtype != kI32 || P.ent[e2].dtype != kI32 ||
      (P.ent[e3].dtype != kU8 && P.ent[e3].dtype != kI8) || (lay != "NCHW" && lay != "NHWC"))
    return {};
  const int N = (int)X.shape[0], C = (int)X.shape[lay.find('C')], HW = (int)(X.shape[lay.find('H')] * X.shape[lay.find('W')]);
  if (C % 4 != 0 || N > 65535 || HW <= 0) return {};
  const int phys = lay == "NCHW" ? kNHWC : kLogical;  // channels-last in memory either way
  P.need(x_eid, phys);
  P.produce(e1, phys);
  P.produce(e2, phys);
  P.produce(e3, phys);
  const int xdt = X.dtype;
  P.rt->hbm_bytes += X.bytes() + P.ent[e2].bytes() + P.ent[e3].bytes();
  P.add_step("cast -> nn.global_avg_pool2d -> cast in one launch (int32 copy of the input written for traces only)", [=](RunCtx& c) {
    launch_upcast_gap(c.ptr(x_eid, phys), xdt, c.mode > 0 ? c.ptr(e1, phys) : nullptr, c.ptr(e2, phys), c.ptr(e3, phys), N, HW, C, c.s);
  });
  return {pool_nid, c2_nid};
}

// ------------------------------------------------------------------------------------------------
// float32 operators (the other half of the reference's op list, tachikoma_json_runtime.cc:240-290): plain fp32 kernels
// (kernels_float.cu), always in the logical (NC[D]HW / row-major) layout of the JSON graph
// ------------------------------------------------------------------------------------------------
static bool name_has(const std::string& s, const char* sub) { return s.find(sub) != std::string::npos; }

// post-ops from the composite's name, as ParseAttrs does with its regular expressions (:185-233)
static FPost float_post_from_name(const GNode& n) {
  FPost post = {kFActNone, 0.f, 0.f, nullptr, nullptr};
  const std::string& op = n.name;
  if (name_has(op, "_relu")) post.act = kFActRelu;
  if (name_has(op, "_tanh")) post.act = kFActTanh;
  if (name_has(op, "_clip")) { post.act = kFActClip; post.a = (float)n.f64("a_min"); post.b = (float)n.f64("a_max"); }
  if (name_has(op, "_sigmoid")) post.act = kFActSigmoid;
  if (name_has(op, "_swish")) post.act = kFActSwish;
  if (name_has(op, "_gelu")) post.act = kFActGelu;
  if (name_has(op, "_mish")) post.act = kFActMish;
  return post;
}

static void require_f32(Planner& P, const GNode& n, std::initializer_list<int> eids) {
  for (int e : eids)
    if (P.ent[P.root(e)].dtype != kF32) fail("Unsupported op: %s on tensors that are not float32", n.name.c_str());
}

// bias (input 2 of *_bias* composites) and sum (last input of *_sum* composites): tachikoma_json_runtime.cc:231, :372-376
static void float_bias_sum(Planner& P, const GNode& n, int channels, long long out_elems, int* bias_eid, int* sum_eid) {
  *bias_eid = *sum_eid = -1;
  if (name_has(n.name, "_bias")) {
    if (n.inputs.size() < 3) fail("%s: missing bias input", n.name.c_str());
    *bias_eid = (int)P.g.eid(n.inputs[2]);
    if (numel(P.ent[P.root(*bias_eid)].shape) != channels) fail("%s: bias of %lld elements for %d channels", n.name.c_str(), numel(P.ent[P.root(*bias_eid)].shape), channels);
  }
  if (name_has(n.name, "_sum")) {
    *sum_eid = (int)P.g.eid(n.inputs.back());
    if (numel(P.ent[P.root(*sum_eid)].shape) != out_elems) fail("%s: sum operand must have the output's shape", n.name.c_str());
  }
}

static void plan_float_conv(Planner& P, int nid, bool transposed) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), w_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  const Entry& Wt = P.ent[P.root(w_eid)];
  const Entry& Y = P.ent[y_eid];
  require_f32(P, n, {x_eid, w_eid, y_eid});
  const int sp = (int)X.shape.size() - 2;  // spatial rank
  if (sp < 1 || sp > 3 || Wt.shape.size() != X.shape.size() || Y.shape.size() != X.shape.size()) fail("Unsupported op: %s on rank-%d data", n.name.c_str(), (int)X.shape.size());
  static const char* dls[3] = {"NCW", "NCHW", "NCDHW"};
  static const char* kls[3] = {"OIW", "OIHW", "OIDHW"};
  static const char* tls[3] = {"IOW", "IOHW", "IODHW"};
  const std::string dl = n.str("data_layout", dls[sp - 1]), kl = n.str("kernel_layout", transposed ? tls[sp - 1] : kls[sp - 1]);
  std::string ol = n.str("out_layout", "");
  if (ol.empty()) ol = dl;
  // conv3d_transpose says OIDHW and means IODHW (the reference's own work-around, tachikoma_json_runtime.cc:420-423)
  const bool kl_ok = transposed ? (kl == tls[sp - 1] || (sp == 3 && kl == "OIDHW")) : kl == kls[sp - 1];
  if (dl != dls[sp - 1] || ol != dl || !kl_ok) fail("Unsupported op: %s with data_layout %s / kernel_layout %s / out_layout %s", n.name.c_str(), dl.c_str(), kl.c_str(), ol.c_str());
  auto st = n.ints("strides"), di = n.ints("dilation"), pd = n.ints("padding");
  if (st.empty()) st.assign(sp, 1);
  if (di.empty()) di.assign(sp, 1);
  if (pd.empty()) pd.assign(2 * sp, 0);
  if ((int)pd.size() == sp) { auto q = pd; pd.insert(pd.end(), q.begin(), q.end()); }
  if ((int)pd.size() == 1) pd.assign(2 * sp, pd[0]);
  if ((int)st.size() != sp || (int)di.size() != sp || (int)pd.size() != 2 * sp) fail("%s: strides / dilation / padding do not match %d spatial axes", n.name.c_str(), sp);
  FConvGeom g = {};
  g.groups = (int)n.i64("groups", 1);
  g.transposed = transposed ? 1 : 0;
  g.N = (int)X.shape[0]; g.C = (int)X.shape[1]; g.O = (int)Y.shape[1];
  int in[3] = {1, 1, 1}, out[3] = {1, 1, 1}, k[3] = {1, 1, 1}, s3[3] = {1, 1, 1}, d3[3] = {1, 1, 1}, p3[3] = {0, 0, 0}, pe[3] = {0, 0, 0};
  for (int i = 0; i < sp; ++i) {
    const int j = 3 - sp + i;
    in[j] = (int)X.shape[2 + i]; out[j] = (int)Y.shape[2 + i]; k[j] = (int)Wt.shape[2 + i];
    s3[j] = (int)st[i]; d3[j] = (int)di[i]; p3[j] = (int)pd[i]; pe[j] = (int)pd[sp + i];
  }
  g.D = in[0]; g.H = in[1]; g.W = in[2]; g.OD = out[0]; g.OH = out[1]; g.OW = out[2]; g.KD = k[0]; g.KH = k[1]; g.KW = k[2];
  g.sd = s3[0]; g.sh = s3[1]; g.sw = s3[2]; g.dd = d3[0]; g.dh = d3[1]; g.dw = d3[2]; g.pd = p3[0]; g.ph = p3[1]; g.pw = p3[2];
  if (g.groups < 1 || g.C % g.groups || g.O % g.groups) fail("%s: %d groups for %d -> %d channels", n.name.c_str(), g.groups, g.C, g.O);
  if (transposed ? (Wt.shape[0] != g.C || Wt.shape[1] != g.O / g.groups) : (Wt.shape[0] != g.O || Wt.shape[1] != g.C / g.groups))
    fail("%s: weight shape does not match %d -> %d channels in %d groups", n.name.c_str(), g.C, g.O, g.groups);
  auto opad = n.ints("output_padding");
  for (int j = 0; j < 3; ++j) {
    const int i = j - (3 - sp);
    const int op_ = transposed && i >= 0 && i < (int)opad.size() ? (int)opad[i] : 0;
    const int want = transposed ? (in[j] - 1) * s3[j] - p3[j] - pe[j] + d3[j] * (k[j] - 1) + 1 + op_
                                : (in[j] + p3[j] + pe[j] - d3[j] * (k[j] - 1) - 1) / s3[j] + 1;
    if (i >= 0 && want != out[j]) fail("%s: output extent %d on spatial axis %d, expected %d", n.name.c_str(), out[j], i, want);
  }
  FPost post = float_post_from_name(n);
  int bias_eid, sum_eid;
  float_bias_sum(P, n, g.O, numel(Y.shape), &bias_eid, &sum_eid);
  P.need(x_eid, kLogical); P.need(w_eid, kLogical);
  if (bias_eid >= 0) P.need(bias_eid, kLogical);
  if (sum_eid >= 0) P.need(sum_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += X.bytes() + Wt.bytes() + Y.bytes();
  P.add_step(n.name + (transposed ? " -> fp32 direct transposed convolution" : fconv_gemm_form(g) ? " -> fp32 implicit-GEMM convolution (64x64x16 tiles)" : " -> fp32 direct convolution"), [=](RunCtx& c) {
    FPost q = post;
    q.bias = bias_eid >= 0 ? (const float*)c.ptr(bias_eid, kLogical) : nullptr;
    q.sum = sum_eid >= 0 ? (const float*)c.ptr(sum_eid, kLogical) : nullptr;
    launch_fconv((const float*)c.ptr(x_eid, kLogical), (const float*)c.ptr(w_eid, kLogical), (float*)c.ptr(y_eid, kLogical), g, q, c.s);
  });
}

static void plan_float_dense(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), w_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  const Entry& Wt = P.ent[P.root(w_eid)];
  const Entry& Y = P.ent[y_eid];
  require_f32(P, n, {x_eid, w_eid, y_eid});
  if (Wt.shape.size() != 2 || X.shape.empty() || X.shape.back() != Wt.shape[1]) fail("%s: data / weight shapes do not match", n.name.c_str());
  const int K = (int)Wt.shape[1], N = (int)Wt.shape[0];
  const int M = (int)(numel(X.shape) / K);
  if (numel(Y.shape) != (long long)M * N) fail("%s: output shape mismatch", n.name.c_str());
  FPost post = float_post_from_name(n);
  int bias_eid, sum_eid;
  float_bias_sum(P, n, N, numel(Y.shape), &bias_eid, &sum_eid);
  P.need(x_eid, kLogical); P.need(w_eid, kLogical);
  if (bias_eid >= 0) P.need(bias_eid, kLogical);
  if (sum_eid >= 0) P.need(sum_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += X.bytes() + Wt.bytes() + Y.bytes();
  P.add_step(n.name + " -> fp32 tiled GEMM", [=](RunCtx& c) {
    FPost q = post;
    q.bias = bias_eid >= 0 ? (const float*)c.ptr(bias_eid, kLogical) : nullptr;
    q.sum = sum_eid >= 0 ? (const float*)c.ptr(sum_eid, kLogical) : nullptr;
    launch_fgemm((const float*)c.ptr(x_eid, kLogical), (const float*)c.ptr(w_eid, kLogical), (float*)c.ptr(y_eid, kLogical), 1, M, N, K,
                 0, K, 1, 0, K, 1, q, c.s);
  });
}

// nn.batch_matmul on float32 (tachikoma_json_runtime.cc:504-548): [B,M,K] x [B,N,K]^T by default, either side transposable
static void plan_float_batch_matmul(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), w_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  const Entry& Wt = P.ent[P.root(w_eid)];
  const Entry& Y = P.ent[y_eid];
  require_f32(P, n, {x_eid, w_eid, y_eid});
  if (X.shape.size() != 3 || Wt.shape.size() != 3 || Y.shape.size() != 3) fail("%s: operands must have rank 3", n.name.c_str());
  const bool ta = n.i64("transpose_a", 0) != 0, tb = n.i64("transpose_b", 1) != 0;
  const int B = (int)Y.shape[0], M = (int)Y.shape[1], N = (int)Y.shape[2];
  const int K = (int)(ta ? X.shape[1] : X.shape[2]);
  if ((ta ? X.shape[2] : X.shape[1]) != M || (tb ? Wt.shape[1] : Wt.shape[2]) != N || (tb ? Wt.shape[2] : Wt.shape[1]) != K ||
      (X.shape[0] != B && X.shape[0] != 1) || (Wt.shape[0] != B && Wt.shape[0] != 1))
    fail("%s: operand shapes do not match", n.name.c_str());
  const long long a_b = X.shape[0] == 1 ? 0 : (long long)M * K, b_b = Wt.shape[0] == 1 ? 0 : (long long)N * K;
  const long long a_m = ta ? 1 : K, a_k = ta ? M : 1, b_n = tb ? K : 1, b_k = tb ? 1 : N;
  P.need(x_eid, kLogical); P.need(w_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += X.bytes() + Wt.bytes() + Y.bytes();
  P.add_step("nn.batch_matmul -> fp32 tiled GEMM (batched)", [=](RunCtx& c) {
    FPost q = {kFActNone, 0.f, 0.f, nullptr, nullptr};
    launch_fgemm((const float*)c.ptr(x_eid, kLogical), (const float*)c.ptr(w_eid, kLogical), (float*)c.ptr(y_eid, kLogical), B, M, N, K,
                 a_b, a_m, a_k, b_b, b_n, b_k, q, c.s);
  });
}

static void plan_float_unary(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  require_f32(P, n, {x_eid, y_eid});
  static const std::map<std::string, int> acts = {{"abs", kFActAbs}, {"exp", kFActExp}, {"log", kFActLog}, {"sqrt", kFActSqrt},
      {"round", kFActRound}, {"nn.relu", kFActRelu}, {"nn.leaky_relu", kFActLeakyRelu}, {"tanh", kFActTanh},
      {"sigmoid", kFActSigmoid}, {"clip", kFActClip}};
  const int act = acts.at(n.name);
  float a = 0.f, b = 0.f;
  if (act == kFActClip) { a = (float)n.f64("a_min"); b = (float)n.f64("a_max"); }
  if (act == kFActLeakyRelu) a = (float)n.f64("alpha", 0.01);
  const long long cnt = numel(X.shape);
  P.need(x_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += 2 * X.bytes();
  P.add_step(n.name + " -> fp32 elementwise kernel", [=](RunCtx& c) {
    launch_funary((const float*)c.ptr(x_eid, kLogical), (float*)c.ptr(y_eid, kLogical), cnt, act, a, b, c.s);
  });
}

static void plan_float_binary(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int l_eid = (int)P.g.eid(n.inputs[0]), r_eid = (int)P.g.eid(n.inputs[1]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& L = P.ent[P.root(l_eid)];
  const Entry& Rr = P.ent[P.root(r_eid)];
  const Entry& Y = P.ent[y_eid];
  require_f32(P, n, {l_eid, r_eid, y_eid});
  const int rank = (int)Y.shape.size();
  if (rank > 5 || (int)L.shape.size() > rank || (int)Rr.shape.size() > rank) fail("Unsupported op: %s of rank %d", n.name.c_str(), rank);
  FBcast5 bc;
  for (int d = 0; d < 5; ++d) { bc.shape[d] = 1; bc.ls[d] = bc.rs[d] = 0; }
  auto fill = [&](const Entry& e, long long* strides) {
    long long stride = 1;
    for (int i = (int)e.shape.size() - 1, d = 4; i >= 0; --i, --d) {
      const long long ext = e.shape[i], oext = Y.shape[rank - ((int)e.shape.size() - i)];
      if (ext != oext && ext != 1) fail("%s: operand shapes do not broadcast", n.name.c_str());
      strides[d] = ext == 1 ? 0 : stride;
      stride *= ext;
    }
  };
  for (int i = rank - 1, d = 4; i >= 0; --i, --d) bc.shape[d] = Y.shape[i];
  fill(L, bc.ls);
  fill(Rr, bc.rs);
  const int mul = n.name == "multiply";
  P.need(l_eid, kLogical); P.need(r_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += L.bytes() + Rr.bytes() + Y.bytes();
  P.add_step(n.name + " -> fp32 broadcasting binary kernel", [=](RunCtx& c) {
    launch_fbinary((const float*)c.ptr(l_eid, kLogical), (const float*)c.ptr(r_eid, kLogical), (float*)c.ptr(y_eid, kLogical), bc, mul, c.s);
  });
}

// nn.batch_norm (inference; tachikoma_json_runtime.cc:550-596): output 0 only, axis = 1, center and scale
static void plan_float_batch_norm(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  if (n.inputs.size() != 5) fail("nn.batch_norm: expected 5 inputs");
  int e[5];
  for (int i = 0; i < 5; ++i) e[i] = (int)P.g.eid(n.inputs[i]);
  const int y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(e[0])];
  require_f32(P, n, {e[0], e[1], e[2], e[3], e[4], y_eid});
  if (n.i64("axis", 1) != 1 || !n.i64("center", 1) || !n.i64("scale", 1)) fail("Unimplemented BatchNorm case");  // the reference's ICHECK
  if (X.shape.size() < 2) fail("nn.batch_norm: rank-%d data", (int)X.shape.size());
  const int C = (int)X.shape[1];
  for (int i = 1; i < 5; ++i) if (numel(P.ent[P.root(e[i])].shape) != C) fail("nn.batch_norm: statistics must have %d elements", C);
  const long long outer = X.shape[0], inner = numel(X.shape) / (outer * C);
  const float eps = (float)n.f64("epsilon", 1e-5);
  for (int i = 0; i < 5; ++i) P.need(e[i], kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += 2 * X.bytes();
  const int x = e[0], ga = e[1], be = e[2], me = e[3], va = e[4];
  P.add_step("nn.batch_norm -> fp32 per-channel normalisation kernel", [=](RunCtx& c) {
    launch_fbatchnorm((const float*)c.ptr(x, kLogical), (float*)c.ptr(y_eid, kLogical), outer, C, inner, (const float*)c.ptr(ga, kLogical),
                      (const float*)c.ptr(be, kLogical), (const float*)c.ptr(me, kLogical), (const float*)c.ptr(va, kLogical), eps, c.s);
  });
}

static void plan_float_layer_norm(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  if (n.inputs.size() != 3) fail("nn.layer_norm: expected 3 inputs");
  const int x_eid = (int)P.g.eid(n.inputs[0]), g_eid = (int)P.g.eid(n.inputs[1]), b_eid = (int)P.g.eid(n.inputs[2]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  require_f32(P, n, {x_eid, g_eid, b_eid, y_eid});
  const int rank = (int)X.shape.size();
  const int axis = (int)n.i64("axis", -1);
  if ((axis != -1 && axis != rank - 1) || !n.i64("center", 1) || !n.i64("scale", 1)) fail("Unimplemented LayerNorm case");  // the reference's ICHECK
  const int C = (int)X.shape.back();
  if (numel(P.ent[P.root(g_eid)].shape) != C || numel(P.ent[P.root(b_eid)].shape) != C) fail("nn.layer_norm: gamma / beta must have %d elements", C);
  const long long rows = numel(X.shape) / C;
  const float eps = (float)n.f64("epsilon", 1e-5);
  P.need(x_eid, kLogical); P.need(g_eid, kLogical); P.need(b_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += 2 * X.bytes();
  P.add_step("nn.layer_norm -> fp32 warp-per-row kernel", [=](RunCtx& c) {
    launch_flayernorm((const float*)c.ptr(x_eid, kLogical), (float*)c.ptr(y_eid, kLogical), rows, C, (const float*)c.ptr(g_eid, kLogical),
                      (const float*)c.ptr(b_eid, kLogical), eps, c.s);
  });
}

static void plan_float_softmax(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  require_f32(P, n, {x_eid, y_eid});
  const int rank = (int)X.shape.size();
  int axis = (int)n.i64("axis", -1);
  if (axis < 0) axis += rank;
  if (axis < 0 || axis >= rank) fail("nn.softmax: axis %d of a rank-%d tensor", axis, rank);
  long long outer = 1, inner = 1;
  for (int i = 0; i < axis; ++i) outer *= X.shape[i];
  for (int i = axis + 1; i < rank; ++i) inner *= X.shape[i];
  const int C = (int)X.shape[axis];
  P.need(x_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += 2 * X.bytes();
  P.add_step("nn.softmax -> fp32 warp-per-vector kernel", [=](RunCtx& c) {
    launch_fsoftmax((const float*)c.ptr(x_eid, kLogical), (float*)c.ptr(y_eid, kLogical), outer, C, inner, c.s);
  });
}

// float max / average pooling, 2-D and 3-D (tachikoma_json_runtime.cc:641-701)
static void plan_float_pool(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[P.root(x_eid)];
  const Entry& Y = P.ent[y_eid];
  require_f32(P, n, {x_eid, y_eid});
  const int sp = (int)X.shape.size() - 2;
  if (sp < 2 || sp > 3) fail("Unsupported op: %s on rank-%d data", n.name.c_str(), (int)X.shape.size());
  const std::string lay = n.str("layout", sp == 2 ? "NCHW" : "NCDHW");
  std::string ol = n.str("out_layout", "");
  if (ol.empty()) ol = lay;
  if (lay != (sp == 2 ? "NCHW" : "NCDHW") || ol != lay) fail("Unsupported op: %s with layout %s", n.name.c_str(), lay.c_str());
  const bool global = name_has(n.name, "global");
  const bool is_avg = name_has(n.name, "avg_pool");
  std::vector<int64_t> ps, st, di, pd;
  if (global) {
    for (int i = 0; i < sp; ++i) ps.push_back(X.shape[2 + i]);
    st.assign(sp, 1); di.assign(sp, 1); pd.assign(2 * sp, 0);
  } else {
    ps = n.ints("pool_size"); st = n.ints("strides"); di = n.ints("dilation"); pd = n.ints("padding");
    if (di.empty()) di.assign(sp, 1);
    if ((int)pd.size() == sp) { auto q = pd; pd.insert(pd.end(), q.begin(), q.end()); }
    if ((int)pd.size() == 1) pd.assign(2 * sp, pd[0]);
    if (n.i64("ceil_mode", 0)) fail("Unsupported op: pooling with ceil_mode");  // tachikoma.py:77-80 keeps these off the backend
  }
  if ((int)ps.size() != sp || (int)st.size() != sp || (int)di.size() != sp || (int)pd.size() != 2 * sp) fail("%s: pool_size / strides / padding do not match %d spatial axes", n.name.c_str(), sp);
  FConvGeom g = {};
  g.N = (int)X.shape[0]; g.C = (int)X.shape[1]; g.O = g.C; g.groups = 1;
  int in[3] = {1, 1, 1}, out[3] = {1, 1, 1}, k[3] = {1, 1, 1}, s3[3] = {1, 1, 1}, d3[3] = {1, 1, 1}, p3[3] = {0, 0, 0};
  for (int i = 0; i < sp; ++i) {
    const int j = 3 - sp + i;
    in[j] = (int)X.shape[2 + i]; out[j] = (int)Y.shape[2 + i]; k[j] = (int)ps[i]; s3[j] = (int)st[i]; d3[j] = (int)di[i]; p3[j] = (int)pd[i];
    const int want = (in[j] + (int)pd[i] + (int)pd[sp + i] - d3[j] * (k[j] - 1) - 1) / s3[j] + 1;
    if (want != out[j]) fail("%s: output extent %d on spatial axis %d, expected %d", n.name.c_str(), out[j], i, want);
  }
  g.D = in[0]; g.H = in[1]; g.W = in[2]; g.OD = out[0]; g.OH = out[1]; g.OW = out[2]; g.KD = k[0]; g.KH = k[1]; g.KW = k[2];
  g.sd = s3[0]; g.sh = s3[1]; g.sw = s3[2]; g.dd = d3[0]; g.dh = d3[1]; g.dw = d3[2]; g.pd = p3[0]; g.ph = p3[1]; g.pw = p3[2];
  const int cip = global ? 1 : (int)n.i64("count_include_pad", 0);
  P.need(x_eid, kLogical);
  P.produce(y_eid, kLogical);
  P.rt->hbm_bytes += X.bytes() + Y.bytes();
  P.add_step(n.name + " -> fp32 pooling kernel", [=](RunCtx& c) {
    launch_fpool((const float*)c.ptr(x_eid, kLogical), (float*)c.ptr(y_eid, kLogical), g, is_avg ? 1 : 0, cip, c.s);
  });
}

// Is this node one of the float operators?  (Every operand float32; integer graphs never get here.)
static bool is_float_node(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  if (P.ent[P.g.eid(nid, 0)].dtype != kF32) return false;
  for (auto& in : n.inputs)
    if (P.ent[P.root((int)P.g.eid(in))].dtype != kF32) return false;
  return !n.inputs.empty();
}

static void plan_transpose(Planner& P, int nid) {
  const GNode& n = P.g.nodes[nid];
  const int x_eid = (int)P.g.eid(n.inputs[0]), y_eid = (int)P.g.eid(nid, 0);
  const Entry& X = P.ent[x_eid];
  const int rank = (int)X.shape.size();
  auto axes = n.ints("axes");
  if (rank > 4 || (int)axes.size() != rank) fail("Unsupported op: transpose of rank %d", rank);
  std::vector<long long> istr(rank, 1);
  for (int i = rank - 2; i >= 0; --i) istr[i] = istr[i + 1] * X.shape[i + 1];
  std::array<long long, 4> oshape = {1, 1, 1, 1}, ostr = {0, 0, 0, 0};
  for (int i = 0; i < rank; ++i) {
    int a = (int)axes[i];
    if (a < 0) a += rank;
    oshape[4 - rank + i] = X.shape[a];
    ostr[4 - rank + i] = istr[a];
  }
  P.need(x_eid, kLogical);
  P.produce(y_eid, kLogical);
  const int elem = dtype_size(X.dtype);
  P.rt->hbm_bytes += 2 * X.bytes();
  P.add_step("transpose -> permute kernel", [=](RunCtx& c) {
    launch_permute4(c.ptr(x_eid, kLogical), c.ptr(y_eid, kLogical), elem, oshape.data(), ostr.data(), c.s);
  });
}

// ------------------------------------------------------------------------------------------------
// serialisation helpers: dmlc stream encoding (u64 length prefixes, 3rdparty/dmlc-core serializer.h:128-193)
// ------------------------------------------------------------------------------------------------
static void put_u64(std::vector<uint8_t>& b, uint64_t v) { for (int i = 0; i < 8; ++i) b.push_back((uint8_t)(v >> (8 * i))); }
static void put_i32(std::vector<uint8_t>& b, int32_t v) { for (int i = 0; i < 4; ++i) b.push_back((uint8_t)((uint32_t)v >> (8 * i))); }
static void put_str(std::vector<uint8_t>& b, const std::string& s) { put_u64(b, s.size()); b.insert(b.end(), s.begin(), s.end()); }

static void dl_dtype(int dt, uint8_t* code, uint8_t* bits) {
  switch (dt) {
    case kI8: *code = 0; *bits = 8; break;
    case kU8: *code = 1; *bits = 8; break;
    case kI16: *code = 0; *bits = 16; break;
    case kI32: *code = 0; *bits = 32; break;
    case kI64: *code = 0; *bits = 64; break;
    default: *code = 2; *bits = 32; break;
  }
}


static int dl_code(int dt) { return dt == kU8 ? 1 : (dt == kF32 ? 2 : 0); }

// ------------------------------------------------------------------------------------------------
// chained tcgen05 launches: merge runs of consecutive candidate steps (see kernels_tc_chain.cu)
// ------------------------------------------------------------------------------------------------
static void run_chain(RunCtx& c, int id) {
  tkb_runtime* rt = c.rt;
  Workspace* ws = c.ws;
  const ChainPlan& ch = rt->chains[id];
  if (ws->chain_state.size() != rt->chains.size()) ws->chain_state.assign(rt->chains.size(), {});
  std::vector<uint8_t>& state = ws->chain_state[id];
  if (state.empty()) {
    // First run on this stream: the operands live at fixed offsets of this workspace's arena, so the tensor maps and
    // per-layer parameters are encoded once and reused by every later run.
    std::vector<TcChainLayer> layers(ch.cands.size());
    for (size_t i = 0; i < ch.cands.size(); ++i) {
      const ChainCand& cc = rt->chain_cands[ch.cands[i]];
      TcChainLayer& L = layers[i];
      L.a = cc.ta;
      L.a.a = c.ptr(cc.x_eid, cc.phys);
      L.a.y = c.ptr(cc.y_eid, cc.phys);
      L.a.e.sum = cc.sum_eid >= 0 ? c.ptr(cc.sum_eid, cc.phys) : nullptr;
      L.rows_per_img = cc.rows_per_img;
      L.dep_a = ch.dep_a[i];
      L.dep_sum = ch.dep_sum[i];
    }
    std::vector<int> tgt, items;
    std::vector<uint8_t> st(tc_chain_state_bytes());
    int* flags = ws->chain_flags + ch.flag_off;
    int* tgt_dev = ws->chain_flags + rt->chain_flag_ints + ch.flag_off;
    int* items_dev = ws->chain_flags + 2 * rt->chain_flag_ints + ch.item_off;
    if (tc_chain_build(layers.data(), (int)layers.size(), ch.bn, flags, tgt_dev, &tgt, items_dev, &items, st.data()) != 0)
      fail("internal: could not encode the chained tcgen05 launch %d", id);
    if ((int)tgt.size() != ch.flag_ints || (int)items.size() != ch.item_count) fail("internal: chain %d layout mismatch", id);
    CUDA_OK(cudaMemcpyAsync(tgt_dev, tgt.data(), tgt.size() * sizeof(int), cudaMemcpyHostToDevice, c.s));
    CUDA_OK(cudaMemcpyAsync(items_dev, items.data(), items.size() * sizeof(int), cudaMemcpyHostToDevice, c.s));
    CUDA_OK(cudaStreamSynchronize(c.s));  // tgt / items are locals; one-time cost
    state.swap(st);
  }
  c.tc_begin();
  if (launch_tc_chain(state.data(), ch.bn, c.s) != 0) fail("chained tcgen05 launch failed: %s", cudaGetErrorString(cudaGetLastError()));
  c.tc_end();
}

static void form_chains(tkb_runtime* rt, Planner& P) {
  rt->chains.clear();
  rt->chain_flag_ints = 0;
  rt->chain_item_ints = 0;
  std::vector<Step> out;
  auto nimg_of = [&](int cand) {
    const ChainCand& cc = rt->chain_cands[cand];
    return (cc.ta.M + cc.rows_per_img - 1) / cc.rows_per_img;
  };
  size_t i = 0;
  while (i < rt->steps.size()) {
    if (rt->steps[i].cand < 0) { out.push_back(std::move(rt->steps[i])); ++i; continue; }
    ChainPlan ch;
    ch.bn = rt->chain_cands[rt->steps[i].cand].bn;
    const int nimg0 = nimg_of(rt->steps[i].cand);
    // Which layers share a launch (TKB_CHAIN_POLICY / conv_impl): 1 (default) = only layers that read the same
    // activation (a residual block's 3x3 convolution and its 1x1 projection, the Q / K / V projections): they are
    // independent, fill the SMs together and nothing ever waits; every other layer keeps its single-layer kernel.
    // 3 = every eligible layer runs on the chained kernel, a new launch at every dependency; 2 (conv_impl
    // "tc_chain_all") = dependent layers are merged too (per-image completion counters): measured slower on ResNet-18,
    // where a dependent step is bounded by one tile's latency (operands streamed at the SM's L2 bandwidth + epilogue +
    // publish), not by the launch boundary -- see profiles/r02_notes.md.
    size_t j = i;
    while (j < rt->steps.size() && rt->steps[j].cand >= 0 && (int)(j - i) < TKB_CHAIN_MAX_SEG &&
           rt->chain_cands[rt->steps[j].cand].bn == ch.bn && nimg_of(rt->steps[j].cand) == nimg0) {
      const ChainCand& cj = rt->chain_cands[rt->steps[j].cand];
      bool depends = false;
      for (int m : ch.cands) {
        const ChainCand& pm = rt->chain_cands[m];
        depends = depends || P.root(pm.y_eid) == P.root(cj.x_eid) || (cj.sum_eid >= 0 && P.root(pm.y_eid) == P.root(cj.sum_eid));
      }
      if (depends && rt->chain_policy != 2) break;
      ch.cands.push_back(rt->steps[j].cand);
      ++j;
    }
    std::vector<TcChainLayer> layers(ch.cands.size());
    for (size_t k = 0; k < ch.cands.size(); ++k) {
      const ChainCand& cc = rt->chain_cands[ch.cands[k]];
      int da = -1, ds = -1;
      for (size_t m = 0; m < k; ++m) {
        const ChainCand& pm = rt->chain_cands[ch.cands[m]];
        if (P.root(pm.y_eid) == P.root(cc.x_eid)) da = (int)m;
        if (cc.sum_eid >= 0 && P.root(pm.y_eid) == P.root(cc.sum_eid)) ds = (int)m;
      }
      ch.dep_a.push_back(da);
      ch.dep_sum.push_back(ds);
      layers[k].a = cc.ta;
      layers[k].rows_per_img = cc.rows_per_img;
    }
    ch.flag_ints = tc_chain_flag_ints(layers.data(), (int)layers.size());
    ch.flag_off = rt->chain_flag_ints;
    rt->chain_flag_ints += ch.flag_ints;
    ch.item_count = tc_chain_item_count(layers.data(), (int)layers.size(), ch.bn);
    if (ch.item_count < 0) fail("internal: chained launch with too many tiles");
    ch.item_off = rt->chain_item_ints;
    rt->chain_item_ints += ch.item_count;
    const int id = (int)rt->chains.size();
    Step st;
    st.nid = rt->steps[j - 1].nid;
    st.desc = "chained tcgen05 launch " + std::to_string(id);
    st.fn = [id](RunCtx& c) { run_chain(c, id); };
    P.plan << "  chain " << id << ": " << ch.cands.size() << " layers in one persistent launch (BN=" << ch.bn << "):";
    for (size_t k = 0; k < ch.cands.size(); ++k) {
      const ChainCand& cc = rt->chain_cands[ch.cands[k]];
      P.plan << " [" << k << "] M=" << cc.ta.M << " N=" << cc.ta.N << " K=" << cc.ta.K << " dep=" << ch.dep_a[k] << "/" << ch.dep_sum[k];
    }
    P.plan << "\n";
    rt->chains.push_back(std::move(ch));
    out.push_back(std::move(st));
    i = j;
  }
  rt->steps.swap(out);
}

// ------------------------------------------------------------------------------------------------
// BuildEngine (tachikoma_json_runtime.cc:240-290): walk the nodes, dispatch on the operator name
// ------------------------------------------------------------------------------------------------
static void build_plan(tkb_runtime* rt) {
  Planner P(rt);
  Graph& g = rt->g;
  rt->steps.clear();
  rt->chain_cands.clear();
  // Chained launches are an opt-in experiment (conv_impl "tc_chain_all" / "tc_chain_siblings" / "tc_chain_split", or
  // TKB_CHAIN_POLICY=2/1/3 with "tc"): bit-exact, but on B200 slower than the single-layer kernels for every network
  // of BASELINE.json (profiles/r02_notes.md), so the default plan does not use them.
  rt->chain_policy = rt->conv_impl == "tc_chain_all" ? 2 : (rt->conv_impl == "tc_chain_siblings" ? 1 : (rt->conv_impl == "tc_chain_split" ? 3 : 0));
  if (const char* cpol = getenv("TKB_CHAIN_POLICY")) rt->chain_policy = atoi(cpol);
  if (rt->chain_policy < 0 || rt->chain_policy > 3) rt->chain_policy = 0;
  rt->chain_enabled = rt->chain_policy != 0 && rt->conv_impl != "simt" && rt->conv_impl != "tc_nochain" && !getenv("TKB_NO_CHAIN");
  rt->arena_bytes = 0;
  rt->macs = rt->hbm_bytes = 0;
  rt->tc_macs = 0;
  rt->trace_recs.clear();
  for (auto& e : rt->ent) { e.off[0] = e.off[1] = -1; e.alias_of = -1; }
  for (uint32_t eid : g.input_eids) rt->ent[eid].off[kLogical] = P.arena_alloc(rt->ent[eid].bytes());  // H2D staging
  P.plan << "plan for " << rt->symbol << " (conv_impl=" << rt->conv_impl << ")\n";
  static const std::regex conv_pat(".*conv2d.*"), dense_pat(".*dense.*");
  std::set<int> fused_nodes;
  for (size_t nid = 0; nid < g.nodes.size(); ++nid) {
    const GNode& n = g.nodes[nid];
    if (n.op_type != "kernel") continue;
    P.cur_nid = (int)nid;
    const std::string& op = n.name;
    if (n.num_outputs != 1 && op != "nn.batch_norm") fail("Unsupported op: %s with %u outputs", op.c_str(), n.num_outputs);
    // float32 half of the op list, dispatched by name exactly as tachikoma_json_runtime.cc:255-285 does
    static const std::set<std::string> float_unary = {"abs", "exp", "log", "sqrt", "round", "nn.relu", "nn.leaky_relu", "tanh", "sigmoid"};  // float clip: plan_unary
    auto plan_float = [&]() -> bool {
      if (name_has(op, "qnn.") || !is_float_node(P, (int)nid)) return false;
      const bool deconv = name_has(op, "deconv1d") || name_has(op, "deconv2d") || name_has(op, "deconv3d") ||
                          name_has(op, "conv1d_transpose") || name_has(op, "conv2d_transpose") || name_has(op, "conv3d_transpose");
      const bool conv = name_has(op, "conv1d") || name_has(op, "conv2d") || name_has(op, "conv3d");
      if (deconv) plan_float_conv(P, (int)nid, true);
      else if (conv) plan_float_conv(P, (int)nid, false);
      else if (name_has(op, "dense")) plan_float_dense(P, (int)nid);
      else if (op == "nn.batch_norm") plan_float_batch_norm(P, (int)nid);
      else if (op == "nn.max_pool2d" || op == "nn.max_pool3d" || op == "nn.avg_pool2d" || op == "nn.avg_pool3d" || op == "nn.global_avg_pool2d") plan_float_pool(P, (int)nid);
      else if (float_unary.count(op)) plan_float_unary(P, (int)nid);
      else if (op == "nn.softmax") plan_float_softmax(P, (int)nid);
      else if (op == "add" || op == "multiply") plan_float_binary(P, (int)nid);
      else if (op == "nn.layer_norm") plan_float_layer_norm(P, (int)nid);
      else if (op == "nn.batch_matmul") plan_float_batch_matmul(P, (int)nid);
      else return false;
      return true;
    };
    std::vector<int> fused_tail;
    if (fused_nodes.count((int)nid)) {}  // its work is part of an earlier node's fused step (plan_upcast_gap)
    else if (op == "cast" && !(fused_tail = plan_upcast_gap(P, (int)nid)).empty()) fused_nodes.insert(fused_tail.begin(), fused_tail.end());
    else if (plan_float()) {}
    else if (op == "tachikoma.qnn.conv2d") plan_conv(P, (int)nid, true);
    else if (op == "tachikoma.qnn.dense") plan_dense(P, (int)nid, true);
    else if (op == "qnn.conv2d") plan_conv(P, (int)nid, false);
    else if (op == "qnn.dense") plan_dense(P, (int)nid, false);
    else if (op == "qnn.batch_matmul") plan_batch_matmul(P, (int)nid);
    else if (op == "qnn.requantize") plan_requantize(P, (int)nid);
    else if (op == "qnn.add" || op == "qnn.subtract") plan_qnn_add(P, (int)nid);
    else if (op == "qnn.mul") plan_qnn_mul(P, (int)nid);
    else if (op == "qnn.quantize") plan_quantize(P, (int)nid, false);
    else if (op == "qnn.dequantize") plan_quantize(P, (int)nid, true);
    else if (op == "qnn.concatenate") plan_qnn_concatenate(P, (int)nid);
    else if (op == "clip" || op == "cast") plan_unary(P, (int)nid);
    else if (op == "nn.bias_add" || op == "add") plan_add(P, (int)nid);
    else if (op == "nn.max_pool2d" || op == "nn.avg_pool2d" || op == "nn.global_avg_pool2d") plan_pool(P, (int)nid);
    else if (op == "nn.pad") plan_pad(P, (int)nid);
    else if (op == "reshape" || op == "nn.batch_flatten") plan_reshape(P, (int)nid);
    else if (op == "transpose") plan_transpose(P, (int)nid);
    else if (op == "reinterpret") plan_reinterpret(P, (int)nid);
    else if (op == "take") plan_take(P, (int)nid);
    else if (op == "qnn.leaky_relu") plan_qnn_leaky_relu(P, (int)nid);
    else fail("Unsupported op: %s", op.c_str());  // same message as tachikoma_json_runtime.cc:286
    // trace records of this node: inputs then outputs, in the logical layout
    auto add_rec = [&](const std::string& name, int eid) {
      TraceRec tr;
      tr.name = name; tr.eid = eid; tr.nid = (int)nid;
      const int r = P.root(eid);
      tr.logical = rt->ent[r].is_const || P.has(r, kLogical);  // as of this node: later conversions do not count
      rt->trace_recs.push_back(tr);
    };
    for (size_t i = 0; i < n.inputs.size(); ++i)
      add_rec(n.name + "____topo-index:" + std::to_string(nid) + "____input-num:" + std::to_string(i), (int)g.eid(n.inputs[i]));
    add_rec(n.name + "____topo-index:" + std::to_string(nid) + "____output-num:0", (int)g.eid((uint32_t)nid, 0));
  }
  // graph outputs are delivered in the logical layout
  P.cur_nid = -1;
  for (auto& h : g.heads) P.need((int)g.eid(h), kLogical);
  form_chains(rt, P);
  // trace staging plan: tensors that exist in logical layout are copied in place; NHWC-physical ones
  // are permuted into a device ring first
  // Every tensor is staged ONCE per run: an activation appears as the output of its producer and again as
  // an input of each consumer; the later records alias the first one's host slot (the serializer reads the
  // same bytes), so neither the D2D permute nor the D2H copy is repeated.
  // The pinned host buffer of a traced run IS the serialized NDArray-list blob (src/runtime/file_utils.cc:210-229,
  // include/tvm/runtime/ndarray.h:449-494): names and per-array headers are written once when the buffer is
  // allocated, the device -> host copies land directly at the payload offsets, and tkb_trace_blob hands the buffer out
  // without another pass over the data (repeated records are filled by a parallel host copy on request).
  rt->trace_host_off.assign(rt->trace_recs.size(), 0);
  rt->trace_ring_off.assign(rt->trace_recs.size(), -1);
  rt->trace_dup_of.assign(rt->trace_recs.size(), -1);
  rt->trace_d2h_bytes = 0;
  rt->blob_headers.clear();
  long long hoff = 0, roff = 0;
  {
    std::vector<uint8_t> pre;
    put_u64(pre, 0xF7E58D4F05049CB7ull);
    put_u64(pre, 0);
    put_u64(pre, rt->trace_recs.size());
    for (auto& r : rt->trace_recs) put_str(pre, r.name);
    put_u64(pre, rt->trace_recs.size());
    hoff = (long long)pre.size();
    rt->blob_headers.emplace_back(0, std::move(pre));
  }
  std::unordered_map<int, int> first_of_root;
  for (size_t i = 0; i < rt->trace_recs.size(); ++i) {
    const Entry& re = rt->ent[rt->trace_recs[i].eid];  // shape / dtype as this record sees the tensor (reshape aliases)
    std::vector<uint8_t> hd;
    put_u64(hd, 0xDD5E40F096B4A13Full);
    put_u64(hd, 0);
    put_i32(hd, 1);  // kDLCPU
    put_i32(hd, 0);
    put_i32(hd, (int32_t)re.shape.size());
    uint8_t code, bits;
    dl_dtype(re.dtype, &code, &bits);
    hd.push_back(code); hd.push_back(bits); hd.push_back(1); hd.push_back(0);  // lanes = 1 (u16 LE)
    for (auto d : re.shape) put_u64(hd, (uint64_t)d);
    put_u64(hd, (uint64_t)re.bytes());
    const long long hlen = (long long)hd.size();
    rt->blob_headers.emplace_back(hoff, std::move(hd));
    hoff += hlen;
    rt->trace_host_off[i] = hoff;
    hoff += re.bytes();
    const int r = P.root(rt->trace_recs[i].eid);
    const Entry& e = rt->ent[r];
    auto it = first_of_root.find(r);
    if (it != first_of_root.end()) {
      rt->trace_dup_of[i] = it->second;
      continue;
    }
    first_of_root[r] = (int)i;
    if (!e.is_const) rt->trace_d2h_bytes += e.bytes();
    if (!e.is_const && !rt->trace_recs[i].logical) {
      rt->trace_ring_off[i] = roff;
      roff += (e.bytes() + 255) & ~255LL;
    }
  }
  rt->trace_host_bytes = hoff;
  rt->trace_ring_bytes = roff;
  P.plan << "arena " << rt->arena_bytes << " B, int8 MACs " << rt->macs << ", algorithmic HBM bytes " << rt->hbm_bytes
         << ", serialized trace " << hoff << " B (" << rt->trace_d2h_bytes << " B of distinct activations)\n";
  rt->plan_text = P.plan.str();
}

// ------------------------------------------------------------------------------------------------
// run
// ------------------------------------------------------------------------------------------------
static Workspace* get_workspace(tkb_runtime* rt, cudaStream_t s) {
  std::lock_guard<std::mutex> lk(rt->ws_mu);
  auto& slot = rt->ws[s];
  if (!slot) {
    slot.reset(new Workspace());
    slot->arena_bytes = rt->arena_bytes + 256;
    CUDA_OK(cudaMalloc((void**)&slot->arena, slot->arena_bytes));
  }
  return slot.get();
}

static void ensure_trace_stream(Workspace* ws) {
  if (ws->side) return;
  // The trace stream gets the highest priority: its transpose blocks are then placed ahead of the next tensor-core
  // kernel's (programmatically launched) CTAs whenever SMs free up instead of starving behind them
  // (traced ResNet-18: 144.7k -> 149.1k img/s; TKB_SIDE_PRIO=0 restores the default priority)
  int prio_lo = 0, prio_hi = 0;
  const char* sp = getenv("TKB_SIDE_PRIO");
  if (!(sp && atoi(sp) == 0) && cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi) == cudaSuccess)
    CUDA_OK(cudaStreamCreateWithPriority(&ws->side, cudaStreamNonBlocking, prio_hi));
  else
    CUDA_OK(cudaStreamCreateWithFlags(&ws->side, cudaStreamNonBlocking));
  CUDA_OK(cudaEventCreateWithFlags(&ws->ev, cudaEventDisableTiming));
  CUDA_OK(cudaEventCreateWithFlags(&ws->ev_join, cudaEventDisableTiming));
}

// Pinned host image of the blob: headers and constant payloads are written here once per workspace.
static void ensure_trace_host(tkb_runtime* rt, Workspace* ws) {
  if (ws->trace_host && ws->trace_host_bytes >= rt->trace_host_bytes) return;
  if (ws->trace_host) cudaFreeHost(ws->trace_host);
  ws->trace_host = nullptr;
  CUDA_OK(cudaMallocHost((void**)&ws->trace_host, (size_t)rt->trace_host_bytes + 256));
  ws->trace_host_bytes = rt->trace_host_bytes;
  for (auto& h : rt->blob_headers) memcpy(ws->trace_host + h.first, h.second.data(), h.second.size());
  for (size_t i = 0; i < rt->trace_recs.size(); ++i) {  // constants never change: every record of one is final
    int r = rt->trace_recs[i].eid;
    while (rt->ent[r].alias_of >= 0) r = rt->ent[r].alias_of;
    const Entry& e = rt->ent[r];
    if (e.is_const) memcpy(ws->trace_host + rt->trace_host_off[i], e.host.data(), (size_t)rt->ent[rt->trace_recs[i].eid].bytes());
  }
}

static void ensure_trace_buffers(tkb_runtime* rt, Workspace* ws, int mode) {
  ensure_trace_stream(ws);
  if (ws->trace_ring_bytes < rt->trace_ring_bytes || !ws->trace_ring) {
    if (ws->trace_ring) cudaFree(ws->trace_ring);
    ws->trace_ring = nullptr;
    CUDA_OK(cudaMalloc((void**)&ws->trace_ring, rt->trace_ring_bytes + 256));
    ws->trace_ring_bytes = rt->trace_ring_bytes;
  }
  if (mode >= 2) ensure_trace_host(rt, ws);
}

// Device address a trace record is read from after the run: the ring for NHWC-physical tensors, else the tensor itself.
static const void* trace_src(tkb_runtime* rt, Workspace* ws, const std::vector<void*>& in_ptr, size_t i) {
  if (rt->trace_ring_off[i] >= 0) return ws->trace_ring + rt->trace_ring_off[i];
  int r = rt->trace_recs[i].eid;
  while (rt->ent[r].alias_of >= 0) r = rt->ent[r].alias_of;
  const Entry& e = rt->ent[r];
  if (e.is_input) return in_ptr[e.input_index];
  if (e.off[kLogical] < 0) fail("internal: trace record %d has no logical copy", (int)i);
  return ws->arena + e.off[kLogical];
}

// Stage the trace records of node `nid` (records [lo, hi)): D2D permute for NHWC-physical tensors,
// then (mode 2) D2H on the side stream, ordered after the producing kernels by an event.
static void trace_node(RunCtx& c, size_t lo, size_t hi) {
  tkb_runtime* rt = c.rt;
  // All trace work runs on the side stream, ordered after the producing kernels by an event: the layout
  // transposes are HBM-bound and overlap with the following tensor-core kernels instead of delaying them.
  bool any = false;
  for (size_t i = lo; i < hi && !any; ++i) {
    if (rt->trace_dup_of[i] >= 0) continue;
    int r = rt->trace_recs[i].eid;
    while (rt->ent[r].alias_of >= 0) r = rt->ent[r].alias_of;
    any = rt->trace_ring_off[i] >= 0 || (c.mode >= 2 && !rt->ent[r].is_const);
  }
  if (!any) return;
  cudaStream_t side = c.ws->side;
  c.tc_flush();
  CUDA_OK(cudaEventRecord(c.ws->ev, c.s));
  CUDA_OK(cudaStreamWaitEvent(side, c.ws->ev, 0));
  c.ws->side_dirty = true;
  for (size_t i = lo; i < hi; ++i) {
    if (rt->trace_ring_off[i] < 0 || rt->trace_dup_of[i] >= 0) continue;
    const int eid = rt->trace_recs[i].eid;
    int r = eid;
    while (rt->ent[r].alias_of >= 0) r = rt->ent[r].alias_of;
    const Entry& e = rt->ent[r];
    const long long N = e.shape[0], C = e.shape[1], H = e.shape[2], W = e.shape[3];
    long long oshape[4] = {N, C, H, W}, istr[4] = {H * W * C, 1, W * C, C};
    void* dstp = c.ws->trace_ring + rt->trace_ring_off[i];
    if (dtype_size(e.dtype) == 1 && launch_nhwc_to_nchw8(c.ptr(r, kNHWC), dstp, N, H * W, C, side)) continue;
    if (launch_transpose_tile(c.ptr(r, kNHWC), dstp, dtype_size(e.dtype), N, H * W, C, side)) continue;
    launch_permute4(c.ptr(r, kNHWC), c.ws->trace_ring + rt->trace_ring_off[i], dtype_size(e.dtype), oshape, istr, side);
  }
  if (c.mode < 2) return;
  for (size_t i = lo; i < hi; ++i) {
    const int eid = rt->trace_recs[i].eid;
    int r = eid;
    while (rt->ent[r].alias_of >= 0) r = rt->ent[r].alias_of;
    const Entry& e = rt->ent[r];
    uint8_t* dst = c.ws->trace_host + rt->trace_host_off[i];
    const long long bytes = rt->ent[eid].bytes();
    if (e.is_const || rt->trace_dup_of[i] >= 0) continue;  // constants were written once; duplicates alias a slot
    CUDA_OK(cudaMemcpyAsync(dst, trace_src(rt, c.ws, c.in_ptr, i), bytes, cudaMemcpyDeviceToHost, side));
  }
}

// Join the side stream back into the caller's stream: everything enqueued on `s` after a traced run (the next
// run's kernels, which overwrite the arena and the ring; the caller's own work) is ordered after the trace work.
static void trace_join(RunCtx& c) {
  if (!c.ws->side || !c.ws->side_dirty) return;
  CUDA_OK(cudaEventRecord(c.ws->ev_join, c.ws->side));
  CUDA_OK(cudaStreamWaitEvent(c.s, c.ws->ev_join, 0));
  c.ws->side_dirty = false;
}

static void check_arg(const Entry& e, const tkb_DLTensor* t, int i) {
  if (!t || !t->data) fail("argument %d is null", i);
  if (t->strides != nullptr) fail("argument %d: only compact tensors are supported (strides must be NULL)", i);
  if (t->ndim != (int)e.shape.size()) fail("argument %d: rank %d, expected %d", i, t->ndim, (int)e.shape.size());
  for (int d = 0; d < t->ndim; ++d)
    if (t->shape[d] != e.shape[d]) fail("argument %d: dim %d is %lld, expected %lld", i, d, (long long)t->shape[d], (long long)e.shape[d]);
  if (t->dtype.bits / 8 != dtype_size(e.dtype) || t->dtype.lanes != 1) fail("argument %d: dtype width mismatch", i);
  // the reference's NDArray copy checks the full DLDataType (code 0 int, 1 uint, 2 float), not only the width
  if (t->dtype.code != dl_code(e.dtype))
    fail("argument %d: dtype code %d (0 int, 1 uint, 2 float), expected %d", i, (int)t->dtype.code, dl_code(e.dtype));
  if (t->device.device_type != tkbDLCPU && t->device.device_type != tkbDLCUDA && t->device.device_type != tkbDLCUDAHost)
    fail("argument %d: unsupported device type %d", i, t->device.device_type);
}

static void run_impl(tkb_runtime* rt, const tkb_DLTensor* const* args, int n_args, cudaStream_t s, int mode) {
  if (!rt->initialized) fail("The module has not been initialized");
  if (mode < 0 || mode > 2) fail("trace mode must be 0, 1 or 2");
  Graph& g = rt->g;
  const int n_in = (int)g.input_eids.size(), n_out = (int)g.heads.size();
  if (n_args != n_in + n_out) fail("Found mismatch in the number of provided data entries and required.");
  RunCtx c;
  c.rt = rt;
  c.s = s;
  c.ws = get_workspace(rt, s);
  c.mode = mode;
  c.time_tc = rt->time_tc.load();
  c.in_ptr.resize(n_in);
  c.ws->tc_used = 0;
  c.ws->tc_launches = 0;
  c.ws->trace_mode = 0;  // until this run has been enqueued completely
  c.ws->payload_on_host = c.ws->dups_filled = false;

  const long long before = g_launches;
  // inputs: device tensors are used in place, host tensors are staged (H2D on the caller's stream)
  for (int i = 0; i < n_in; ++i) {
    Entry& e = rt->ent[g.input_eids[i]];
    check_arg(e, args[i], i);
    uint8_t* p = (uint8_t*)args[i]->data + args[i]->byte_offset;
    if (args[i]->device.device_type == tkbDLCUDA) {
      c.in_ptr[i] = p;
    } else {
      void* stage = c.ws->arena + rt->ent[g.input_eids[i]].off[kLogical];
      CUDA_OK(cudaMemcpyAsync(stage, p, e.bytes(), cudaMemcpyHostToDevice, s));
      c.in_ptr[i] = stage;
    }
  }
  if (rt->chain_flag_ints > 0) {  // completion counters of the chained launches: one memset ahead of the first kernel
    if (!c.ws->chain_flags)  // [counters | targets | work lists]
      CUDA_OK(cudaMalloc((void**)&c.ws->chain_flags, ((size_t)rt->chain_flag_ints * 2 + rt->chain_item_ints) * sizeof(int)));
    CUDA_OK(cudaMemsetAsync(c.ws->chain_flags, 0, (size_t)rt->chain_flag_ints * sizeof(int), s));
  }
  if (mode > 0) ensure_trace_buffers(rt, c.ws, mode);
  size_t rec = 0;
  for (size_t si = 0; si < rt->steps.size(); ++si) {
    if (!rt->steps[si].is_tc) c.tc_flush();
    rt->steps[si].fn(c);
    const int nid = rt->steps[si].nid;
    const bool last_of_node = si + 1 == rt->steps.size() || rt->steps[si + 1].nid != nid;
    if (mode > 0 && nid >= 0 && last_of_node) {
      // records are grouped per node in node order; nodes without steps (reshape aliases) in between
      size_t hi = rec;
      while (hi < rt->trace_recs.size() && rt->trace_recs[hi].nid <= nid) ++hi;
      trace_node(c, rec, hi);
      rec = hi;
    }
  }
  if (mode > 0 && rec < rt->trace_recs.size()) trace_node(c, rec, rt->trace_recs.size());
  c.tc_flush();
  bool host_out = false;
  for (int j = 0; j < n_out; ++j) {
    const int eid = (int)g.eid(g.heads[j]);
    const Entry& e = rt->ent[eid];
    check_arg(e, args[n_in + j], n_in + j);
    uint8_t* p = (uint8_t*)args[n_in + j]->data + args[n_in + j]->byte_offset;
    const bool dev = args[n_in + j]->device.device_type == tkbDLCUDA;
    CUDA_OK(cudaMemcpyAsync(p, c.ptr(eid, kLogical), e.bytes(), dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
    host_out = host_out || !dev;
  }
  CUDA_OK(cudaGetLastError());
  // Host results are valid on return, like the reference's.  The trace is NOT waited for here: the side stream keeps
  // copying after the call has returned (tkb_trace_sync / tkb_trace_blob wait for it), so a caller that alternates two
  // streams overlaps one batch's trace transfer with the next batch's upload and compute.  The join comes after the
  // output copies: everything enqueued on `s` later (the next run, which overwrites the arena and the ring) is still
  // ordered after the trace work.
  const bool side_busy = c.ws->side && c.ws->side_dirty;
  if (host_out && side_busy) {
    CUDA_OK(cudaEventRecord(c.ws->ev, s));
    trace_join(c);
    CUDA_OK(cudaEventSynchronize(c.ws->ev));
  } else {
    trace_join(c);
    if (host_out) CUDA_OK(cudaStreamSynchronize(s));
  }
  c.ws->trace_mode = mode;
  c.ws->payload_on_host = mode >= 2;
  c.ws->in_ptr = c.in_ptr;
  rt->launches += g_launches - before;
  std::lock_guard<std::mutex> lk(rt->ws_mu);
  rt->last_stream = s;
}

// Copy `bytes` from src to dst with up to `nthreads` host threads (the repeated records of a ResNet-18 batch-64 trace
// are ~250 MB: one thread moves that in ~25 ms, eight in ~5).
static void parallel_copy(const std::vector<std::array<long long, 3>>& jobs, uint8_t* base, int nthreads) {
  const long long chunk = 4ll << 20;
  std::vector<std::array<long long, 3>> parts;
  for (auto& j : jobs)
    for (long long o = 0; o < j[2]; o += chunk) parts.push_back({j[0] + o, j[1] + o, std::min(chunk, j[2] - o)});
  if (parts.empty()) return;
  nthreads = std::max(1, std::min<int>(nthreads, (int)parts.size()));
  std::atomic<size_t> next{0};
  auto work = [&] {
    for (size_t i = next.fetch_add(1); i < parts.size(); i = next.fetch_add(1))
      memcpy(base + parts[i][0], base + parts[i][1], (size_t)parts[i][2]);
  };
  std::vector<std::thread> th;
  for (int t = 1; t < nthreads; ++t) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();
}

// Make ws->trace_host the complete serialized blob of the last run on this workspace.  Mode 2 runs already streamed
// every distinct activation to its payload offset; after a mode 1 run (device-resident capture) the read-out happens
// here, from the ring / arena, which stay valid until the next run on the same stream.
static void finish_trace_blob(tkb_runtime* rt, Workspace* ws) {
  if (ws->trace_mode < 1) fail("no trace captured: run with trace mode 1 or 2 (tkb_trace_enable / tkb_run_traced) first");
  ensure_trace_stream(ws);
  if (!ws->payload_on_host) {
    ensure_trace_host(rt, ws);
    for (size_t i = 0; i < rt->trace_recs.size(); ++i) {
      if (rt->trace_dup_of[i] >= 0) continue;
      int r = rt->trace_recs[i].eid;
      while (rt->ent[r].alias_of >= 0) r = rt->ent[r].alias_of;
      if (rt->ent[r].is_const) continue;
      CUDA_OK(cudaMemcpyAsync(ws->trace_host + rt->trace_host_off[i], trace_src(rt, ws, ws->in_ptr, i),
                              (size_t)rt->ent[rt->trace_recs[i].eid].bytes(), cudaMemcpyDeviceToHost, ws->side));
    }
    ws->payload_on_host = true;
    ws->dups_filled = false;
  }
  CUDA_OK(cudaStreamSynchronize(ws->side));
  if (!ws->dups_filled) {
    std::vector<std::array<long long, 3>> jobs;
    for (size_t i = 0; i < rt->trace_recs.size(); ++i) {
      const int d = rt->trace_dup_of[i];
      if (d < 0) continue;
      int r = rt->trace_recs[i].eid;
      while (rt->ent[r].alias_of >= 0) r = rt->ent[r].alias_of;
      if (rt->ent[r].is_const) continue;  // written when the buffer was allocated
      jobs.push_back({rt->trace_host_off[i], rt->trace_host_off[d], rt->ent[rt->trace_recs[i].eid].bytes()});
    }
    int nt = (int)std::thread::hardware_concurrency();
    if (const char* e = getenv("TKB_BLOB_THREADS")) nt = atoi(e);
    parallel_copy(jobs, ws->trace_host, std::max(1, std::min(nt, 16)));
    ws->dups_filled = true;
  }
}

// ------------------------------------------------------------------------------------------------
// serialisation: dmlc stream encoding (u64 length prefixes, 3rdparty/dmlc-core serializer.h:128-193)
// ------------------------------------------------------------------------------------------------
struct Reader {
  const uint8_t* p;
  size_t n, pos = 0;
  uint64_t u64() {
    if (pos + 8 > n) fail("binary blob truncated");
    uint64_t v = 0;
    for (int i = 0; i < 8; ++i) v |= (uint64_t)p[pos + i] << (8 * i);
    pos += 8;
    return v;
  }
  std::string str() {
    uint64_t len = u64();
    if (pos + len > n) fail("binary blob truncated");
    std::string s((const char*)p + pos, len);
    pos += len;
    return s;
  }
};

static tkb_runtime* create_impl(const std::string& symbol, const std::string& json, const std::vector<std::string>& names) {
  std::unique_ptr<tkb_runtime> rt(new tkb_runtime());
  rt->symbol = symbol;
  rt->graph_json = json;
  rt->const_names = names;
  for (auto& s : rt->const_names) rt->const_name_ptrs.push_back(s.c_str());
  rt->g = parse_graph(json, names);
  rt->ent.resize(rt->g.num_entries());
  for (size_t nid = 0; nid < rt->g.nodes.size(); ++nid) {
    const GNode& n = rt->g.nodes[nid];
    if (n.shapes.size() != rt->g.node_row_ptr[nid + 1] - rt->g.node_row_ptr[nid])
      fail("node %s: %d shapes for %d entries", n.name.c_str(), (int)n.shapes.size(),
           (int)(rt->g.node_row_ptr[nid + 1] - rt->g.node_row_ptr[nid]));
    for (size_t j = 0; j < n.shapes.size(); ++j) {
      Entry& e = rt->ent[rt->g.eid((uint32_t)nid, (uint32_t)j)];
      e.shape = n.shapes[j];
      e.dtype = dtype_from_string(n.dtypes[j]);
      e.is_const = n.op_type == "const";
      e.is_input = n.op_type == "input";
    }
  }
  for (size_t i = 0; i < rt->g.input_eids.size(); ++i) rt->ent[rt->g.input_eids[i]].input_index = (int)i;
  return rt.release();
}

}  // namespace tkb

// ================================================================================================
// C ABI
// ================================================================================================
using namespace tkb;

#define TKB_TRY try {
#define TKB_CATCH                                   \
  }                                                 \
  catch (const std::exception& ex) {                \
    g_last_error = ex.what();                       \
    return -1;                                      \
  }                                                 \
  return 0;

extern "C" {

const char* tkb_last_error(void) { return g_last_error.c_str(); }
const char* tkb_type_key(void) { return "tachikoma_json"; }
const char* tkb_version(void) { return "tkb200 0.1 sm_100a tcgen05.kind::i8+TMA"; }

int tkb_create(const char* symbol, const char* graph_json, const char* const* const_names, int n_consts, tkb_handle* out) {
  TKB_TRY
  if (!symbol || !graph_json || !out) fail("tkb_create: null argument");
  std::vector<std::string> names;
  for (int i = 0; i < n_consts; ++i) names.emplace_back(const_names[i]);
  *out = create_impl(symbol, graph_json, names);
  TKB_CATCH
}

static void destroy_workspaces(tkb_handle h) {
  for (auto& kv : h->ws) {
    Workspace* w = kv.second.get();
    if (w->side) cudaStreamSynchronize(w->side);
    if (w->arena) cudaFree(w->arena);
    if (w->trace_ring) cudaFree(w->trace_ring);
    if (w->trace_host) cudaFreeHost(w->trace_host);
    if (w->chain_flags) cudaFree(w->chain_flags);
    if (w->side) cudaStreamDestroy(w->side);
    if (w->ev) cudaEventDestroy(w->ev);
    if (w->ev_join) cudaEventDestroy(w->ev_join);
    for (cudaEvent_t e : w->tc_events) cudaEventDestroy(e);
  }
  h->ws.clear();
}

static void free_plan_allocs(tkb_handle h) {
  for (void* p : h->dev_allocs) cudaFree(p);
  h->dev_allocs.clear();
  h->host_tabs.clear();
  for (auto& e : h->ent) e.dev_const = nullptr;  // uploaded constants were part of dev_allocs
}

int tkb_free(tkb_handle h) {
  TKB_TRY
  if (!h) return 0;
  destroy_workspaces(h);
  free_plan_allocs(h);
  delete h;
  TKB_CATCH
}

int tkb_save_binary(tkb_handle h, const void** blob, size_t* len) {
  TKB_TRY
  h->blob.clear();
  put_str(h->blob, h->symbol);
  put_str(h->blob, h->graph_json);
  put_u64(h->blob, h->const_names.size());
  for (auto& s : h->const_names) put_str(h->blob, s);
  *blob = h->blob.data();
  *len = h->blob.size();
  TKB_CATCH
}

int tkb_load_binary(const void* blob, size_t len, tkb_handle* out) {
  TKB_TRY
  Reader r{(const uint8_t*)blob, len};
  std::string symbol = r.str(), json = r.str();
  uint64_t n = r.u64();
  std::vector<std::string> names;
  for (uint64_t i = 0; i < n; ++i) names.push_back(r.str());
  *out = create_impl(symbol, json, names);
  TKB_CATCH
}

const char* tkb_get_symbol(tkb_handle h) { return h->symbol.c_str(); }

int tkb_get_const_vars(tkb_handle h, const char* const** names, int* n) {
  *names = h->const_name_ptrs.data();
  *n = (int)h->const_name_ptrs.size();
  return 0;
}

int tkb_num_inputs(tkb_handle h) { return (int)h->g.input_eids.size(); }
int tkb_num_outputs(tkb_handle h) { return (int)h->g.heads.size(); }

int tkb_init(tkb_handle h, const tkb_DLTensor* const* consts, int n_consts) {
  TKB_TRY
  std::lock_guard<std::mutex> guard(h->init_mu);
  if (h->initialized) return 0;
  if (n_consts != (int)h->g.const_nids.size()) fail("The number of input constants must match the number of required.");
  for (int i = 0; i < n_consts; ++i) {
    Entry& e = h->ent[h->g.eid(h->g.const_nids[i], 0)];
    check_arg(e, consts[i], i);
    e.host.resize((size_t)e.bytes());
    const uint8_t* p = (const uint8_t*)consts[i]->data + consts[i]->byte_offset;
    if (consts[i]->device.device_type == tkbDLCUDA) CUDA_OK(cudaMemcpy(e.host.data(), p, e.host.size(), cudaMemcpyDeviceToHost));
    else memcpy(e.host.data(), p, e.host.size());
  }
  build_plan(h);
  h->initialized = true;
  TKB_CATCH
}

int tkb_run(tkb_handle h, const tkb_DLTensor* const* args, int n_args, void* cuda_stream) {
  TKB_TRY
  run_impl(h, args, n_args, (cudaStream_t)cuda_stream, h->trace_mode.load());
  TKB_CATCH
}

int tkb_run_traced(tkb_handle h, const tkb_DLTensor* const* args, int n_args, void* cuda_stream, int trace_mode) {
  TKB_TRY
  run_impl(h, args, n_args, (cudaStream_t)cuda_stream, trace_mode);
  TKB_CATCH
}

int tkb_packed_run(void* args, int* type_codes, int num_args, void* out_ret_value, int* out_ret_tcode, void* resource_handle) {
  TKB_TRY
  (void)out_ret_value; (void)out_ret_tcode; (void)resource_handle;
  union Value { int64_t v_int64; double v_float64; void* v_handle; const char* v_str; };
  Value* a = (Value*)args;
  if (num_args < 1 || type_codes[0] != 3) fail("tkb_packed_run: args[0] must be the module handle");
  std::vector<const tkb_DLTensor*> ts;
  for (int i = 1; i < num_args; ++i) {
    if (type_codes[i] != 7 && type_codes[i] != 13) fail("Expect NDArray or DLTensor as inputs");
    ts.push_back((const tkb_DLTensor*)a[i].v_handle);  // an NDArray handle points at its DLTensor
  }
  tkb_handle h = (tkb_handle)a[0].v_handle;
  if (!h) fail("tkb_packed_run: null module handle");
  run_impl(h, ts.data(), (int)ts.size(), nullptr, h->trace_mode.load());
  TKB_CATCH
}

int tkb_set_option(tkb_handle h, const char* key, const char* value) {
  TKB_TRY
  const std::string k = key, v = value;
  if (k == "conv_impl") {
    if (v != "auto" && v != "simt" && v != "tc" && v != "tc_explicit" && v != "tc_im2col" && v != "tc_nochain" && v != "tc_chain_all" &&
        v != "tc_chain_siblings" && v != "tc_chain_split" && v != "tc_pair")
      fail("conv_impl must be auto|simt|tc|tc_nochain|tc_chain_all|tc_chain_siblings|tc_chain_split|tc_explicit|tc_im2col|tc_pair");
    if (h->initialized && v != h->conv_impl) {
      // Re-planning is a configuration step like tkb_init: no run may be in flight.  The arena layout and every
      // uploaded / repacked constant belong to the old plan: drop the workspaces (streams and events included) and
      // the plan's device allocations before building the new one.
      h->conv_impl = v;
      std::lock_guard<std::mutex> guard(h->init_mu);
      std::lock_guard<std::mutex> lk(h->ws_mu);
      CUDA_OK(cudaDeviceSynchronize());
      destroy_workspaces(h);
      free_plan_allocs(h);
      build_plan(h);
    }
    h->conv_impl = v;
  } else if (k == "time_tc") {
    h->time_tc = v == "1";
  } else {
    fail("unknown option %s", key);
  }
  TKB_CATCH
}

int64_t tkb_launch_count(tkb_handle h) { return h->launches.load(); }
const char* tkb_plan_summary(tkb_handle h) { return h->plan_text.c_str(); }
int tkb_work(tkb_handle h, int64_t* macs, int64_t* hbm_bytes) {
  if (macs) *macs = h->macs;
  if (hbm_bytes) *hbm_bytes = h->hbm_bytes;
  return 0;
}

int64_t tkb_tc_macs(tkb_handle h) { return h ? h->tc_macs : 0; }

static Workspace* stream_ws(tkb_handle h, cudaStream_t s) {
  std::lock_guard<std::mutex> lk(h->ws_mu);
  auto it = h->ws.find(s);
  if (it == h->ws.end()) fail("no run has been executed on this stream yet");
  return it->second.get();
}
static Workspace* last_ws(tkb_handle h) {
  cudaStream_t s;
  {
    std::lock_guard<std::mutex> lk(h->ws_mu);
    s = h->last_stream;
  }
  return stream_ws(h, s);
}

int tkb_tc_time(tkb_handle h, double* ms, int* launches) {
  TKB_TRY
  Workspace* ws = last_ws(h);
  double total = 0;
  for (size_t i = 0; i + 1 < ws->tc_used; i += 2) {
    CUDA_OK(cudaEventSynchronize(ws->tc_events[i + 1]));
    float t = 0;
    CUDA_OK(cudaEventElapsedTime(&t, ws->tc_events[i], ws->tc_events[i + 1]));
    total += t;
  }
  *ms = total;
  *launches = (int)ws->tc_launches;
  TKB_CATCH
}

int tkb_trace_enable(tkb_handle h, int mode) {
  TKB_TRY
  if (mode < 0 || mode > 2) fail("trace mode must be 0, 1 or 2");
  h->trace_mode = mode;
  TKB_CATCH
}

int tkb_trace_sync_on(tkb_handle h, void* cuda_stream) {
  TKB_TRY
  Workspace* ws = stream_ws(h, (cudaStream_t)cuda_stream);
  if (ws->side) CUDA_OK(cudaStreamSynchronize(ws->side));
  TKB_CATCH
}

int tkb_trace_sync(tkb_handle h) {
  TKB_TRY
  Workspace* ws = last_ws(h);
  if (ws->side) CUDA_OK(cudaStreamSynchronize(ws->side));
  TKB_CATCH
}

int tkb_trace_blob_on(tkb_handle h, void* cuda_stream, const void** blob, size_t* len) {
  TKB_TRY
  Workspace* ws = stream_ws(h, (cudaStream_t)cuda_stream);
  if (ws->trace_mode == 1) CUDA_OK(cudaStreamSynchronize((cudaStream_t)cuda_stream));  // the run itself must be complete
  finish_trace_blob(h, ws);
  *blob = ws->trace_host;
  *len = (size_t)h->trace_host_bytes;
  TKB_CATCH
}

int tkb_trace_blob(tkb_handle h, const void** blob, size_t* len) {
  TKB_TRY
  cudaStream_t s;
  {
    std::lock_guard<std::mutex> lk(h->ws_mu);
    s = h->last_stream;
  }
  if (tkb_trace_blob_on(h, (void*)s, blob, len) != 0) fail("%s", g_last_error.c_str());
  TKB_CATCH
}

static void write_file(const char* path, const void* data, size_t len) {
  FILE* f = fopen(path, "wb");
  if (!f) fail("cannot open %s for writing", path);
  const size_t w = fwrite(data, 1, len, f);
  fclose(f);
  if (w != len) fail("short write to %s", path);
}

int tkb_trace_save(tkb_handle h, const char* path) {
  TKB_TRY
  const void* blob;
  size_t len;
  if (tkb_trace_blob(h, &blob, &len) != 0) fail("%s", g_last_error.c_str());
  write_file(path, blob, len);
  TKB_CATCH
}

// Debug-executor-compatible dump of the last run (python/tvm/contrib/debugger/debug_result.py): the reference's
// DebugResult writes (i) output_tensors.params = save_tensors({"{name}____topo-index:{i}____output-num:{j}": array}) over
// ALL nodes of the graph, parameters included (:114-149), and (ii) _tvmdbg_graph_dump.json = the graph with, per node,
// "inputs" replaced by the producers' names, "op" = "param" or the function name, attrs["T"] = "type: <dtype>" and
// "shape" (:74-90, :191-202).  Same files here, from the tachikoma JSON graph and the captured trace.
static std::string json_escape(const std::string& in) {
  std::string o;
  for (char ch : in) {
    if (ch == '"' || ch == '\\') { o.push_back('\\'); o.push_back(ch); }
    else if (ch == '\n') o += "\\n";
    else o.push_back(ch);
  }
  return o;
}

int tkb_trace_dump_debug(tkb_handle h, const char* dir) {
  TKB_TRY
  if (!dir) fail("tkb_trace_dump_debug: null directory");
  Workspace* ws = last_ws(h);
  {
    cudaStream_t s;
    { std::lock_guard<std::mutex> lk(h->ws_mu); s = h->last_stream; }
    if (ws->trace_mode == 1) CUDA_OK(cudaStreamSynchronize(s));
  }
  finish_trace_blob(h, ws);
  const Graph& g = h->g;
  auto root_of = [&](int eid) { while (h->ent[eid].alias_of >= 0) eid = h->ent[eid].alias_of; return eid; };
  std::unordered_map<int, size_t> rec_of_root;  // root entry -> a trace record holding its bytes
  for (size_t i = 0; i < h->trace_recs.size(); ++i) rec_of_root.emplace(root_of(h->trace_recs[i].eid), i);
  // ---- output_tensors.params ----
  std::vector<uint8_t> out;
  std::vector<std::pair<std::string, int>> keys;  // (key, eid)
  for (size_t nid = 0; nid < g.nodes.size(); ++nid)
    for (uint32_t j = 0; j < g.node_row_ptr[nid + 1] - g.node_row_ptr[nid]; ++j)
      keys.emplace_back(g.nodes[nid].name + "____topo-index:" + std::to_string(nid) + "____output-num:" + std::to_string(j),
                        (int)g.eid((uint32_t)nid, j));
  put_u64(out, 0xF7E58D4F05049CB7ull);
  put_u64(out, 0);
  put_u64(out, keys.size());
  for (auto& k : keys) put_str(out, k.first);
  put_u64(out, keys.size());
  for (auto& k : keys) {
    const Entry& e = h->ent[k.second];
    const Entry& re = h->ent[root_of(k.second)];
    const uint8_t* src = nullptr;
    if (re.is_const) src = re.host.data();
    else {
      auto it = rec_of_root.find(root_of(k.second));
      if (it == rec_of_root.end()) fail("tkb_trace_dump_debug: entry %d (%s) is not part of the trace", k.second, k.first.c_str());
      src = ws->trace_host + h->trace_host_off[it->second];
    }
    put_u64(out, 0xDD5E40F096B4A13Full);
    put_u64(out, 0);
    put_i32(out, 1);
    put_i32(out, 0);
    put_i32(out, (int32_t)e.shape.size());
    uint8_t code, bits;
    dl_dtype(e.dtype, &code, &bits);
    out.push_back(code); out.push_back(bits); out.push_back(1); out.push_back(0);
    for (auto d : e.shape) put_u64(out, (uint64_t)d);
    put_u64(out, (uint64_t)e.bytes());
    out.insert(out.end(), src, src + e.bytes());
  }
  write_file((std::string(dir) + "/output_tensors.params").c_str(), out.data(), out.size());
  // ---- _tvmdbg_graph_dump.json ----
  std::ostringstream js;
  auto shape_str = [](const std::vector<int64_t>& shp) {
    std::string o = "[";
    for (size_t i = 0; i < shp.size(); ++i) o += (i ? ", " : "") + std::to_string(shp[i]);
    return o + "]";
  };
  js << "{\n    \"nodes\": [";
  for (size_t nid = 0; nid < g.nodes.size(); ++nid) {
    const GNode& n = g.nodes[nid];
    js << (nid ? "," : "") << "\n        {\n            \"op\": \"" << json_escape(n.op_type == "kernel" ? n.name : std::string("param"))
       << "\",\n            \"name\": \"" << json_escape(n.name) << "\",\n            \"inputs\": [";
    for (size_t i = 0; i < n.inputs.size(); ++i) js << (i ? ", " : "") << "\"" << json_escape(g.nodes[n.inputs[i][0]].name) << "\"";
    js << "],\n            \"attrs\": {";
    for (auto& kv : n.attrs) {
      js << "\n                \"" << json_escape(kv.first) << "\": [";
      for (size_t i = 0; i < kv.second.size(); ++i) js << (i ? ", " : "") << "\"" << json_escape(kv.second[i]) << "\"";
      js << "],";
    }
    js << "\n                \"num_inputs\": \"" << n.inputs.size() << "\",\n                \"num_outputs\": \"" << n.num_outputs
       << "\",\n                \"T\": \"type: " << (n.dtypes.empty() ? "" : n.dtypes[0]) << "\"\n            },\n            \"shape\": "
       << (n.shapes.empty() ? "[]" : shape_str(n.shapes[0])) << "\n        }";
  }
  js << "\n    ],\n    \"arg_nodes\": [";
  for (size_t i = 0; i < g.arg_nodes.size(); ++i) js << (i ? ", " : "") << g.arg_nodes[i];
  js << "],\n    \"node_row_ptr\": [";
  for (size_t i = 0; i < g.node_row_ptr.size(); ++i) js << (i ? ", " : "") << g.node_row_ptr[i];
  js << "],\n    \"heads\": [";
  for (size_t i = 0; i < g.heads.size(); ++i) js << (i ? ", " : "") << "[" << g.heads[i][0] << ", " << g.heads[i][1] << ", " << g.heads[i][2] << "]";
  js << "],\n    \"attrs\": {\n        \"shape\": [\"list_shape\", [";
  for (size_t e = 0; e < h->ent.size(); ++e) js << (e ? ", " : "") << shape_str(h->ent[e].shape);
  js << "]],\n        \"dltype\": [\"list_str\", [";
  {
    size_t e = 0;
    for (size_t nid = 0; nid < g.nodes.size(); ++nid)
      for (auto& d : g.nodes[nid].dtypes) js << (e++ ? ", " : "") << "\"" << d << "\"";
  }
  js << "]]\n    },\n    \"symbol\": \"" << json_escape(h->symbol) << "\"\n}\n";
  const std::string text = js.str();
  write_file((std::string(dir) + "/_tvmdbg_graph_dump.json").c_str(), text.data(), text.size());
  TKB_CATCH
}

int64_t tkb_trace_d2h_bytes(tkb_handle h) { return h->trace_d2h_bytes; }

int64_t tkb_trace_bytes(tkb_handle h) { return h->trace_host_bytes; }

void tkb_fixed_point_multiplier_shift(double multiplier, int32_t* significand, int32_t* exponent) {
  fixed_point_multiplier_shift(multiplier, significand, exponent);
}

int tkb_debug_chain_stats(uint64_t* out8, int reset) {
  TKB_TRY
  CUDA_OK(cudaDeviceSynchronize());
  tc_chain_stats((unsigned long long*)out8, reset != 0);
  TKB_CATCH
}

int tkb_spin(double microseconds, void* cuda_stream) {
  TKB_TRY
  if (microseconds < 0 || microseconds > 1e6) fail("tkb_spin: 0 .. 1e6 microseconds");
  launch_spin(microseconds, (cudaStream_t)cuda_stream);
  CUDA_OK(cudaGetLastError());
  TKB_CATCH
}

int tkb_probe_int8_shape(int iters, int n, int commit_every, double* tops, void* cuda_stream) {
  TKB_TRY
  if (n < 16 || n > 256 || (n % 16)) fail("probe: N must be a multiple of 16 in [16, 256]");
  if (probe_int8_shape(iters, n, commit_every, tops, (cudaStream_t)cuda_stream) != 0)
    fail("peak probe failed: %s", cudaGetErrorString(cudaGetLastError()));
  TKB_CATCH
}

int tkb_probe_int8_peak(int iters, double* tops, void* cuda_stream) {
  TKB_TRY
  if (probe_int8_peak(iters, tops, (cudaStream_t)cuda_stream) != 0)
    fail("peak probe failed: %s", cudaGetErrorString(cudaGetLastError()));
  TKB_CATCH
}

}  // extern "C"
