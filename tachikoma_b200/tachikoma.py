"""Host-side mirror of the reference's tachikoma BYOC wrapper and codegen for the QNN hot path.

Reference interfaces mirrored here (paths relative to /root/reference):
  python/tvm/relay/op/contrib/tachikoma.py
      pattern_table()                 :418-458   composite names
      make_qnn_conv2d_pattern()       :356-385   qnn.conv2d -> add -> requantize -> clip -> cast [-> qnn.add -> clip]
      make_qnn_dense_pattern()        :388-415
      LegalizeQnnOpForTachikoma       :1122-1305 float post-op constants (path "B")
      legalize_qnn_for_tachikoma()    :1308-1322
  tests/python/contrib/test_tachikoma.py
      partition_for_tachikoma()       :61-132
  src/relay/backend/contrib/tachikoma/comp_op_matcher.h
      ArgPacker / ParseQnnConvComp / ParseQnnDenseComp :78-221  argument order + *_idx attributes
  src/relay/backend/contrib/tachikoma/codegen.cc
      TachikomaJSONSerializer         :464-541   node emission
      TachikomaCompiler               :548-570   (symbol, graph_json, const_names)
  src/relay/backend/contrib/codegen_json/codegen_json.h
      OpAttrExtractor                 :57-145    attrs -> list<string>, floats with max_digits10
      JSONSerializer::Save            :329-349   {"nodes","arg_nodes","heads","node_row_ptr"}

Two epilogue modes exist:
  mode="int"        (default) the composite keeps the original QNN constants, so the runtime executes
                    TVM's exact integer lowering (fixed-point multiplier + shift, UPWARD/int64).
                    Constants are packed with the ArgPacker convention under the names the
                    reference's own pattern gives them (tachikoma.py:1146-1164): src_zp, wgh_zp,
                    src_scl, wgh_scl, rq_in_scl, rq_in_zp, rq_out_scl, rq_out_zp, sum_*.
  mode="dnnl_float" the reference's literal legalization: float32 o_scl/act_scl/sum_scl/dst_zp and a
                    float bias, executed as oneDNN-style float post-ops (judged at +-1 like the
                    reference's own test, tests/python/contrib/test_tachikoma.py:1615-1616).
"""
from __future__ import annotations

import json
import logging

import numpy as np

from . import relay as R

logger = logging.getLogger("Tachikoma")

_ACTS = ["relu", "tanh", "sigmoid", "clip", "gelu", "swish", "mish", None]
_FLOAT_OPS = ["conv1d", "conv2d", "conv3d", "deconv2d", "deconv3d", "dense"]


def pattern_table():
    """Composite names in the reference's order (tachikoma.py:418-458): the two QNN composites (the tensor-core integer
    path) and the float patterns (conv / deconv / dense [+ bias] [+ sum] [+ activation], executed as fp32 kernels)."""
    names = ["tachikoma.qnn.conv2d", "tachikoma.qnn.dense", "tachikoma.conv2d_bias_sum_relu",
             "tachikoma.conv2d_bias_sum"]
    for op in _FLOAT_OPS:
        for with_bias in (True, False):
            for act in _ACTS:
                n = f"tachikoma.{op}"
                if with_bias:
                    n += "_bias"
                if act:
                    n += f"_{act}"
                names.append(n)
    return names


def get_optimal_layout_for_conv(data_layout, kernel_layout, *_args, **_kw):
    """Reference: tachikoma.py:461-487 asks oneDNN for a blocked layout.  The CUDA runtime has one
    internal layout (NHWC activations, OHWI weights) and converts at the subgraph boundary."""
    return "NHWC,OHWI,NHWC"


def get_optimal_layout_for_conv_transpose(data_layout, kernel_layout, *_args, **_kw):
    return "NHWC,OHWI,NHWC"


# ------------------------------------------------------------------------------------------------
# pattern matching on the mini IR
# ------------------------------------------------------------------------------------------------

def _users(body):
    users = {}
    for e in R.post_order(body):
        if isinstance(e, R.Call):
            for a in e.args:
                users.setdefault(id(a), []).append(e)
    return users


def _is_const(e):
    return isinstance(e, R.Constant)


def _match_qnn_chain(tail, users):
    """Match, ending at ``tail``:
         root(qnn.conv2d|qnn.dense) -> [add|nn.bias_add] -> qnn.requantize -> clip -> cast
         [-> qnn.add(cast, sum_src) -> clip]
    Returns a dict of the matched nodes or None.  Interior nodes must have a single user."""
    m = {}
    e = tail

    def single(x):
        return len(users.get(id(x), [])) == 1

    if isinstance(e, R.Call) and e.op == "clip" and isinstance(e.args[0], R.Call) and e.args[0].op == "qnn.add":
        qadd = e.args[0]
        if not single(qadd) or not all(_is_const(a) for a in qadd.args[2:]):
            return None
        m["sum_clip"], m["qadd"] = e, qadd
        e = qadd.args[0]
        m["sum_src"] = qadd.args[1]
        if not single(e):
            return None
    if not (isinstance(e, R.Call) and e.op == "cast"):
        return None
    m["cast"] = e
    e = e.args[0]
    if not (isinstance(e, R.Call) and e.op == "clip" and single(e)):
        return None
    m["clip"] = e
    e = e.args[0]
    if not (isinstance(e, R.Call) and e.op == "qnn.requantize" and single(e)):
        return None
    if not all(_is_const(a) for a in e.args[1:]):
        return None
    m["rq"] = e
    e = e.args[0]
    if isinstance(e, R.Call) and e.op in ("add", "nn.bias_add") and single(e):
        m["bias_op"], m["bias"] = e, e.args[1]
        e = e.args[0]
    if not (isinstance(e, R.Call) and e.op in ("qnn.conv2d", "qnn.dense") and single(e)):
        return None
    if not all(_is_const(a) for a in e.args[2:]):
        return None
    # reference pattern requires a zero weight zero point (tachikoma.py:1152)
    if np.any(e.args[3].data != 0):
        return None
    m["root"] = e
    return m


def _pattern_string(m):
    ops = [m["root"].op]
    if "bias_op" in m:
        ops.append(m["bias_op"].op)
    ops += ["qnn.requantize", "clip", "cast"]
    if "qadd" in m:
        ops += ["qnn.add", "clip"]
    return "_".join(ops) + "_"


def _f32(x):
    return np.asarray(x, dtype=np.float32)


def _legalize_float(m):
    """LegalizeQnnOpForTachikoma.callback (tachikoma.py:1195-1277), all in float32 like the
    constant-folded Relay expressions."""
    root = m["root"]
    wgh = root.args[1].data
    src_zp = root.args[2].data.astype(np.int32)
    rq = m["rq"]
    rq_in_scl, rq_in_zp, rq_out_scl, rq_out_zp = [a.data for a in rq.args[1:5]]
    if "qadd" in m:
        lhs_scl, lhs_zp, rhs_scl, rhs_zp, out_scl, out_zp = [a.data for a in m["qadd"].args[2:8]]
    else:
        lhs_scl, lhs_zp, rhs_scl, rhs_zp, out_scl, out_zp = (_f32(1), np.int32(0), _f32(0), np.int32(0), _f32(1),
                                                             np.int32(0))
    o_scl = _f32(rq_in_scl) / _f32(rq_out_scl)
    act_scl = _f32(lhs_scl) / _f32(out_scl)
    sum_scl = _f32(rhs_scl) / _f32(out_scl)
    dst_zp = _f32(out_zp) - _f32(lhs_zp) * _f32(lhs_scl) / _f32(out_scl) - _f32(rhs_zp) * _f32(rhs_scl) / _f32(out_scl)
    if root.op == "qnn.conv2d":
        wl = root.attrs["kernel_layout"]
        if wl == "HWOI":
            red = wgh.astype(np.int32).sum(axis=(0, 1)).reshape(-1)  # [C, mult] -> out channel c*mult+m
        else:
            red = wgh.astype(np.int32).sum(axis=tuple(i for i in range(4) if i != wl.index("O")))
    else:
        red = wgh.astype(np.int32).sum(axis=1)
    bias = m["bias"].data.reshape(-1) if "bias" in m else np.zeros(1, dtype=np.int32)
    # fake_op (tachikoma.py:1279-1288): zp * reduced kernel, int32
    bias_f = _f32(bias) - _f32((src_zp.reshape(-1)[0] * red).astype(np.int32)) - _f32(rq_in_zp) \
        + _f32(rq_out_zp) * _f32(rq_out_scl) / _f32(rq_in_scl)
    return dict(bias=_f32(bias_f).reshape(-1), o_scl=_f32(o_scl).reshape(-1), act_scl=_f32(act_scl).reshape(()),
                sum_scl=_f32(sum_scl).reshape(()), dst_zp=_f32(dst_zp).reshape(()))


def _make_composite(m, mode):
    root = m["root"]
    name = "tachikoma.qnn.conv2d" if root.op == "qnn.conv2d" else "tachikoma.qnn.dense"
    tail = m.get("sum_clip", m["cast"])
    args, attrs = [], dict(root.attrs)

    def put(expr, tag=None):
        if expr is None:
            return -1
        args.append(expr)
        if tag:
            attrs[tag] = len(args) - 1
        return len(args) - 1

    put(root.args[0])
    put(root.args[1])
    if mode == "int":
        if "bias" in m and not _is_const(m["bias"]) and "bias_op" in m:
            put(m["bias"], "bias_idx")
        elif "bias" in m:
            put(R.Constant(m["bias"].data.reshape(-1)), "bias_idx")
        put(m.get("sum_src"), "sum_idx")
        for tag, e in zip(("src_zp", "wgh_zp", "src_scl", "wgh_scl"), root.args[2:6]):
            put(e, tag + "_idx")
        for tag, e in zip(("rq_in_scl", "rq_in_zp", "rq_out_scl", "rq_out_zp"), m["rq"].args[1:5]):
            put(e, tag + "_idx")
        if "qadd" in m:
            for tag, e in zip(("sum_lhs_scl", "sum_lhs_zp", "sum_rhs_scl", "sum_rhs_zp", "sum_out_scl", "sum_out_zp"),
                              m["qadd"].args[2:8]):
                put(e, tag + "_idx")
            attrs["sum_clip_a_min"] = m["sum_clip"].attrs["a_min"]
            attrs["sum_clip_a_max"] = m["sum_clip"].attrs["a_max"]
        rqa = m["rq"].attrs
        attrs.update(rq_axis=rqa["axis"], rounding=rqa["rounding"], compute_dtype=rqa["compute_dtype"],
                     rq_out_dtype=rqa["out_dtype"], clip_a_min=m["clip"].attrs["a_min"],
                     clip_a_max=m["clip"].attrs["a_max"], cast_dtype=m["cast"].attrs["dtype"])
    else:
        if "bias" in m and not _is_const(m["bias"]):
            return None  # the reference folds the bias into a float constant; it must be constant
        lg = _legalize_float(m)
        put(R.Constant(lg["bias"]), "bias_idx")
        put(m.get("sum_src"), "sum_idx")
        put(R.Constant(lg["o_scl"]), "o_scl_idx")
        if "qadd" in m:
            put(R.Constant(lg["act_scl"]), "act_scl_idx")
            put(R.Constant(lg["sum_scl"]), "sum_scl_idx")
        put(R.Constant(lg["dst_zp"]), "dst_zp_idx")
        # Activation triple: act_scale, alpha, beta as extra constants (comp_op_matcher.h:151-156)
        i0 = put(R.Constant(lg["act_scl"] if "qadd" in m else _f32(1.0).reshape(())))
        i1 = put(R.Constant(_f32(0.0).reshape(())))
        i2 = put(R.Constant(_f32(255.0).reshape(())))
        attrs["activation"] = ["clip", str(i0), str(i1), str(i2)]
    attrs["PartitionedFromPattern"] = _pattern_string(m)
    return R.Call(name, args, attrs, tail.shape, tail.dtype)


# ------------------------------------------------------------------------------------------------
# partitioning
# ------------------------------------------------------------------------------------------------

#: operators the CUDA runtime executes outside a composite (single-op JSON kernel nodes)
SUPPORTED_OPS = {
    "qnn.conv2d", "qnn.dense", "qnn.batch_matmul", "qnn.requantize", "qnn.add", "qnn.subtract", "qnn.mul", "qnn.concatenate",
    "qnn.quantize", "qnn.dequantize", "qnn.leaky_relu", "reinterpret", "take", "nn.bias_add", "add", "clip",
    "cast", "nn.max_pool2d", "nn.avg_pool2d", "nn.global_avg_pool2d", "nn.pad", "reshape", "nn.batch_flatten",
    "transpose",
}

#: float32 operators the CUDA runtime executes as single-operator nodes (tachikoma.py:88-114, float operands only)
FLOAT_SUPPORTED_OPS = {
    "nn.batch_norm", "nn.conv1d", "nn.conv2d", "nn.conv3d", "nn.conv2d_transpose", "nn.conv3d_transpose", "nn.dense",
    "nn.max_pool2d", "nn.avg_pool2d", "nn.global_avg_pool2d", "nn.max_pool3d", "nn.avg_pool3d", "abs", "clip", "exp", "log",
    "sqrt", "round", "nn.relu", "nn.leaky_relu", "tanh", "sigmoid", "nn.softmax", "add", "multiply", "nn.layer_norm",
    "nn.batch_matmul",
}


def _hardswish(x):
    return x * np.clip(x + 3.0, 0.0, 6.0) / 6.0


def _unary_float_funcs():
    """python/tvm/relay/qnn/op/legalizations.py:68-86 (the numpy functions the reference registers)."""
    from scipy import special
    return {"qnn.sqrt": np.sqrt, "qnn.rsqrt": lambda a: 1 / np.sqrt(a), "qnn.exp": np.exp, "qnn.erf": special.erf,
            "qnn.sigmoid": lambda a: 1 / (1 + np.exp(-a)), "qnn.hardswish": _hardswish, "qnn.tanh": np.tanh,
            "qnn.log": np.log, "qnn.abs": np.abs}


def create_integer_lookup_table(floating_point_func, input_scale, input_zero_point, output_scale, output_zero_point,
                                in_dtype="uint8", out_dtype="uint8"):
    """python/tvm/relay/qnn/op/canonicalizations.py:31-103: out[b] = quantize(f(dequantize(b))) for all 256 bit
    patterns b of the input dtype, indexed by the pattern read as uint8.  dequantize = float32(int32(b) - zp) * scale
    (dequantize.cc:95-127); quantize = clip(round(x / scale) + zp) in float32, round half away from zero
    (quantize.cc:113-150)."""
    if np.dtype(in_dtype).itemsize != 1 or np.dtype(out_dtype).itemsize != 1:
        raise ValueError(f"Only 8-bit integer dtypes are supported, got {in_dtype} and {out_dtype}")
    pat = np.arange(256, dtype=np.uint8).view(in_dtype)
    with np.errstate(all="ignore"):
        deq = (pat.astype(np.int32) - np.int32(input_zero_point)).astype(np.float32) * np.float32(input_scale)
        f = np.asarray(floating_point_func(deq), dtype=np.float32)
        v = (f / np.float32(output_scale)).astype(np.float32)
        t = np.trunc(v)
        r = np.where(np.isfinite(v), t + np.sign(v) * (np.abs(v - t) >= 0.5), v).astype(np.float32)
        r = (r + np.float32(output_zero_point)).astype(np.float32)
        info = np.iinfo(out_dtype)
        r = np.where(r < info.min, info.min, np.where(r > info.max, info.max, r))
        r = np.where(np.isnan(r), 0.0, r)  # undefined in the reference (float -> int cast of NaN)
    return r.astype(np.int64).astype(out_dtype)


def legalize_qnn_unary(func):
    """relay.qnn.transform.Legalize for the table-lookup unary operators (legalizations.py:54-71,
    canonicalizations.py:106-160): qnn.<f>(x, s_in, zp_in, s_out, zp_out) -> take(table, reinterpret(x, uint8))."""
    funcs = None
    memo = {}

    def visit(e):
        nonlocal funcs
        if id(e) in memo:
            return memo[id(e)]
        out = e
        if isinstance(e, R.Call):
            args = [visit(x) for x in e.args]
            if e.op in R.QNN_UNARY_OPS:
                if not all(_is_const(a) and np.asarray(a.data).size == 1 for a in args[1:5]):
                    raise ValueError(f"Unsupported op: {e.op} with non-scalar quantization parameters")
                funcs = funcs or _unary_float_funcs()
                sc = lambda a: np.asarray(a.data).reshape(-1)[0]
                table = create_integer_lookup_table(funcs[e.op], np.float32(sc(args[1])), int(sc(args[2])),
                                                    np.float32(sc(args[3])), int(sc(args[4])), e.args[0].dtype, e.dtype)
                out = R.take(R.const(table), R.reinterpret(args[0], "uint8"), axis=0, mode="fast")
            elif any(a is not b for a, b in zip(args, e.args)):
                out = R.Call(e.op, args, dict(e.attrs), e.shape, e.dtype)
        memo[id(e)] = out
        return out

    body = visit(func.body)
    return func if body is func.body else R.Function(func.params, body, dict(func.attrs))


def legalize_qnn_for_tachikoma(func, mode="dnnl_float"):
    """Rewrite every QNN chain into a tachikoma composite call (MergeComposite of the QNN patterns;
    with mode="dnnl_float" also the reference's LegalizeQnnOpForTachikoma constant derivation)."""
    users = _users(func.body)
    memo = {}

    def visit(e):
        if id(e) in memo:
            return memo[id(e)]
        r = e
        if isinstance(e, R.Call):
            m = _match_qnn_chain(e, users) if e.op in ("cast", "clip") else None
            # a cast that feeds a fusable qnn.add -> clip is fused by the longer match at the clip
            if m is not None and "qadd" not in m and e.op == "cast":
                us = users.get(id(e), [])
                if len(us) == 1 and us[0].op == "qnn.add" and us[0].args[0] is e:
                    uu = users.get(id(us[0]), [])
                    if len(uu) == 1 and uu[0].op == "clip" and _match_qnn_chain(uu[0], users) is not None:
                        m = None
            comp = None
            if m is not None:
                comp = _make_composite(m, mode)
                if comp is not None:
                    # re-point the data inputs (src, sum_src, non-constant bias) at rewritten producers
                    comp.args = [visit(a) if not isinstance(a, R.Constant) else a for a in comp.args]
            if comp is not None:
                r = comp
            else:
                r = R.Call(e.op, [visit(a) for a in e.args], e.attrs, e.shape, e.dtype)
        memo[id(e)] = r
        return r

    return R.Function(func.params, visit(func.body), func.attrs)


# ------------------------------------------------------------------------------------------------
# float patterns (tachikoma.py:131-160 append_eltwise_ops, :163-187 make_conv_pattern, :190-213 conv_bias_sum[_relu],
# :296-321 make_dense_pattern, :324-353 names): conv / deconv / dense [+ bias] [+ sum] [+ activation] -> one composite
# ------------------------------------------------------------------------------------------------
_FLOAT_ROOTS = {"nn.conv1d": "tachikoma.conv1d", "nn.conv2d": "tachikoma.conv2d", "nn.conv3d": "tachikoma.conv3d",
                "nn.conv2d_transpose": "tachikoma.deconv2d", "nn.conv3d_transpose": "tachikoma.deconv3d",
                "nn.dense": "tachikoma.dense"}


def _is_call(e, op):
    return isinstance(e, R.Call) and e.op == op


def _strip_activation(e, users):
    """Returns (inner, act, attrs, shared) if ``e`` is the last node of an activation sub-pattern over ``inner``; ``shared``
    = how many users ``inner`` has inside the pattern.  Interior nodes must have no user outside the pattern."""
    def n_users(x):
        return len(users.get(id(x), []))

    for op, act in (("nn.relu", "relu"), ("tanh", "tanh"), ("sigmoid", "sigmoid"), ("clip", "clip")):
        if _is_call(e, op):
            return e.args[0], act, ({"a_min": e.attrs["a_min"], "a_max": e.attrs["a_max"]} if act == "clip" else {}), 1
    if _is_call(e, "multiply"):
        a, b = e.args
        # swish: multiply(op, sigmoid(op))
        if _is_call(b, "sigmoid") and b.args[0] is a and n_users(b) == 1:
            return a, "swish", {}, 2
        # mish: multiply(op, tanh(log(add(exp(op), c))))
        if _is_call(b, "tanh") and n_users(b) == 1 and _is_call(b.args[0], "log") and n_users(b.args[0]) == 1:
            ad = b.args[0].args[0]
            if _is_call(ad, "add") and n_users(ad) == 1 and _is_call(ad.args[0], "exp") and n_users(ad.args[0]) == 1 and \
                    ad.args[0].args[0] is a and _is_const(ad.args[1]):
                return a, "mish", {}, 2
        # gelu: multiply(multiply(op, add(erf(divide(op, c1)), c2)), c3)
        if _is_call(a, "multiply") and n_users(a) == 1 and _is_const(b):
            x, added = a.args
            if _is_call(added, "add") and n_users(added) == 1 and _is_const(added.args[1]) and _is_call(added.args[0], "erf") and \
                    n_users(added.args[0]) == 1:
                dv = added.args[0].args[0]
                if _is_call(dv, "divide") and n_users(dv) == 1 and dv.args[0] is x and _is_const(dv.args[1]):
                    return x, "gelu", {}, 2
    return None


def _match_float_chain(tail, users):
    """Longest float pattern that ends at ``tail``; returns (name, root, bias, sum_src, extra_attrs) or None."""
    if not isinstance(tail, R.Call) or tail.dtype != "float32":
        return None

    def n_users(x):
        return len(users.get(id(x), []))

    cands = []
    st = _strip_activation(tail, users)
    if st is not None:
        cands.append(st)
    cands.append((tail, None, {}, None))
    for inner, act, extra, shared in cands:
        cur = inner
        if shared is not None and n_users(cur) != shared:
            continue
        # conv2d + bias + sum [+ relu] (tachikoma.py:190-213, predicate :251-264)
        if act in (None, "relu") and _is_call(cur, "add") and cur.args[0].shape == cur.args[1].shape and \
                _is_call(cur.args[0], "add") and n_users(cur.args[0]) == 1:
            b = cur.args[0]
            if _is_call(b.args[0], "nn.conv2d") and n_users(b.args[0]) == 1 and \
                    int(np.prod(b.args[1].shape)) == b.args[0].attrs["channels"]:
                name = "tachikoma.conv2d_bias_sum" + ("_relu" if act else "")
                return name, b.args[0], b.args[1], cur.args[1], {}
        bias = None
        root = cur
        if isinstance(cur, R.Call) and cur.op in ("add", "nn.bias_add"):
            # add is commutative for the dataflow matcher: (root, bias) or (bias, root)
            for r0, b0, first in ((cur.args[0], cur.args[1], True), (cur.args[1], cur.args[0], False)):
                if not (isinstance(r0, R.Call) and r0.op in _FLOAT_ROOTS and n_users(r0) == 1) or (cur.op == "nn.bias_add" and not first):
                    continue
                ok = True
                if "conv" in r0.op and act in (None, "relu"):
                    # make_bias_add_pattren_predicate / add_checker (:267-293): evaluated when the pattern ends in the add
                    # or in relu(add); it looks at args[0] / args[1] as written and accepts nn.conv2d with a bias of
                    # `channels` elements only
                    ok = first and r0.op == "nn.conv2d" and int(np.prod(b0.shape)) == r0.attrs["channels"]
                if ok:
                    bias, root = b0, r0
                    break
        if not (isinstance(root, R.Call) and root.op in _FLOAT_ROOTS):
            continue
        if root.dtype != "float32" or any(a.dtype != "float32" for a in root.args):
            continue
        if bias is None and root is not cur:
            continue
        name = _FLOAT_ROOTS[root.op] + ("_bias" if bias is not None else "") + (f"_{act}" if act else "")
        return name, root, bias, None, extra  # a bare conv / dense is the pattern without bias and activation
    return None


def merge_float_composites(func):
    """MergeComposite of the float rows of pattern_table() on the mini IR."""
    users = _users(func.body)
    memo = {}

    def visit(e):
        if id(e) in memo:
            return memo[id(e)]
        r = e
        if isinstance(e, R.Call):
            m = _match_float_chain(e, users)
            if m is not None:
                name, root, bias, sum_src, extra = m
                args = [visit(root.args[0]), visit(root.args[1])]
                if bias is not None:
                    args.append(R.Constant(bias.data.reshape(-1)) if _is_const(bias) else visit(bias))
                if sum_src is not None:
                    args.append(visit(sum_src))
                attrs = dict(root.attrs)
                attrs.update(extra)
                r = R.Call(name, args, attrs, e.shape, e.dtype)
            else:
                r = R.Call(e.op, [visit(a) for a in e.args], e.attrs, e.shape, e.dtype)
        memo[id(e)] = r
        return r

    return R.Function(func.params, visit(func.body), func.attrs)


#: tachikoma.py:711-723 (IsComputeIntensiveGraph) plus the QNN forms the integer path offloads
COMPUTE_INTENSIVE_OPS = {
    "nn.conv1d", "nn.conv2d", "nn.conv2d_transpose", "nn.conv3d", "nn.conv3d_transpose", "nn.dense", "nn.layer_norm",
    "nn.batch_matmul", "nn.global_avg_pool2d", "qnn.conv2d", "qnn.dense", "qnn.batch_matmul",
    "tachikoma.qnn.conv2d", "tachikoma.qnn.dense",
}


def is_valid_subgraph(body):
    """tachikoma.py:739-741: a subgraph is worth offloading only if it holds a compute-intensive operator."""
    return any(isinstance(e, R.Call) and e.op in COMPUTE_INTENSIVE_OPS for e in R.post_order(body))


def prune_tachikoma_subgraphs(func):
    """tachikoma.py:744-793 reverts offloaded subgraphs without a compute-intensive operator to TVM, and does nothing
    when the module holds a single subgraph (:775-777).  partition_for_tachikoma here always produces exactly one
    subgraph (the whole function, or an error), so the faithful behaviour is the reference's early return; the
    validity check is still evaluated so that a caller can log what the reference would have pruned."""
    if not is_valid_subgraph(func.body):
        logger.info("subgraph without a compute-intensive operator kept (single-subgraph module, tachikoma.py:775)")
    return func


# rewrite_layer_norm, rewrite_dense_bias_gelu_reshape_last, rewrite_resnetv1 (tachikoma.py:796-1119): rewrites.py
from .rewrites import rewrite_dense_bias_gelu_reshape_last, rewrite_layer_norm, rewrite_resnetv1  # noqa: E402,F401


def partition_for_tachikoma(func, params=None, mode="int", symbol="tvmgen_default_tachikoma_main_0"):
    """tests/python/contrib/test_tachikoma.py:61-132 on the mini IR: bind params, fuse the QNN
    patterns into composites, check that every remaining op is supported (the reference's
    ``check_fully_annotated``, :1347-1364) and mark the function for the tachikoma compiler."""
    if params:
        func = R.bind_params(func, params)
    from .frontend import fold_explicit_padding  # SimplifyExpr / FoldExplicitPadding, test_tachikoma.py:88
    func = fold_explicit_padding(func)
    func = legalize_qnn_unary(func)  # relay.qnn.transform.Legalize: table-lookup unary operators
    func = legalize_qnn_for_tachikoma(func, mode=mode)
    func = merge_float_composites(func)
    for e in R.post_order(func.body):
        if isinstance(e, R.Call) and e.op == "tuple" and e is func.body:
            continue  # several outputs: one graph head per field
        if isinstance(e, R.Call) and not e.op.startswith("tachikoma.") and e.op not in SUPPORTED_OPS:
            if e.op in FLOAT_SUPPORTED_OPS and e.dtype == "float32" and all(a.dtype == "float32" for a in e.args):
                continue
            logger.info("%s is not supported by the tachikoma CUDA runtime", e.op)
            raise ValueError(f"Unsupported op: {e.op}")
    func = prune_tachikoma_subgraphs(func)
    func.attrs.update(Compiler="tachikoma", global_symbol=symbol)
    return func


# ------------------------------------------------------------------------------------------------
# codegen: function -> (symbol, graph_json, const_names, consts)
# ------------------------------------------------------------------------------------------------

def _fp2str(v):
    # OpAttrExtractor::Fp2String: ostringstream with precision max_digits10 (17 for double)
    return "%.17g" % float(v)


def _attr_strings(v):
    if isinstance(v, bool):
        return [str(int(v))]
    if isinstance(v, (int, np.integer)):
        return [str(int(v))]
    if isinstance(v, (float, np.floating)):
        return [_fp2str(v)]
    if isinstance(v, str):
        return [v]
    if isinstance(v, (list, tuple)):
        out = []
        for x in v:
            if isinstance(x, (list, tuple)):  # pad_width: list of pairs -> flattened
                out += [s for y in x for s in _attr_strings(y)]
            else:
                out += _attr_strings(x)
        return out
    raise TypeError(f"attribute of type {type(v)}")


def serialize(func, symbol=None):
    """TachikomaJSONSerializer + TachikomaCompiler.  Returns (symbol, graph_json, const_names,
    consts) where consts are numpy arrays in const_names order; const names are
    ``"tachikoma_" + symbol + "_const_" + i`` in visit order (codegen_json.h:261-269)."""
    symbol = symbol or func.attrs.get("global_symbol", "tvmgen_default_tachikoma_main_0")
    nodes, entry, const_names, consts = [], {}, [], []

    def add_node(node, e):
        node.setdefault("attrs", {})
        node["attrs"]["shape"] = [[list(e.shape)]]
        node["attrs"]["dtype"] = [[e.dtype]]
        nodes.append(node)
        entry[id(e)] = [len(nodes) - 1, 0, 0]

    for p in func.params:
        add_node(dict(op="input", name=p.name), p)
    for e in R.post_order(func.body):
        if id(e) in entry:
            continue
        if isinstance(e, R.Var):
            raise ValueError(f"free variable {e.name} is not a function parameter")
        if isinstance(e, R.Call) and e.op == "tuple":
            if e is not func.body:
                raise ValueError("Unsupported op: tuple (only as the function body)")
            continue
        if isinstance(e, R.Constant):
            name = f"tachikoma_{symbol}_const_{len(const_names)}"
            const_names.append(name)
            consts.append(e.data)
            add_node(dict(op="const", name=name), e)
            continue
        attrs = {k: [_attr_strings(v)] for k, v in e.attrs.items()}
        attrs["num_inputs"] = str(len(e.args))
        attrs["num_outputs"] = "1"
        add_node(dict(op="kernel", name=e.op, inputs=[entry[id(a)] for a in e.args], attrs=attrs), e)
        if e.op == "nn.batch_norm":
            # the operator is a 3-tuple (normalised data, moving mean, moving variance): the JSON node declares all three
            # entries (codegen_json.h:203-224), inference reads entry 0 only (tachikoma_json_runtime.cc:556)
            c = [int(e.args[1].shape[0])]
            attrs["num_outputs"] = "3"
            attrs["shape"] = [[list(e.shape), c, c]]
            attrs["dtype"] = [[e.dtype] * 3]
    row_ptr = [0]
    for n in nodes:
        row_ptr.append(row_ptr[-1] + int(n["attrs"].get("num_outputs", "1")))
    graph = dict(nodes=nodes, arg_nodes=[i for i, n in enumerate(nodes) if n["op"] != "kernel"],
                 heads=[entry[id(f)] for f in func.body.args] if getattr(func.body, "op", "") == "tuple" else [entry[id(func.body)]],
                 node_row_ptr=row_ptr)
    return symbol, json.dumps(graph), const_names, consts
