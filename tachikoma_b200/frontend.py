"""Importer for PREQUANTIZED TorchScript models.

Restates what the reference's PyTorch frontend does for quantized graphs
(python/tvm/relay/frontend/qnn_torch.py) on the mini IR of this package, emitting the exact operator chains the
reference emits, so that a real eager-mode-quantized model (torch.ao.quantization: fuse -> prepare -> calibrate ->
convert -> torch.jit.trace) runs on the integer path:

  quantized::conv2d[_relu]   nn.pad(pad_value = input zero point) -> qnn.conv2d(padding 0) -> nn.bias_add(int32 bias)
                             -> qnn.requantize(axis=1, out_dtype=int32) -> clip(zp_out if relu else 0, 255) -> cast(uint8)
                             (qnn_torch.py:685-766, :649-682; bias quantized at import time with scale
                             s_in * s_w and zero point 0, :157-164, :205-210)
  quantized::linear[_relu]   qnn.dense -> the same tail                                   (:768-797)
  quantized::add[_relu]      qnn.add [-> clip(zp_out, 255)]                               (:800-893)
  quantized::mul[_relu]      qnn.mul [-> clip(zp_out, 255)]                               (:800-893)
  quantized::cat             qnn.concatenate                                              (:895-931)
  aten::quantize_per_tensor  qnn.quantize(out_dtype=uint8, axis=1)                        (:616-628)
  aten::dequantize           qnn.dequantize                                               (:631-638)
  aten::adaptive_avg_pool2d / aten::avg_pool2d / aten::mean on a quantized tensor:
                             cast(int32) -> pool -> cast(uint8)  (apply_with_upcast, :587-590)
  aten::max_pool2d, aten::flatten, aten::reshape / view, aten::relu (= clip(zp, 255), :599-602)
  quantized::leaky_relu -> qnn.leaky_relu (:982-1012); quantized::hardswish -> qnn.hardswish (:1041-1078);
  aten::sigmoid / aten::tanh on a quantized tensor -> qnn.sigmoid / qnn.tanh with torch's fixed output parameters
  (:399-423); aten::hardtanh / relu6 -> clip with quantized bounds (:964-979)

Input quantization parameters are not operands in TorchScript (they live in the QTensor); like the reference
(:240-296, :426-526) they are found by walking back to the producing op's output scale / zero point, here simply
carried along with every imported value.

torch is imported only inside from_torchscript (it is needed to read the model: packed parameters are unpacked with
torch.ops.quantized.*_unpack); nothing here runs torch math.
"""
from __future__ import annotations

import numpy as np

from . import relay as R


class _Q:
    """An imported value: expression + (scale, zero_point) when it is a quantized tensor."""

    def __init__(self, expr, scale=None, zp=None):
        self.expr, self.scale, self.zp = expr, scale, zp

    @property
    def quantized(self):
        return self.scale is not None


def _quantize_numpy(w, scale, zero_point, dtype):
    """qnn_torch.py:157-164."""
    info = np.iinfo(dtype)
    scale = np.asarray(scale, dtype=np.float64)
    if scale.ndim > 0:
        scale = scale.reshape([w.shape[0]] + [1] * (w.ndim - 1))
    return np.clip(np.round(zero_point + w / scale), info.min, info.max).astype(dtype)


def _weight_params(qweight):
    """(int8 weights, scale (float32 scalar or [O]), zero point) of a quantized weight tensor (qnn_torch.py:69-82)."""
    import torch
    w = qweight.int_repr().numpy().astype(np.int8)
    if qweight.qscheme() in (torch.per_tensor_affine, torch.per_tensor_symmetric):
        return w, np.float32(qweight.q_scale()), int(qweight.q_zero_point())
    zps = qweight.q_per_channel_zero_points().numpy()
    if not np.all(zps == 0):
        raise ValueError("The values of zero points should be all zero for per channel")
    return w, qweight.q_per_channel_scales().numpy().astype(np.float32), 0


def _bias_and_requantize(out, bias, s_in, s_w, s_out, zp_out, relu, axis):
    """qnn_torch.py:649-682."""
    rq_in = (np.float32(s_in) * np.asarray(s_w, dtype=np.float32)).astype(np.float32)
    if bias is not None:
        qb = _quantize_numpy(bias.astype(np.float64), rq_in.astype(np.float64), 0, np.int32)
        out = R.bias_add(out, R.const(qb), axis=axis)
    rq_scale = R.const(rq_in) if rq_in.ndim else np.float32(rq_in)
    out = R.qnn_requantize(out, rq_scale, 0, np.float32(s_out), int(zp_out), axis=axis, out_dtype="int32")
    out = R.clip(out, float(zp_out if relu else 0), 255.0)
    return R.cast(out, "uint8")


def _conv2d(x, packed, s_out, zp_out, relu):
    import torch
    qweight, bias = torch.ops.quantized.conv2d_unpack(packed)
    w, s_w, zp_w = _weight_params(qweight)
    stride, padding, dilation, groups = packed.stride(), packed.padding(), packed.dilation(), packed.groups()
    e = x.expr
    if padding[0] != 0 or padding[1] != 0:  # explicit pad with the input zero point (qnn_torch.py:732-740)
        e = R.pad(e, ((0, 0), (0, 0), (padding[0], padding[0]), (padding[1], padding[1])), pad_value=float(x.zp))
    s_w_arg = R.const(s_w) if np.ndim(s_w) else np.float32(s_w)
    e = R.qnn_conv2d(e, R.const(w), int(x.zp), int(zp_w), np.float32(x.scale), s_w_arg, (w.shape[2], w.shape[3]),
                     w.shape[0], strides=tuple(stride), padding=(0, 0), dilation=tuple(dilation), groups=int(groups))
    b = None if bias is None else bias.detach().numpy()
    return _Q(_bias_and_requantize(e, b, x.scale, s_w, s_out, zp_out, relu, 1), np.float32(s_out), int(zp_out))


def _linear(x, packed, s_out, zp_out, relu):
    import torch
    qweight, bias = torch.ops.quantized.linear_unpack(packed)
    w, s_w, zp_w = _weight_params(qweight)
    s_w_arg = R.const(s_w) if np.ndim(s_w) else np.float32(s_w)
    e = R.qnn_dense(x.expr, R.const(w), int(x.zp), int(zp_w), np.float32(x.scale), s_w_arg, w.shape[0])
    b = None if bias is None else bias.detach().numpy()
    return _Q(_bias_and_requantize(e, b, x.scale, s_w, s_out, zp_out, relu, 1), np.float32(s_out), int(zp_out))


def _upcast(x, fn):
    """apply_with_upcast (qnn_torch.py:587-590)."""
    return _Q(R.cast(fn(R.cast(x.expr, "int32")), "uint8"), x.scale, x.zp)


def _ints(v):
    return [int(a) for a in v]


def from_torchscript(script_module, input_shapes, input_names=None):
    """script_module: a traced / scripted eager-mode-quantized model (eval mode).  input_shapes: list of shapes of the
    float32 graph inputs.  Returns a mini-IR Function over float32 inputs."""
    import torch
    mod = script_module
    if not hasattr(mod, "graph"):
        raise TypeError("from_torchscript expects a torch.jit.ScriptModule (torch.jit.trace / torch.jit.script output)")
    try:
        mod = torch.jit.freeze(mod.eval())  # GetAttr chains (packed parameters, scales) become constants
    except Exception:  # already frozen
        pass
    graph = mod.graph
    torch._C._jit_pass_inline(graph)
    env = {}
    g_inputs = [v for v in graph.inputs() if v.type().kind() != "ClassType"]
    if len(g_inputs) != len(input_shapes):
        raise ValueError("the graph has %d tensor inputs, %d shapes given" % (len(g_inputs), len(input_shapes)))
    params = []
    for i, (v, shp) in enumerate(zip(g_inputs, input_shapes)):
        name = input_names[i] if input_names else "input%d" % i
        var = R.var(name, tuple(int(s) for s in shp), "float32")
        params.append(var)
        env[v.debugName()] = _Q(var)

    def val(v):
        return env[v.debugName()]

    for node in graph.nodes():
        kind = node.kind()
        outs = list(node.outputs())
        ins = list(node.inputs())
        if kind == "prim::Constant":
            env[outs[0].debugName()] = outs[0].toIValue() if outs[0].type().kind() != "NoneType" else None
            continue
        if kind in ("prim::ListConstruct", "prim::TupleConstruct"):
            env[outs[0].debugName()] = [val(v) for v in ins]
            continue
        a = [val(v) for v in ins]
        if kind == "aten::quantize_per_tensor":
            x, scale, zp = a[0], np.float32(a[1]), int(a[2])
            axis = 1 if len(x.expr.shape) > 1 else 0
            r = _Q(R.qnn_quantize(x.expr, scale, zp, axis=axis, out_dtype="uint8"), scale, zp)
        elif kind == "aten::dequantize":
            r = _Q(R.qnn_dequantize(a[0].expr, np.float32(a[0].scale), int(a[0].zp)))
        elif kind in ("quantized::conv2d", "quantized::conv2d_relu"):
            r = _conv2d(a[0], a[1], a[2], a[3], kind.endswith("_relu"))
        elif kind in ("quantized::linear", "quantized::linear_relu"):
            r = _linear(a[0], a[1], a[2], a[3], kind.endswith("_relu"))
        elif kind in ("quantized::add", "quantized::add_relu", "quantized::mul", "quantized::mul_relu"):
            op = R.qnn_add if "add" in kind else R.qnn_mul
            s_out, zp_out = np.float32(a[2]), int(a[3])
            e = op(a[0].expr, a[1].expr, np.float32(a[0].scale), int(a[0].zp), np.float32(a[1].scale), int(a[1].zp),
                   s_out, zp_out)
            if kind.endswith("_relu"):
                e = R.clip(e, float(zp_out), 255.0)
            r = _Q(e, s_out, zp_out)
        elif kind == "quantized::cat":
            xs, axis, s_out, zp_out = a[0], int(a[1]), np.float32(a[2]), int(a[3])
            e = R.qnn_concatenate([x.expr for x in xs], [np.float32(x.scale) for x in xs], [int(x.zp) for x in xs],
                                  s_out, zp_out, axis)
            r = _Q(e, s_out, zp_out)
        elif kind == "aten::max_pool2d":
            k, st, pd, dil, ceil = _ints(a[1]), _ints(a[2]) or _ints(a[1]), _ints(a[3]), _ints(a[4]), bool(a[5])
            if dil != [1, 1] or ceil:
                raise ValueError("Unsupported op: aten::max_pool2d with dilation / ceil_mode")
            r = _Q(R.max_pool2d(a[0].expr, tuple(k), tuple(st), tuple(pd)), a[0].scale, a[0].zp)
        elif kind == "aten::adaptive_avg_pool2d":
            osz, (h, w) = _ints(a[1]), a[0].expr.shape[2:4]
            if osz == [1, 1]:
                fn = R.global_avg_pool2d
            elif h % osz[0] == 0 and w % osz[1] == 0:
                kk = (h // osz[0], w // osz[1])
                fn = lambda e: R.avg_pool2d(e, kk, kk)  # noqa: E731
            else:
                raise ValueError("Unsupported op: aten::adaptive_avg_pool2d with a non-dividing output size")
            r = _upcast(a[0], fn) if a[0].quantized else _Q(fn(a[0].expr))
        elif kind == "aten::avg_pool2d":
            k, st, pd = _ints(a[1]), _ints(a[2]) or _ints(a[1]), _ints(a[3])
            fn = lambda e: R.avg_pool2d(e, tuple(k), tuple(st), tuple(pd))  # noqa: E731
            r = _upcast(a[0], fn) if a[0].quantized else _Q(fn(a[0].expr))
        elif kind == "aten::mean":
            dims = sorted(d % len(a[0].expr.shape) for d in _ints(a[1]))
            if dims != [2, 3]:
                raise ValueError("Unsupported op: aten::mean over dims %s" % dims)
            r = _upcast(a[0], R.global_avg_pool2d)
            if not bool(a[2]):  # keepdim = False
                r = _Q(R.reshape(r.expr, r.expr.shape[:2]), r.scale, r.zp)
        elif kind == "aten::flatten":
            start, end = int(a[1]), int(a[2])
            shp = list(a[0].expr.shape)
            end = end % len(shp)
            e = R.batch_flatten(a[0].expr) if (start, end) == (1, len(shp) - 1) else \
                R.reshape(a[0].expr, shp[:start] + [int(np.prod(shp[start:end + 1]))] + shp[end + 1:])
            r = _Q(e, a[0].scale, a[0].zp)
        elif kind in ("aten::reshape", "aten::view"):
            r = _Q(R.reshape(a[0].expr, _ints(a[1])), a[0].scale, a[0].zp)
        elif kind == "quantized::leaky_relu":  # (x, negative_slope, inplace, output_scale, output_zero_point), :982-1012
            s_out, zp_out = np.float32(a[3]), int(a[4])
            r = _Q(R.qnn_leaky_relu(a[0].expr, float(a[1]), np.float32(a[0].scale), int(a[0].zp), s_out, zp_out), s_out, zp_out)
        elif kind == "quantized::hardswish":  # (x, output_scale, output_zero_point), :1041-1078
            s_out, zp_out = np.float32(a[1]), int(a[2])
            r = _Q(R.qnn_unary("qnn.hardswish", a[0].expr, np.float32(a[0].scale), int(a[0].zp), s_out, zp_out), s_out, zp_out)
        elif kind in ("aten::sigmoid", "aten::tanh") and a[0].quantized:
            # fixed output parameters of torch's quantized kernels (qsigmoid.cpp / qtanh.cpp; :399-423 for sigmoid)
            s_out, zp_out = (np.float32(1.0 / 256), 0) if kind == "aten::sigmoid" else (np.float32(2.0 / 256), 128)
            r = _Q(R.qnn_unary("qnn." + kind[6:], a[0].expr, np.float32(a[0].scale), int(a[0].zp), s_out, zp_out), s_out, zp_out)
        elif kind in ("aten::hardtanh", "aten::hardtanh_", "aten::relu6", "aten::relu6_", "quantized::relu6") and a[0].quantized:
            # clip in the quantized domain with quantized bounds (quantize_scalar, :964-979)
            lo_f, hi_f = (float(a[1]), float(a[2])) if "hardtanh" in kind else (0.0, 6.0)
            qs = lambda v: max(0, min(int(round(a[0].zp + v / float(a[0].scale))), 255))
            r = _Q(R.clip(a[0].expr, float(qs(lo_f)), float(qs(hi_f))), a[0].scale, a[0].zp)
        elif kind in ("aten::relu", "aten::relu_"):
            if not a[0].quantized:
                raise ValueError("Unsupported op: aten::relu on a float tensor")
            r = _Q(R.clip(a[0].expr, float(a[0].zp), 255.0), a[0].scale, a[0].zp)  # maximum(x, zp), qnn_torch.py:599-602
        elif kind in ("aten::contiguous", "aten::detach", "aten::dropout", "aten::dropout_", "aten::feature_dropout"):
            r = a[0]
        else:
            raise ValueError("Unsupported op: %s" % kind)
        env[outs[0].debugName()] = r
    rets = [val(v) for v in graph.outputs()]
    if len(rets) != 1 or not isinstance(rets[0], _Q):
        raise ValueError("from_torchscript supports graphs with one tensor output")
    return R.Function(params, rets[0].expr)


def fold_explicit_padding(func):
    """The part of the reference's SimplifyExpr that partition_for_tachikoma relies on (FoldExplicitPadding,
    tests/python/contrib/test_tachikoma.py:88): nn.pad(constant, pad_value == input zero point, spatial axes only)
    feeding a qnn.conv2d with padding 0 becomes the convolution's padding attribute."""
    users = {}
    for e in R.post_order(func.body):
        if isinstance(e, R.Call):
            for x in e.args:
                users[id(x)] = users.get(id(x), 0) + 1
    memo = {}

    def visit(e):
        if id(e) in memo:
            return memo[id(e)]
        out = e
        if isinstance(e, R.Call):
            args = [visit(x) for x in e.args]
            out = e if all(x is y for x, y in zip(args, e.args)) else R.Call(e.op, args, dict(e.attrs), e.shape, e.dtype)
            p = e.args[0] if e.args else None
            if (e.op == "qnn.conv2d" and isinstance(p, R.Call) and p.op == "nn.pad" and users.get(id(p), 0) == 1
                    and p.attrs.get("pad_mode", "constant") == "constant" and e.attrs.get("data_layout", "NCHW") == "NCHW"
                    and all(int(v) == 0 for v in e.attrs["padding"]) and isinstance(e.args[2], R.Constant)
                    and isinstance(p.args[1], R.Constant)):
                pw = p.attrs["pad_width"]
                zp = np.asarray(e.args[2].data).reshape(-1)
                if (len(pw) == 4 and list(pw[0]) == [0, 0] and list(pw[1]) == [0, 0] and zp.size == 1
                        and float(np.asarray(p.args[1].data).reshape(-1)[0]) == float(zp[0])):
                    attrs = dict(e.attrs)
                    attrs["padding"] = [int(pw[2][0]), int(pw[3][0]), int(pw[2][1]), int(pw[3][1])]
                    out = R.Call(e.op, [visit(p.args[0])] + args[1:], attrs, e.shape, e.dtype)
            # float operators (src/relay/transforms/fold_explicit_padding.cc:45-180): zero padding folds into
            # nn.conv1d/2d/3d and into nn.avg_pool2d/3d (which then counts the padding), lowest-value padding into max pools
            elif (e.op in ("nn.conv1d", "nn.conv2d", "nn.conv3d", "nn.avg_pool2d", "nn.avg_pool3d", "nn.max_pool2d", "nn.max_pool3d")
                    and e.dtype == "float32" and isinstance(p, R.Call) and p.op == "nn.pad" and users.get(id(p), 0) == 1
                    and p.attrs.get("pad_mode", "constant") == "constant" and isinstance(p.args[1], R.Constant)
                    and e.attrs.get("data_layout", e.attrs.get("layout", "NC")).startswith("NC")):
                pw = [list(map(int, w)) for w in p.attrs["pad_width"]]
                sp = len(pw) - 2
                pv = float(np.asarray(p.args[1].data).reshape(-1)[0])
                is_max = "max_pool" in e.op
                value_ok = (pv == float("-inf") or pv <= float(np.finfo(np.float32).min)) if is_max else pv == 0.0
                old = [int(v) for v in e.attrs["padding"]]
                if len(old) == sp:
                    old = old + old
                pool_ok = "pool" not in e.op or is_max or e.attrs.get("count_include_pad", False) or all(v == 0 for v in old)
                if pw[0] == [0, 0] and pw[1] == [0, 0] and value_ok and pool_ok and len(old) == 2 * sp:
                    attrs = dict(e.attrs)
                    attrs["padding"] = [old[i] + pw[2 + i][0] for i in range(sp)] + [old[sp + i] + pw[2 + i][1] for i in range(sp)]
                    if "avg_pool" in e.op:
                        attrs["count_include_pad"] = True
                    out = R.Call(e.op, [visit(p.args[0])] + args[1:], attrs, e.shape, e.dtype)
        memo[id(e)] = out
        return out

    return R.Function(func.params, visit(func.body), dict(func.attrs))
