"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the float32 operators the reference's tachikoma backend hands to
oneDNN (src/runtime/contrib/tachikoma/tachikoma_json_runtime.cc:240-290 dispatch, :292-764 primitives), written from the
operators' Relay definitions, not from oneDNN:
  nn.conv1d/2d/3d          include/tvm/topi/nn.h + python/tvm/topi/nn/conv*.py   (cross-correlation, NC[D]HW x OI[D]HW)
  nn.conv*_transpose       python/tvm/topi/nn/conv2d_transpose.py:75-115 (dilate, pad, flipped-kernel convolution) -- here
                           as the equivalent scatter: y[o, i*s + k*d - p] += x[c, i] * w[c, o, k]
  nn.dense / batch_matmul  python/tvm/topi/nn/dense.py, batch_matmul.py
  nn.batch_norm            src/relay/transforms/simplify_inference.cc:35-73  ((x - mean) / sqrt(var + eps) * gamma + beta)
  nn.layer_norm            src/relay/transforms/simplify_inference.cc:75-105
  nn.softmax, pooling, elementwise, gelu / swish / mish compositions  python/tvm/relay/op/contrib/tachikoma.py:131-160
Sums are accumulated in float64 and rounded once to float32, so the oracle is the more accurate side of the comparison;
the bar is the reference's own rtol = atol = 1e-3 (tests/python/contrib/test_tachikoma.py:155).  Parity for this half is
"unpinned" in the strict sense: the reference holds no golden vectors for it (its tests compare two builds of TVM), and
neither TVM nor oneDNN can be built here.  Only tests/ may import this module.
"""
from __future__ import annotations

import itertools
import math

import numpy as np

_erf = np.vectorize(math.erf, otypes=[np.float64])


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).astype(np.float32))


def _norm_pad(padding, sp):
    p = [int(v) for v in (padding if isinstance(padding, (list, tuple)) else [padding])]
    if len(p) == 1:
        p = p * (2 * sp)
    elif len(p) == sp:
        p = p + p
    assert len(p) == 2 * sp, (padding, sp)
    return p[:sp], p[sp:]


def _tuple(v, sp):
    v = [int(x) for x in (v if isinstance(v, (list, tuple)) else [v])]
    return v * sp if len(v) == 1 else v


def conv_nd(x, w, strides, padding, dilation, groups=1):
    x, w = np.asarray(x, np.float64), np.asarray(w, np.float64)
    sp = x.ndim - 2
    st, di = _tuple(strides, sp), _tuple(dilation, sp)
    pb, pe = _norm_pad(padding, sp)
    xp = np.pad(x, [(0, 0), (0, 0)] + list(zip(pb, pe)))
    n, c = x.shape[:2]
    o, cg = w.shape[:2]
    og = o // groups
    assert cg * groups == c
    ks = w.shape[2:]
    out = [(xp.shape[2 + i] - di[i] * (ks[i] - 1) - 1) // st[i] + 1 for i in range(sp)]
    acc = np.zeros([n, o] + out, np.float64)
    for tap in itertools.product(*[range(k) for k in ks]):
        sl = tuple(slice(tap[i] * di[i], tap[i] * di[i] + (out[i] - 1) * st[i] + 1, st[i]) for i in range(sp))
        xs = xp[(slice(None), slice(None)) + sl]
        for g in range(groups):
            wt = w[(slice(g * og, (g + 1) * og), slice(None)) + tap]  # [og, cg]
            acc[:, g * og:(g + 1) * og] += np.tensordot(wt, xs[:, g * cg:(g + 1) * cg], axes=([1], [1])).swapaxes(0, 1)
    return acc


def conv_transpose_nd(x, w, strides, padding, dilation, output_padding, groups=1):
    x, w = np.asarray(x, np.float64), np.asarray(w, np.float64)
    sp = x.ndim - 2
    st, di = _tuple(strides, sp), _tuple(dilation, sp)
    op = _tuple(output_padding if output_padding else 0, sp)
    pb, pe = _norm_pad(padding, sp)
    n, c = x.shape[:2]
    og = w.shape[1]
    cg = c // groups
    ks = w.shape[2:]
    ins = x.shape[2:]
    full = [(ins[i] - 1) * st[i] + di[i] * (ks[i] - 1) + 1 + op[i] for i in range(sp)]
    acc = np.zeros([n, og * groups] + full, np.float64)
    for tap in itertools.product(*[range(k) for k in ks]):
        sl = tuple(slice(tap[i] * di[i], tap[i] * di[i] + (ins[i] - 1) * st[i] + 1, st[i]) for i in range(sp))
        for g in range(groups):
            wt = w[(slice(g * cg, (g + 1) * cg), slice(None)) + tap]  # [cg, og]
            acc[(slice(None), slice(g * og, (g + 1) * og)) + sl] += \
                np.tensordot(wt, x[:, g * cg:(g + 1) * cg], axes=([0], [1])).swapaxes(0, 1)
    crop = tuple(slice(pb[i], full[i] - pe[i]) for i in range(sp))
    return acc[(slice(None), slice(None)) + crop]


def pool_nd(x, pool_size, strides, padding, dilation, is_avg, count_include_pad):
    x = np.asarray(x, np.float64)
    sp = x.ndim - 2
    ks, st, di = _tuple(pool_size, sp), _tuple(strides, sp), _tuple(dilation if dilation else 1, sp)
    pb, pe = _norm_pad(padding, sp)
    fill = 0.0 if is_avg else -np.inf
    xp = np.pad(x, [(0, 0), (0, 0)] + list(zip(pb, pe)), constant_values=fill)
    ones = np.pad(np.ones(x.shape[2:]), list(zip(pb, pe)))
    out = [(xp.shape[2 + i] - di[i] * (ks[i] - 1) - 1) // st[i] + 1 for i in range(sp)]
    acc = np.full(list(x.shape[:2]) + out, fill, np.float64)
    cnt = np.zeros(out, np.float64)
    for tap in itertools.product(*[range(k) for k in ks]):
        sl = tuple(slice(tap[i] * di[i], tap[i] * di[i] + (out[i] - 1) * st[i] + 1, st[i]) for i in range(sp))
        xs = xp[(slice(None), slice(None)) + sl]
        acc = acc + xs if is_avg else np.maximum(acc, xs)
        cnt += ones[sl]
    if is_avg:
        acc = acc / (float(np.prod(ks)) if count_include_pad else np.maximum(cnt, 1.0))
    return acc


def activation(v, act, a=0.0, b=0.0):
    v = np.asarray(v, np.float64)
    if act in (None, "none"):
        return v
    if act == "relu":
        return np.maximum(v, 0.0)
    if act == "tanh":
        return np.tanh(v)
    if act == "sigmoid":
        return 1.0 / (1.0 + np.exp(-v))
    if act == "clip":
        return np.clip(v, a, b)
    if act == "gelu":
        return 0.5 * v * (1.0 + _erf(v / math.sqrt(2.0)))
    if act == "swish":
        return v / (1.0 + np.exp(-v))
    if act == "mish":
        return v * np.tanh(np.log1p(np.exp(v)))
    raise NotImplementedError(act)


_UNARY = {
    "abs": np.abs, "exp": np.exp, "log": np.log, "sqrt": np.sqrt, "round": np.rint, "tanh": np.tanh, "erf": _erf,
    "nn.relu": lambda v: np.maximum(v, 0.0), "sigmoid": lambda v: 1.0 / (1.0 + np.exp(-v)),
}

FLOAT_OPS = set(_UNARY) | {
    "nn.conv1d", "nn.conv2d", "nn.conv3d", "nn.conv1d_transpose", "nn.conv2d_transpose", "nn.conv3d_transpose", "nn.dense",
    "nn.batch_matmul", "nn.batch_norm", "nn.layer_norm", "nn.softmax", "nn.leaky_relu", "multiply", "divide",
    "nn.max_pool3d", "nn.avg_pool3d",
}


def eval_float_op(op, attrs, args):
    """One primitive float operator of the mini IR; returns float32."""
    a = [np.asarray(v) for v in args]
    g = attrs.get
    if op in _UNARY:
        with np.errstate(all="ignore"):
            return _f32(_UNARY[op](a[0].astype(np.float64)))
    if op == "nn.leaky_relu":
        x = a[0].astype(np.float64)
        return _f32(np.where(x > 0, x, x * float(g("alpha", 0.01))))
    if op in ("nn.conv1d", "nn.conv2d", "nn.conv3d"):
        return _f32(conv_nd(a[0], a[1], g("strides", 1), g("padding", 0), g("dilation", 1), int(g("groups", 1))))
    if op in ("nn.conv1d_transpose", "nn.conv2d_transpose", "nn.conv3d_transpose"):
        return _f32(conv_transpose_nd(a[0], a[1], g("strides", 1), g("padding", 0), g("dilation", 1), g("output_padding", 0),
                                      int(g("groups", 1))))
    if op == "nn.dense":
        x = a[0].astype(np.float64)
        return _f32(x.reshape(-1, x.shape[-1]) @ a[1].astype(np.float64).T).reshape(tuple(x.shape[:-1]) + (a[1].shape[0],))
    if op == "nn.batch_matmul":
        x, y = a[0].astype(np.float64), a[1].astype(np.float64)
        if g("transpose_a", False):
            x = x.transpose(0, 2, 1)
        if g("transpose_b", True):
            y = y.transpose(0, 2, 1)
        return _f32(np.matmul(x, y))
    if op == "nn.batch_norm":
        x = a[0].astype(np.float64)
        shp = [1] * x.ndim
        shp[int(g("axis", 1))] = -1
        ga, be, me, va = [v.astype(np.float64).reshape(shp) for v in a[1:5]]
        return _f32((x - me) / np.sqrt(va + float(g("epsilon", 1e-5))) * ga + be)
    if op == "nn.layer_norm":
        x = a[0].astype(np.float64)
        ax = int(g("axis", -1))
        mean = x.mean(axis=ax, keepdims=True)
        var = ((x - mean) ** 2).mean(axis=ax, keepdims=True)
        return _f32((x - mean) / np.sqrt(var + float(g("epsilon", 1e-5))) * a[1].astype(np.float64) + a[2].astype(np.float64))
    if op == "nn.softmax":
        x = a[0].astype(np.float64)
        ax = int(g("axis", -1))
        e = np.exp(x - x.max(axis=ax, keepdims=True))
        return _f32(e / e.sum(axis=ax, keepdims=True))
    if op == "multiply":
        return _f32(a[0].astype(np.float64) * a[1].astype(np.float64))
    if op == "divide":
        return _f32(a[0].astype(np.float64) / a[1].astype(np.float64))
    if op in ("nn.max_pool3d", "nn.avg_pool3d"):
        return float_pool(op, attrs, a[0])
    raise NotImplementedError(op)


def float_pool(op, attrs, x):
    g = attrs.get
    x = np.asarray(x)
    if "global" in op:
        return _f32(x.astype(np.float64).mean(axis=tuple(range(2, x.ndim)), keepdims=True))
    return _f32(pool_nd(x, g("pool_size"), g("strides"), g("padding", 0), g("dilation", 1), "avg_pool" in op,
                        bool(g("count_include_pad", False))))


def float_add(a, b):
    return _f32(np.asarray(a, np.float64) + np.asarray(b, np.float64))


def eval_float_composite(name, attrs, ins):
    """A float composite of the pattern table (tachikoma.<op>[_bias][_sum][_<act>]) evaluated as the reference's runtime
    does: primitive, + bias on the channel axis, + sum, activation (tachikoma_json_runtime.cc:185-233)."""
    root = name.split(".", 1)[1]
    base = root.split("_")[0]  # conv2d, deconv2d, dense ...
    if base.startswith("deconv"):
        y = conv_transpose_nd(ins[0], ins[1], attrs.get("strides", 1), attrs.get("padding", 0), attrs.get("dilation", 1),
                              attrs.get("output_padding", 0), int(attrs.get("groups", 1)))
    elif base.startswith("conv"):
        y = conv_nd(ins[0], ins[1], attrs.get("strides", 1), attrs.get("padding", 0), attrs.get("dilation", 1),
                    int(attrs.get("groups", 1)))
    elif base == "dense":
        x = np.asarray(ins[0], np.float64)
        y = (x.reshape(-1, x.shape[-1]) @ np.asarray(ins[1], np.float64).T).reshape(tuple(x.shape[:-1]) + (ins[1].shape[0],))
    else:
        raise NotImplementedError(name)
    if "_bias" in root:
        shp = [1] * y.ndim
        shp[1 if base != "dense" else -1] = -1
        y = y + np.asarray(ins[2], np.float64).reshape(shp)
    if "_sum" in root:
        y = y + np.asarray(ins[-1], np.float64)
    for act in ("relu", "tanh", "sigmoid", "clip", "gelu", "swish", "mish"):
        if root.endswith("_" + act):
            y = activation(y, act, float(attrs.get("a_min", 0.0)), float(attrs.get("a_max", 0.0)))
    return _f32(y)
