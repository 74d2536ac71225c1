"""Minimal Relay-shaped expression IR for the QNN hot path.

The reference consumes a TVM Relay ``IRModule``; TVM cannot be imported in this environment, so
this module provides just enough of Relay's surface (``var`` / ``const`` / ``Call`` with the same
operator names, argument order and attribute names) to describe the integer-only operator chain
that the reference's frontends emit (python/tvm/relay/frontend/qnn_torch.py:649-682, :736-764)
and that its tachikoma partitioner consumes (python/tvm/relay/op/contrib/tachikoma.py:1122-1277).

Operator constructors mirror ``tvm.relay.qnn.op.*`` / ``tvm.relay.nn.*`` signatures and perform
the reference's type relations (shape / dtype inference):
  qnn.conv2d   src/relay/qnn/op/convolution.cc:43-90
  qnn.dense    src/relay/qnn/op/dense.cc:40-74
  qnn.batch_matmul src/relay/qnn/op/batch_matmul.cc:40-74
  qnn.requantize   src/relay/qnn/op/requantize.cc:465-514
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "Expr", "Var", "Constant", "Call", "Function", "Tuple", "var", "const", "post_order",
    "qnn_conv2d", "qnn_dense", "qnn_batch_matmul", "qnn_requantize", "qnn_add", "qnn_mul", "qnn_concatenate",
    "qnn_quantize", "qnn_dequantize", "bias_add", "add",
    "clip", "cast", "max_pool2d", "avg_pool2d", "global_avg_pool2d", "pad", "reshape",
    "batch_flatten", "transpose",
    "conv1d", "conv2d", "conv3d", "conv2d_transpose", "conv3d_transpose", "dense", "batch_matmul", "batch_norm", "layer_norm",
    "softmax", "relu", "tanh", "sigmoid", "abs_", "exp", "log", "sqrt", "round_", "erf", "leaky_relu", "multiply", "divide",
    "max_pool3d", "avg_pool3d",
]


class Expr:
    shape: tuple
    dtype: str

    def __repr__(self):
        return f"{type(self).__name__}({getattr(self, 'name', getattr(self, 'op', ''))}, {self.shape}, {self.dtype})"


class Var(Expr):
    def __init__(self, name, shape, dtype):
        self.name, self.shape, self.dtype = name, tuple(int(s) for s in shape), str(dtype)


class Constant(Expr):
    def __init__(self, data):
        self.data = np.ascontiguousarray(data)
        self.shape, self.dtype = tuple(self.data.shape), str(self.data.dtype)


class Call(Expr):
    """An operator call.  ``op`` is the Relay operator name (e.g. ``"qnn.conv2d"``) or, after
    partitioning, a composite name (e.g. ``"tachikoma.qnn.conv2d"``)."""

    def __init__(self, op, args, attrs, shape, dtype):
        self.op, self.args, self.attrs = op, list(args), dict(attrs)
        self.shape, self.dtype = tuple(int(s) for s in shape), str(dtype)


class Function:
    def __init__(self, params, body, attrs=None):
        self.params, self.body, self.attrs = list(params), body, dict(attrs or {})


def Tuple(fields):
    """relay.Tuple as a function body (several outputs, tests/python/contrib/test_tachikoma.py:640-650): a pseudo call that
    the serializer turns into one graph head per field; it never becomes a node."""
    fields = list(fields)
    assert fields and all(isinstance(f, Expr) for f in fields)
    return Call("tuple", fields, {}, fields[0].shape, fields[0].dtype)


def var(name, shape, dtype="float32"):
    return Var(name, shape, dtype)


def const(value, dtype=None):
    arr = np.asarray(value)
    if dtype is None:
        # tvm.relay.const defaults: python float -> float32, python int -> int32
        if arr.dtype == np.float64:
            dtype = "float32"
        elif arr.dtype == np.int64:
            dtype = "int32"
        else:
            dtype = str(arr.dtype)
    return Constant(arr.astype(dtype))


def _c(x, dtype=None):
    return x if isinstance(x, Expr) else const(x, dtype)


def post_order(expr):
    """Topological (post-order DFS, args left to right) list of distinct nodes."""
    out, seen = [], set()
    stack = [(expr, False)]
    while stack:
        e, done = stack.pop()
        if id(e) in seen:
            continue
        if done:
            seen.add(id(e))
            out.append(e)
            continue
        stack.append((e, True))
        if isinstance(e, Call):
            for a in reversed(e.args):
                if id(a) not in seen:
                    stack.append((a, False))
    return out


def _pair(v):
    if isinstance(v, (int, np.integer)):
        return (int(v), int(v))
    return tuple(int(x) for x in v)


def _pad4(p):
    p = _pair(p) if not isinstance(p, (list, tuple)) or len(p) != 4 else tuple(int(x) for x in p)
    if len(p) == 2:  # (ph, pw) -> (top, left, bottom, right), as relay's GetPaddingHeightWidth
        return (p[0], p[1], p[0], p[1])
    if len(p) == 1:
        return (p[0],) * 4
    return p


def _layout_dims(shape, layout, want):
    return [shape[layout.index(ch)] for ch in want]


def qnn_conv2d(data, kernel, input_zero_point, kernel_zero_point, input_scale, kernel_scale,
               kernel_size, channels, strides=(1, 1), padding=(0, 0), dilation=(1, 1), groups=1,
               data_layout="NCHW", kernel_layout="OIHW", out_layout="", out_dtype="int32"):
    """tvm.relay.qnn.op.conv2d (python/tvm/relay/qnn/op/qnn.py), rel convolution.cc:43-90."""
    strides, dilation, pad = _pair(strides), _pair(dilation), _pad4(padding)
    n, c, h, w = _layout_dims(data.shape, data_layout, "NCHW")
    o, i, kh, kw = _layout_dims(kernel.shape, kernel_layout, "OIHW")
    if kernel_layout == "HWOI" or (groups > 1 and kernel_layout == "HWIO" and i != c // groups):
        # depthwise HWOI / HWIO with multiplier: O axis holds in-channels, I axis the multiplier
        pass
    assert data.dtype in ("int8", "uint8", "int16") and kernel.dtype in ("int8", "uint8")
    oh = (h + pad[0] + pad[2] - (kh - 1) * dilation[0] - 1) // strides[0] + 1
    ow = (w + pad[1] + pad[3] - (kw - 1) * dilation[1] - 1) // strides[1] + 1
    olay = out_layout or data_layout
    oshape = [0] * 4
    for ch, v in zip("NCHW", (n, channels, oh, ow)):
        oshape[olay.index(ch)] = v
    attrs = dict(strides=list(strides), padding=list(pad), dilation=list(dilation), groups=int(groups),
                 channels=int(channels), kernel_size=list(_pair(kernel_size)), data_layout=data_layout,
                 kernel_layout=kernel_layout, out_layout=out_layout, out_dtype=out_dtype)
    args = [data, kernel, _c(input_zero_point, "int32"), _c(kernel_zero_point, "int32"),
            _c(input_scale, "float32"), _c(kernel_scale, "float32")]
    return Call("qnn.conv2d", args, attrs, oshape, out_dtype)


def qnn_dense(data, weight, input_zero_point, kernel_zero_point, input_scale, kernel_scale, units,
              out_dtype="int32"):
    """tvm.relay.qnn.op.dense; weight is [units, in] (dense.cc:40-74)."""
    assert data.shape[-1] == weight.shape[1], (data.shape, weight.shape)
    attrs = dict(units=int(units), out_dtype=out_dtype)
    args = [data, weight, _c(input_zero_point, "int32"), _c(kernel_zero_point, "int32"),
            _c(input_scale, "float32"), _c(kernel_scale, "float32")]
    return Call("qnn.dense", args, attrs, tuple(data.shape[:-1]) + (weight.shape[0],), out_dtype)


def qnn_batch_matmul(x, y, x_zero_point, y_zero_point, x_scale, y_scale, out_dtype="int32"):
    """tvm.relay.qnn.op.batch_matmul: x [B,M,K], y [B,N,K] (transpose_b only, batch_matmul.cc:83-87)."""
    assert x.shape[2] == y.shape[2]
    b = max(x.shape[0], y.shape[0])
    attrs = dict(out_dtype=out_dtype, transpose_a=False, transpose_b=True)
    args = [x, y, _c(x_zero_point, "int32"), _c(y_zero_point, "int32"), _c(x_scale, "float32"),
            _c(y_scale, "float32")]
    return Call("qnn.batch_matmul", args, attrs, (b, x.shape[1], y.shape[1]), out_dtype)


def qnn_requantize(data, input_scale, input_zero_point, output_scale, output_zero_point, axis=-1,
                   rounding="None", compute_dtype="None", out_dtype="int8"):
    """tvm.relay.qnn.op.requantize (requantize.cc:465-514). rounding/compute_dtype "None" resolve to
    the RequantizeConfig defaults UPWARD / int64 (requantize_config.h:53-73)."""
    attrs = dict(axis=int(axis), rounding=rounding, compute_dtype=compute_dtype, out_dtype=out_dtype)
    args = [data, _c(input_scale, "float32"), _c(input_zero_point, "int32"),
            _c(output_scale, "float32"), _c(output_zero_point, "int32")]
    return Call("qnn.requantize", args, attrs, data.shape, out_dtype)


def qnn_add(lhs, rhs, lhs_scale, lhs_zero_point, rhs_scale, rhs_zero_point, output_scale,
            output_zero_point, lhs_axis=-1, rhs_axis=-1):
    """tvm.relay.qnn.op.add (add.cc:40-93); output dtype = input dtype."""
    assert lhs.shape == rhs.shape and lhs.dtype == rhs.dtype
    args = [lhs, rhs, _c(lhs_scale, "float32"), _c(lhs_zero_point, "int32"), _c(rhs_scale, "float32"),
            _c(rhs_zero_point, "int32"), _c(output_scale, "float32"), _c(output_zero_point, "int32")]
    return Call("qnn.add", args, dict(lhs_axis=int(lhs_axis), rhs_axis=int(rhs_axis)), lhs.shape, lhs.dtype)


def qnn_mul(lhs, rhs, lhs_scale, lhs_zero_point, rhs_scale, rhs_zero_point, output_scale,
            output_zero_point, lhs_axis=-1, rhs_axis=-1):
    """tvm.relay.qnn.op.mul (src/relay/qnn/op/mul.cc:43-101): (Qa - zp_a) * (Qb - zp_b) in int32, requantized from
    the float32 product of the two scales to the output qparams; output dtype = input dtype.  Per-tensor form."""
    assert lhs.shape == rhs.shape and lhs.dtype == rhs.dtype
    args = [lhs, rhs, _c(lhs_scale, "float32"), _c(lhs_zero_point, "int32"), _c(rhs_scale, "float32"),
            _c(rhs_zero_point, "int32"), _c(output_scale, "float32"), _c(output_zero_point, "int32")]
    return Call("qnn.mul", args, dict(lhs_axis=int(lhs_axis), rhs_axis=int(rhs_axis)), lhs.shape, lhs.dtype)


def qnn_subtract(lhs, rhs, lhs_scale, lhs_zero_point, rhs_scale, rhs_zero_point, output_scale,
                 output_zero_point, lhs_axis=-1, rhs_axis=-1):
    """tvm.relay.qnn.op.subtract (src/relay/qnn/op/subtract.cc:40-93): both sides requantized to the output qparams,
    Q_a' - Q_b' + zp_out in int32, saturated to the input dtype."""
    assert lhs.shape == rhs.shape and lhs.dtype == rhs.dtype
    args = [lhs, rhs, _c(lhs_scale, "float32"), _c(lhs_zero_point, "int32"), _c(rhs_scale, "float32"),
            _c(rhs_zero_point, "int32"), _c(output_scale, "float32"), _c(output_zero_point, "int32")]
    return Call("qnn.subtract", args, dict(lhs_axis=int(lhs_axis), rhs_axis=int(rhs_axis)), lhs.shape, lhs.dtype)


def qnn_concatenate(data, input_scales, input_zero_points, output_scale, output_zero_point, axis):
    """tvm.relay.qnn.op.concatenate (src/relay/qnn/op/concatenate.cc:138-213): every input whose (scale, zero point)
    differs from the output's is requantized to them (same dtype), then nn.concatenate.  Arguments of the call: the n
    tensors, their n scales, their n zero points, output scale, output zero point (the tuples flattened, as the JSON
    serializer does)."""
    data = list(data)
    n = len(data)
    assert n >= 1 and len(input_scales) == n and len(input_zero_points) == n
    rank = len(data[0].shape)
    ax = axis + rank if axis < 0 else axis
    oshape = list(data[0].shape)
    oshape[ax] = sum(d.shape[ax] for d in data)
    for d in data:
        assert d.dtype == data[0].dtype and len(d.shape) == rank
        assert all(d.shape[i] == oshape[i] for i in range(rank) if i != ax)
    args = data + [_c(s, "float32") for s in input_scales] + [_c(z, "int32") for z in input_zero_points] + \
        [_c(output_scale, "float32"), _c(output_zero_point, "int32")]
    return Call("qnn.concatenate", args, dict(axis=int(axis), num_inputs=n), oshape, data[0].dtype)


def qnn_quantize(data, output_scale, output_zero_point, axis=-1, out_dtype="int8"):
    """tvm.relay.qnn.op.quantize (src/relay/qnn/op/quantize.cc:113-150): clip(round(x / scale) + zp, qmin, qmax)
    in float32, cast to out_dtype; scale / zero point scalar or per-``axis`` vectors."""
    assert data.dtype == "float32"
    return Call("qnn.quantize", [data, _c(output_scale, "float32"), _c(output_zero_point, "int32")],
                dict(axis=int(axis), out_dtype=out_dtype), data.shape, out_dtype)


def qnn_dequantize(data, input_scale, input_zero_point, axis=-1):
    """tvm.relay.qnn.op.dequantize (src/relay/qnn/op/dequantize.cc:95-127): float32(int32(x) - zp) * scale."""
    return Call("qnn.dequantize", [data, _c(input_scale, "float32"), _c(input_zero_point, "int32")],
                dict(axis=int(axis)), data.shape, "float32")


#: LUT-based unary QNN operators (python/tvm/relay/qnn/op/legalizations.py:54-86)
QNN_UNARY_OPS = ("qnn.sqrt", "qnn.rsqrt", "qnn.exp", "qnn.erf", "qnn.sigmoid", "qnn.hardswish", "qnn.tanh", "qnn.log",
                 "qnn.abs")


def qnn_unary(op, x, scale, zero_point, output_scale, output_zero_point):
    """relay.qnn.op.{sqrt,rsqrt,exp,erf,sigmoid,hardswish,tanh,log,abs}(x, scale, zero_point, output_scale,
    output_zero_point): 8-bit in, same dtype out; legalized to a 256-entry table lookup (legalizations.py:54-86)."""
    assert op in QNN_UNARY_OPS and x.dtype in ("int8", "uint8")
    return Call(op, [x, _c(scale, "float32"), _c(zero_point, "int32"), _c(output_scale, "float32"),
                     _c(output_zero_point, "int32")], {}, x.shape, x.dtype)


def qnn_leaky_relu(x, alpha, input_scale, input_zero_point, output_scale, output_zero_point):
    """relay.qnn.op.leaky_relu (src/relay/qnn/op/leaky_relu.cc:70-139)."""
    return Call("qnn.leaky_relu", [x, _c(input_scale, "float32"), _c(input_zero_point, "int32"),
                                   _c(output_scale, "float32"), _c(output_zero_point, "int32")],
                dict(alpha=float(alpha)), x.shape, x.dtype)


def reinterpret(data, dtype):
    """relay.reinterpret: same bits, other dtype of the same width."""
    assert np.dtype(dtype).itemsize == np.dtype(data.dtype).itemsize
    return Call("reinterpret", [data], dict(dtype=dtype), data.shape, dtype)


def take(table, indices, axis=0, mode="fast"):
    """relay.take(table, indices, axis=0, mode="fast") with a 1-D table: out[i...] = table[indices[i...]]."""
    table = _c(table)
    assert len(table.shape) == 1 and axis == 0
    return Call("take", [table, indices], dict(axis=0, mode=mode, batch_dims=0), indices.shape, table.dtype)


def bias_add(data, bias, axis=1):
    return Call("nn.bias_add", [data, _c(bias)], dict(axis=int(axis)), data.shape, data.dtype)


def add(lhs, rhs):
    rhs = _c(rhs)
    shape = np.broadcast_shapes(lhs.shape, rhs.shape)
    return Call("add", [lhs, rhs], {}, shape, lhs.dtype)


def clip(data, a_min, a_max):
    return Call("clip", [data], dict(a_min=float(a_min), a_max=float(a_max)), data.shape, data.dtype)


def cast(data, dtype):
    return Call("cast", [data], dict(dtype=dtype), data.shape, dtype)


def _pool(op, data, pool_size, strides, padding, layout, extra):
    ps, st, pad = _pair(pool_size), _pair(strides), _pad4(padding)
    n, c, h, w = _layout_dims(data.shape, layout, "NCHW")
    oh = (h + pad[0] + pad[2] - ps[0]) // st[0] + 1
    ow = (w + pad[1] + pad[3] - ps[1]) // st[1] + 1
    oshape = [0] * 4
    for ch, v in zip("NCHW", (n, c, oh, ow)):
        oshape[layout.index(ch)] = v
    attrs = dict(pool_size=list(ps), strides=list(st), dilation=[1, 1], padding=list(pad), layout=layout,
                 out_layout="", ceil_mode=False, **extra)
    return Call(op, [data], attrs, oshape, data.dtype)


def max_pool2d(data, pool_size=(1, 1), strides=(1, 1), padding=(0, 0), layout="NCHW"):
    return _pool("nn.max_pool2d", data, pool_size, strides, padding, layout, {})


def avg_pool2d(data, pool_size=(1, 1), strides=(1, 1), padding=(0, 0), layout="NCHW",
               count_include_pad=False):
    return _pool("nn.avg_pool2d", data, pool_size, strides, padding, layout,
                 dict(count_include_pad=bool(count_include_pad)))


def global_avg_pool2d(data, layout="NCHW"):
    oshape = list(data.shape)
    oshape[layout.index("H")] = 1
    oshape[layout.index("W")] = 1
    return Call("nn.global_avg_pool2d", [data], dict(layout=layout, out_layout=""), oshape, data.dtype)


def pad(data, pad_width, pad_value=0.0, pad_mode="constant"):
    pw = [[int(a), int(b)] for a, b in pad_width]
    oshape = [s + a + b for s, (a, b) in zip(data.shape, pw)]
    return Call("nn.pad", [data, _c(float(pad_value), "float32")], dict(pad_width=pw, pad_mode=pad_mode),
                oshape, data.dtype)


def reshape(data, newshape):
    n = int(np.prod(data.shape))
    ns = [int(s) for s in newshape]
    if -1 in ns:
        k = ns.index(-1)
        ns[k] = n // int(-np.prod(ns))
    assert int(np.prod(ns)) == n
    return Call("reshape", [data], dict(newshape=[int(s) for s in newshape]), ns, data.dtype)


def batch_flatten(data):
    return Call("nn.batch_flatten", [data], {}, (data.shape[0], int(np.prod(data.shape[1:]))), data.dtype)


def transpose(data, axes):
    axes = [int(a) for a in axes]
    return Call("transpose", [data], dict(axes=axes), [data.shape[a] for a in axes], data.dtype)


# ----------------------------------------------------------------------------------------------
# float32 operators of the tachikoma float patterns (python/tvm/relay/op/contrib/tachikoma.py:88-114): constructors with
# relay.nn.* / relay.* signatures and type relations, default layouts NC[D]HW x OI[D]HW
# ----------------------------------------------------------------------------------------------

def _ntuple(v, n):
    v = [int(v)] if isinstance(v, (int, np.integer)) else [int(x) for x in v]
    return v * n if len(v) == 1 else v


def _npad(p, n):
    p = _ntuple(p, n)
    if len(p) == n:  # (p_d, p_h, p_w) -> begin + end, as relay's GetPadding*
        p = p + p
    assert len(p) == 2 * n, p
    return p


def _conv_nd(op, sp, data, weight, strides, padding, dilation, groups, channels, kernel_size, data_layout, kernel_layout,
             out_layout, out_dtype):
    assert len(data.shape) == sp + 2 and len(weight.shape) == sp + 2, (op, data.shape, weight.shape)
    st, di, pd = _ntuple(strides, sp), _ntuple(dilation, sp), _npad(padding, sp)
    o = int(weight.shape[0])
    assert weight.shape[1] * groups == data.shape[1], (op, data.shape, weight.shape, groups)
    spatial = [(data.shape[2 + i] + pd[i] + pd[sp + i] - di[i] * (weight.shape[2 + i] - 1) - 1) // st[i] + 1 for i in range(sp)]
    attrs = dict(strides=st, padding=pd, dilation=di, groups=int(groups), channels=int(channels if channels else o),
                 kernel_size=[int(k) for k in (kernel_size if kernel_size is not None else weight.shape[2:])],
                 data_layout=data_layout, kernel_layout=kernel_layout, out_layout=out_layout, out_dtype=out_dtype)
    return Call(op, [data, weight], attrs, [data.shape[0], o] + spatial, data.dtype)


def conv1d(data, weight, strides=1, padding=0, dilation=1, groups=1, channels=None, kernel_size=None, data_layout="NCW",
           kernel_layout="OIW", out_layout="", out_dtype=""):
    """relay.nn.conv1d (python/tvm/relay/op/nn/nn.py:30-117)."""
    return _conv_nd("nn.conv1d", 1, data, weight, strides, padding, dilation, groups, channels, kernel_size, data_layout,
                    kernel_layout, out_layout, out_dtype)


def conv2d(data, weight, strides=(1, 1), padding=(0, 0), dilation=(1, 1), groups=1, channels=None, kernel_size=None,
           data_layout="NCHW", kernel_layout="OIHW", out_layout="", out_dtype=""):
    """relay.nn.conv2d (nn.py:120-227)."""
    return _conv_nd("nn.conv2d", 2, data, weight, strides, padding, dilation, groups, channels, kernel_size, data_layout,
                    kernel_layout, out_layout, out_dtype)


def conv3d(data, weight, strides=(1, 1, 1), padding=(0, 0, 0), dilation=(1, 1, 1), groups=1, channels=None, kernel_size=None,
           data_layout="NCDHW", kernel_layout="OIDHW", out_layout="", out_dtype=""):
    """relay.nn.conv3d (nn.py:230-330)."""
    return _conv_nd("nn.conv3d", 3, data, weight, strides, padding, dilation, groups, channels, kernel_size, data_layout,
                    kernel_layout, out_layout, out_dtype)


def _conv_transpose_nd(op, sp, data, weight, strides, padding, dilation, groups, channels, kernel_size, data_layout,
                       kernel_layout, out_layout, output_padding, out_dtype):
    assert len(data.shape) == sp + 2 and len(weight.shape) == sp + 2
    st, di, pd, opad = _ntuple(strides, sp), _ntuple(dilation, sp), _npad(padding, sp), _ntuple(output_padding, sp)
    assert weight.shape[0] == data.shape[1], (op, data.shape, weight.shape)
    o = int(weight.shape[1]) * int(groups)
    spatial = [(data.shape[2 + i] - 1) * st[i] - pd[i] - pd[sp + i] + di[i] * (weight.shape[2 + i] - 1) + 1 + opad[i] for i in range(sp)]
    attrs = dict(strides=st, padding=pd, dilation=di, groups=int(groups), channels=int(channels if channels else o),
                 kernel_size=[int(k) for k in (kernel_size if kernel_size is not None else weight.shape[2:])],
                 data_layout=data_layout, kernel_layout=kernel_layout, out_layout=out_layout, output_padding=opad,
                 out_dtype=out_dtype)
    return Call(op, [data, weight], attrs, [data.shape[0], o] + spatial, data.dtype)


def conv2d_transpose(data, weight, strides=(1, 1), padding=(0, 0), dilation=(1, 1), groups=1, channels=None, kernel_size=None,
                     data_layout="NCHW", kernel_layout="IOHW", out_layout="", output_padding=(0, 0), out_dtype=""):
    """relay.nn.conv2d_transpose (nn.py:515-590): weight is [in, out / groups, kh, kw]."""
    return _conv_transpose_nd("nn.conv2d_transpose", 2, data, weight, strides, padding, dilation, groups, channels, kernel_size,
                              data_layout, kernel_layout, out_layout, output_padding, out_dtype)


def conv3d_transpose(data, weight, strides=(1, 1, 1), padding=(0, 0, 0), dilation=(1, 1, 1), groups=1, channels=None,
                     kernel_size=None, data_layout="NCDHW", kernel_layout="OIDHW", out_layout="", output_padding=(0, 0, 0),
                     out_dtype=""):
    """relay.nn.conv3d_transpose (nn.py:430-512): the layout string says OIDHW, the tensor is [in, out / groups, d, h, w]
    (the reference runtime's own work-around, tachikoma_json_runtime.cc:420-423)."""
    return _conv_transpose_nd("nn.conv3d_transpose", 3, data, weight, strides, padding, dilation, groups, channels, kernel_size,
                              data_layout, kernel_layout, out_layout, output_padding, out_dtype)


def dense(data, weight, units=None, out_dtype=""):
    """relay.nn.dense: weight is [units, in]."""
    assert data.shape[-1] == weight.shape[1], (data.shape, weight.shape)
    return Call("nn.dense", [data, weight], dict(units=int(units if units else weight.shape[0]), out_dtype=out_dtype),
                tuple(data.shape[:-1]) + (weight.shape[0],), data.dtype)


def batch_matmul(tensor_a, tensor_b, out_dtype="", transpose_a=False, transpose_b=True):
    """relay.nn.batch_matmul (nn.py:2226): NT by default."""
    m, k = (tensor_a.shape[2], tensor_a.shape[1]) if transpose_a else (tensor_a.shape[1], tensor_a.shape[2])
    n, k2 = (tensor_b.shape[1], tensor_b.shape[2]) if transpose_b else (tensor_b.shape[2], tensor_b.shape[1])
    assert k == k2, (tensor_a.shape, tensor_b.shape)
    return Call("nn.batch_matmul", [tensor_a, tensor_b], dict(out_dtype=out_dtype, transpose_a=bool(transpose_a),
                                                               transpose_b=bool(transpose_b)),
                (max(tensor_a.shape[0], tensor_b.shape[0]), m, n), tensor_a.dtype)


def batch_norm(data, gamma, beta, moving_mean, moving_var, axis=1, epsilon=1e-5, center=True, scale=True):
    """relay.nn.batch_norm(...)[0] (nn.py:1963): the normalised tensor.  The reference's operator returns a 3-tuple of which
    inference uses field 0 only; the JSON node still declares the three outputs (serialize())."""
    return Call("nn.batch_norm", [data, _c(gamma), _c(beta), _c(moving_mean), _c(moving_var)],
                dict(axis=int(axis), epsilon=float(epsilon), center=bool(center), scale=bool(scale)), data.shape, data.dtype)


def layer_norm(data, gamma, beta, axis=-1, epsilon=1e-5, center=True, scale=True):
    """relay.nn.layer_norm (nn.py:2112)."""
    return Call("nn.layer_norm", [data, _c(gamma), _c(beta)],
                dict(axis=int(axis), epsilon=float(epsilon), center=bool(center), scale=bool(scale)), data.shape, data.dtype)


def softmax(data, axis=-1):
    return Call("nn.softmax", [data], dict(axis=int(axis)), data.shape, data.dtype)


def _unary(op):
    def f(data):
        return Call(op, [data], {}, data.shape, data.dtype)
    f.__name__ = op.split(".")[-1]
    f.__doc__ = f"relay.{op}"
    return f


relu, tanh, sigmoid, abs_, exp, log, sqrt, round_, erf = (_unary(o) for o in (
    "nn.relu", "tanh", "sigmoid", "abs", "exp", "log", "sqrt", "round", "erf"))


def leaky_relu(data, alpha=0.01):
    return Call("nn.leaky_relu", [data], dict(alpha=float(alpha)), data.shape, data.dtype)


def _binary(op):
    def f(lhs, rhs):
        rhs = _c(rhs, lhs.dtype)
        return Call(op, [lhs, rhs], {}, np.broadcast_shapes(lhs.shape, rhs.shape), lhs.dtype)
    f.__name__ = op
    return f


multiply, divide = _binary("multiply"), _binary("divide")


def max_pool3d(data, pool_size=(1, 1, 1), strides=(1, 1, 1), dilation=(1, 1, 1), padding=(0, 0, 0), layout="NCDHW"):
    return _pool3d("nn.max_pool3d", data, pool_size, strides, dilation, padding, layout, {})


def avg_pool3d(data, pool_size=(1, 1, 1), strides=(1, 1, 1), dilation=(1, 1, 1), padding=(0, 0, 0), layout="NCDHW",
               count_include_pad=False):
    return _pool3d("nn.avg_pool3d", data, pool_size, strides, dilation, padding, layout, dict(count_include_pad=bool(count_include_pad)))


def _pool3d(op, data, pool_size, strides, dilation, padding, layout, extra):
    assert layout == "NCDHW" and len(data.shape) == 5
    ps, st, di, pd = _ntuple(pool_size, 3), _ntuple(strides, 3), _ntuple(dilation, 3), _npad(padding, 3)
    spatial = [(data.shape[2 + i] + pd[i] + pd[3 + i] - di[i] * (ps[i] - 1) - 1) // st[i] + 1 for i in range(3)]
    return Call(op, [data], dict(pool_size=ps, strides=st, dilation=di, padding=pd, layout=layout, out_layout="",
                                 ceil_mode=False, **extra), list(data.shape[:2]) + spatial, data.dtype)


# ----------------------------------------------------------------------------------------------
# (de)serialisation of a Function to plain python data -- used for golden fixtures and for saving
# model descriptions.  Not a reference format.
# ----------------------------------------------------------------------------------------------

def bind_params(func, params):
    """relay.build(..., params=) / bind_params_by_name: turn named vars into constants."""
    params = {k: np.asarray(v) for k, v in (params or {}).items()}
    memo = {}

    def visit(e):
        if id(e) in memo:
            return memo[id(e)]
        if isinstance(e, Var) and e.name in params:
            r = Constant(params[e.name].astype(e.dtype))
        elif isinstance(e, Call):
            r = Call(e.op, [visit(a) for a in e.args], e.attrs, e.shape, e.dtype)
        else:
            r = e
        memo[id(e)] = r
        return r

    body = visit(func.body)
    return Function([p for p in func.params if p.name not in params], body, func.attrs)


def free_vars(expr):
    return [e for e in post_order(expr) if isinstance(e, Var)]


def to_dict(func):
    nodes, index = [], {}
    for e in post_order(func.body):
        index[id(e)] = len(nodes)
        if isinstance(e, Var):
            nodes.append(dict(kind="var", name=e.name, shape=list(e.shape), dtype=e.dtype))
        elif isinstance(e, Constant):
            nodes.append(dict(kind="const", shape=list(e.shape), dtype=e.dtype,
                              data=e.data.reshape(-1).tolist()))
        else:
            nodes.append(dict(kind="call", op=e.op, args=[index[id(a)] for a in e.args], attrs=e.attrs,
                              shape=list(e.shape), dtype=e.dtype))
    for p in func.params:  # unused params still belong to the signature
        if id(p) not in index:
            index[id(p)] = len(nodes)
            nodes.append(dict(kind="var", name=p.name, shape=list(p.shape), dtype=p.dtype))
    return dict(nodes=nodes, params=[index[id(p)] for p in func.params], body=index[id(func.body)])


def from_dict(d):
    objs = []
    for n in d["nodes"]:
        if n["kind"] == "var":
            objs.append(Var(n["name"], n["shape"], n["dtype"]))
        elif n["kind"] == "const":
            objs.append(Constant(np.array(n["data"], dtype=n["dtype"]).reshape(n["shape"])))
        else:
            objs.append(Call(n["op"], [objs[i] for i in n["args"]], n["attrs"], n["shape"], n["dtype"]))
    return Function([objs[i] for i in d["params"]], objs[d["body"]])
