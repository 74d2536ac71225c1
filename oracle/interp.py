"""Oracle graph interpreter -- TEST INFRASTRUCTURE ONLY.

Evaluates (a) an un-partitioned ``tachikoma_b200.relay`` expression op by op, exactly like the
reference's un-offloaded "llvm" build does after QNN canonicalization, and
(b) a tachikoma JSON graph (composite nodes are evaluated by running the primitive oracle ops of
the chain one after the other).  All arithmetic is done by oracle/libqnn_oracle.so
(qnn_oracle.c); numpy is used for reshapes/transposes and trivial elementwise ops only.

May be imported only by tests/, __graft_entry__.smoke() and bench.py (cpu_baseline /
--impl reference).  The product package never imports this module.
"""
from __future__ import annotations

import ctypes
import json
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

DT = {"int8": 0, "uint8": 1, "int32": 2, "int16": 3, "int64": 4, "float32": 5}
ROUNDING = {"UPWARD": 0, "TONEAREST": 1, "None": 0}
COMPUTE = {"int64": 0, "float32": 1, "float64": 2, "None": 0}


def build(force=False):
    so = os.path.join(_HERE, "libqnn_oracle.so")
    src = os.path.join(_HERE, "qnn_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "clean", "all"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.tko_fixed_point_multiply.restype = ctypes.c_int32
        _LIB.tko_fixed_point_multiply.argtypes = [ctypes.c_int32] * 3
    return _LIB


def set_threads(n=0):
    """Set (n > 0) and return the OpenMP thread count the oracle kernels run with."""
    return int(lib().tko_set_threads(int(n)))


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _i64(v):
    return (ctypes.c_int64 * len(v))(*[int(x) for x in v])


def fixed_point_multiplier_shift(m):
    s, e = ctypes.c_int32(), ctypes.c_int32()
    lib().tko_fixed_point_multiplier_shift(ctypes.c_double(m), ctypes.byref(s), ctypes.byref(e))
    return s.value, e.value


def fixed_point_multiply(x, multiplier, shift):
    x = np.asarray(x, dtype=np.int32)
    f = lib().tko_fixed_point_multiply
    return np.array([f(int(v), int(multiplier), int(shift)) for v in x.ravel()], dtype=np.int32).reshape(x.shape)


def requantize(data, input_scale, input_zero_point, output_scale, output_zero_point, axis=-1,
               rounding="None", compute_dtype="None", out_dtype="int8"):
    data = np.asarray(data)
    shape0 = data.shape
    data = np.ascontiguousarray(data)
    sc = np.ascontiguousarray(np.asarray(input_scale, dtype=np.float32).reshape(-1))
    zp = np.ascontiguousarray(np.asarray(input_zero_point, dtype=np.int32).reshape(-1))
    nd = data.ndim
    ax = axis + nd if axis < 0 else axis
    if nd == 0:
        outer, C, inner = 1, 1, 1
    else:
        outer = int(np.prod(data.shape[:ax], dtype=np.int64))
        C = int(data.shape[ax])
        inner = int(np.prod(data.shape[ax + 1:], dtype=np.int64))
    assert sc.size in (1, C) and zp.size in (1, C)
    out = np.empty(shape0, dtype=out_dtype)
    lib().tko_requantize(_p(data), DT[str(data.dtype)], ctypes.c_int64(outer), ctypes.c_int64(C),
                         ctypes.c_int64(inner), _p(sc), int(sc.size), _p(zp), int(zp.size),
                         ctypes.c_float(float(np.asarray(output_scale, dtype=np.float32).reshape(-1)[0])),
                         ctypes.c_int32(int(np.asarray(output_zero_point).reshape(-1)[0])),
                         ROUNDING[rounding], COMPUTE[compute_dtype], DT[out_dtype], _p(out))
    return out


def _strides(shape, layout, want):
    st = [int(s) for s in (np.array(np.zeros(shape, dtype=np.int8).strides))]
    return [st[layout.index(ch)] for ch in want]


def conv2d(data, kernel, input_zero_point, kernel_zero_point, strides=(1, 1), padding=(0, 0, 0, 0),
           dilation=(1, 1), groups=1, data_layout="NCHW", kernel_layout="OIHW", out_layout=""):
    data, kernel = np.ascontiguousarray(data), np.ascontiguousarray(kernel)
    if kernel_layout == "HWOI":  # depthwise NHWC convention: [H, W, C, multiplier] -> O = C*mult
        kh, kw, c_, m_ = kernel.shape
        kernel = np.ascontiguousarray(kernel.transpose(2, 3, 0, 1).reshape(c_ * m_, 1, kh, kw))
        kernel_layout = "OIHW"
    elif kernel_layout == "HWIO" and groups > 1 and kernel.shape[2] * groups != data.shape[data_layout.index("C")]:
        # depthwise HWIO with I = in-channels, O = multiplier is not used by the reference tests
        raise NotImplementedError("grouped HWIO kernel with I != C/groups")
    N, C, H, W = [data.shape[data_layout.index(ch)] for ch in "NCHW"]
    O, I, KH, KW = [kernel.shape[kernel_layout.index(ch)] for ch in "OIHW"]
    assert I * groups == C, (I, groups, C)
    pad = list(padding)
    if len(pad) == 2:
        pad = [pad[0], pad[1], pad[0], pad[1]]
    sh, sw = strides
    dh, dw = dilation
    OH = (H + pad[0] + pad[2] - (KH - 1) * dh - 1) // sh + 1
    OW = (W + pad[1] + pad[3] - (KW - 1) * dw - 1) // sw + 1
    olay = out_layout or data_layout
    oshape = [0] * 4
    for ch, v in zip("NCHW", (N, O, OH, OW)):
        oshape[olay.index(ch)] = v
    out = np.empty(oshape, dtype=np.int32)
    zw = np.ascontiguousarray(np.asarray(kernel_zero_point, dtype=np.int32).reshape(-1))
    assert zw.size in (1, O)
    L = ctypes.c_int64
    lib().tko_conv2d(_p(data), DT[str(data.dtype)], _p(kernel), DT[str(kernel.dtype)], _p(out),
                     L(N), L(C), L(H), L(W), L(O), L(KH), L(KW), L(groups), L(sh), L(sw), L(pad[0]), L(pad[1]),
                     L(pad[2]), L(pad[3]), L(dh), L(dw), _i64(_strides(data.shape, data_layout, "NCHW")),
                     _i64(_strides(kernel.shape, kernel_layout, "OIHW")), _i64(_strides(oshape, olay, "NCHW")),
                     ctypes.c_int32(int(np.asarray(input_zero_point).reshape(-1)[0])), _p(zw), int(zw.size))
    return out


def dense(x, w, zp_x, zp_w):
    x, w = np.ascontiguousarray(x), np.ascontiguousarray(w)
    lead = x.shape[:-1]
    M, K, N = int(np.prod(lead, dtype=np.int64)), x.shape[-1], w.shape[0]
    out = np.empty((M, N), dtype=np.int32)
    L = ctypes.c_int64
    lib().tko_dense(_p(x), DT[str(x.dtype)], _p(w), DT[str(w.dtype)], _p(out), L(M), L(N), L(K),
                    ctypes.c_int32(int(np.asarray(zp_x).reshape(-1)[0])),
                    ctypes.c_int32(int(np.asarray(zp_w).reshape(-1)[0])))
    return out.reshape(tuple(lead) + (N,))


def batch_matmul(x, y, zp_x, zp_y):
    x, y = np.ascontiguousarray(x), np.ascontiguousarray(y)
    B, M, K = x.shape
    By, N, _ = y.shape
    out = np.empty((B, M, N), dtype=np.int32)
    L = ctypes.c_int64
    lib().tko_batch_matmul(_p(x), DT[str(x.dtype)], _p(y), DT[str(y.dtype)], _p(out), L(B), L(M), L(N), L(K),
                           L(By), ctypes.c_int32(int(np.asarray(zp_x).reshape(-1)[0])),
                           ctypes.c_int32(int(np.asarray(zp_y).reshape(-1)[0])))
    return out


def qnn_add(a, b, sa, zpa, sb, zpb, so, zpo):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    assert a.shape == b.shape and a.dtype == b.dtype
    out = np.empty_like(a)
    f = lambda v: ctypes.c_float(float(np.float32(np.asarray(v).reshape(-1)[0])))
    z = lambda v: ctypes.c_int32(int(np.asarray(v).reshape(-1)[0]))
    lib().tko_qnn_add(_p(a), _p(b), DT[str(a.dtype)], ctypes.c_int64(a.size), f(sa), z(zpa), f(sb), z(zpb),
                      f(so), z(zpo), _p(out))
    return out


def qnn_subtract(a, b, sa, zpa, sb, zpb, so, zpo):
    """src/relay/qnn/op/subtract.cc:40-93."""
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    assert a.shape == b.shape and a.dtype == b.dtype
    out = np.empty_like(a)
    f = lambda v: ctypes.c_float(float(np.float32(np.asarray(v).reshape(-1)[0])))
    z = lambda v: ctypes.c_int32(int(np.asarray(v).reshape(-1)[0]))
    lib().tko_qnn_subtract(_p(a), _p(b), DT[str(a.dtype)], ctypes.c_int64(a.size), f(sa), z(zpa), f(sb), z(zpb),
                           f(so), z(zpo), _p(out))
    return out


def _sc(v):
    return np.float32(np.asarray(v, dtype=np.float32).reshape(-1)[0])


def _zp(v):
    return int(np.asarray(v).reshape(-1)[0])


def qnn_mul(a, b, sa, zpa, sb, zpb, so, zpo):
    """src/relay/qnn/op/mul.cc:66-101 (per-tensor form): Q' = (int32(a) - zp_a) * (int32(b) - zp_b) with int32
    wrap-around, then Requantize(Q', scale = float32(sa) * float32(sb), zp 0 -> output qparams, dtype of the inputs)."""
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    with np.errstate(over="ignore"):
        qa = a.astype(np.int32) - np.int32(_zp(zpa))
        qb = b.astype(np.int32) - np.int32(_zp(zpb))
        prod = (qa.astype(np.int64) * qb.astype(np.int64)).astype(np.int32)  # two's-complement wrap of the int32 multiply
    new_scale = np.float32(_sc(sa) * _sc(sb))
    return requantize(prod, new_scale, 0, _sc(so), _zp(zpo), -1, "None", "None", str(a.dtype))


def qnn_concatenate(tensors, scales, zps, so, zpo, axis):
    """src/relay/qnn/op/concatenate.cc:138-213: requantize every input whose qparams differ from the output's
    (IsEqualScalar on the float32 scale and the int32 zero point), keep its dtype, then concatenate."""
    outs = []
    for t, s_, z_ in zip(tensors, scales, zps):
        t = np.ascontiguousarray(t)
        if _sc(s_) != _sc(so) or _zp(z_) != _zp(zpo):
            t = requantize(t, _sc(s_), _zp(z_), _sc(so), _zp(zpo), -1, "None", "None", str(t.dtype))
        outs.append(t)
    return np.ascontiguousarray(np.concatenate(outs, axis=axis))


def qnn_leaky_relu(x, alpha, s_in, zp_in, s_out, zp_out):
    """src/relay/qnn/op/leaky_relu.cc:70-139: Q' = RequantizeOrUpcast(int32(x)) to the output qparams
    (op_common.h:196-210); negative side (x < zp_in): FixedPointMultiply(Q', alpha) + FixedPointMultiply(zp_out, 1 - alpha)
    (src/relay/qnn/utils.cc FixedPointMultiply -> fixed_point_multiply); ConvertDtype saturates to the input dtype."""
    x = np.ascontiguousarray(x)
    if x.size > 4096:  # elementwise: evaluate the distinct values once (the scalar fixed_point_multiply is a Python loop)
        vals, inv = np.unique(x, return_inverse=True)
        return qnn_leaky_relu(vals, alpha, s_in, zp_in, s_out, zp_out)[inv].reshape(x.shape)
    x32 = x.astype(np.int32)
    if _sc(s_in) != _sc(s_out) or _zp(zp_in) != _zp(zp_out):
        q = requantize(x32, _sc(s_in), _zp(zp_in), _sc(s_out), _zp(zp_out), -1, "None", "None", "int32")
    else:
        q = x32
    ma, sa_ = fixed_point_multiplier_shift(float(alpha))
    mz, sz = fixed_point_multiplier_shift(1.0 - float(alpha))
    prod = fixed_point_multiply(q, ma, sa_)
    scaled_z = fixed_point_multiply(np.asarray([_zp(zp_out)], dtype=np.int32), mz, sz)[0]
    neg = _wrap_add_i32(np.asarray(prod, dtype=np.int32), np.full(q.shape, scaled_z, dtype=np.int32))
    out = np.where(x32 < np.int32(_zp(zp_in)), neg, q)
    info = np.iinfo(x.dtype)
    return np.clip(out, info.min, info.max).astype(x.dtype)


_UNARY = None


def _unary_funcs():
    """python/tvm/relay/qnn/op/legalizations.py:74-86."""
    global _UNARY
    if _UNARY is None:
        from scipy import special

        def hardswish(x):
            return x * np.clip(x + 3.0, 0.0, 6.0) / 6.0
        _UNARY = {"qnn.sqrt": np.sqrt, "qnn.rsqrt": lambda a: 1 / np.sqrt(a), "qnn.exp": np.exp, "qnn.erf": special.erf,
                  "qnn.sigmoid": lambda a: 1 / (1 + np.exp(-a)), "qnn.hardswish": hardswish, "qnn.tanh": np.tanh,
                  "qnn.log": np.log, "qnn.abs": np.abs}
    return _UNARY


def qnn_unary(op, x, s_in, zp_in, s_out, zp_out):
    """legalizations.py:54-71 + canonicalizations.py:31-160: the operator IS quantize(f(dequantize(.))) tabulated over
    all 256 bit patterns of the input dtype; evaluated here element by element (no table) with the oracle's own
    dequantize / quantize, f = the numpy function the reference registers, applied to the float32 dequantized values."""
    x = np.ascontiguousarray(x)
    with np.errstate(all="ignore"):
        d = dequantize(x, _sc(s_in), _zp(zp_in))
        f = np.asarray(_unary_funcs()[op](d), dtype=np.float32)
    return quantize(f, _sc(s_out), _zp(zp_out), -1, str(x.dtype))


def _axis_split(shape, axis):
    nd = len(shape)
    if nd == 0:
        return 1, 1, 1
    ax = axis + nd if axis < 0 else axis
    return int(np.prod(shape[:ax], dtype=np.int64)), int(shape[ax]), int(np.prod(shape[ax + 1:], dtype=np.int64))


def quantize(data, scale, zp, axis=-1, out_dtype="int8"):
    shape0 = np.asarray(data).shape  # ascontiguousarray turns rank 0 into rank 1
    data = np.ascontiguousarray(data, dtype=np.float32).reshape(shape0)
    sc = np.ascontiguousarray(np.asarray(scale, dtype=np.float32).reshape(-1))
    z = np.ascontiguousarray(np.asarray(zp, dtype=np.int32).reshape(-1))
    outer, C, inner = _axis_split(data.shape, axis)
    if sc.size == 1 and z.size == 1:
        outer, C, inner = outer * C * inner, 1, 1
    out = np.empty(data.shape, dtype=out_dtype)
    L = ctypes.c_int64
    lib().tko_quantize(_p(data), L(outer), L(C), L(inner), _p(sc), _p(z), int(sc.size), int(z.size), DT[out_dtype], _p(out))
    return out


def dequantize(data, scale, zp, axis=-1):
    shape0 = np.asarray(data).shape
    data = np.ascontiguousarray(data).reshape(shape0)
    sc = np.ascontiguousarray(np.asarray(scale, dtype=np.float32).reshape(-1))
    z = np.ascontiguousarray(np.asarray(zp, dtype=np.int32).reshape(-1))
    outer, C, inner = _axis_split(data.shape, axis)
    if sc.size == 1 and z.size == 1:
        outer, C, inner = outer * C * inner, 1, 1
    out = np.empty(data.shape, dtype=np.float32)
    L = ctypes.c_int64
    lib().tko_dequantize(_p(data), DT[str(data.dtype)], L(outer), L(C), L(inner), _p(sc), _p(z), int(sc.size), int(z.size), _p(out))
    return out


def pool2d(data, pool_size, strides, padding, layout="NCHW", is_avg=False, count_include_pad=False):
    data = np.ascontiguousarray(data)
    N, C, H, W = [data.shape[layout.index(ch)] for ch in "NCHW"]
    pad = list(padding)
    if len(pad) == 2:
        pad = [pad[0], pad[1], pad[0], pad[1]]
    OH = (H + pad[0] + pad[2] - pool_size[0]) // strides[0] + 1
    OW = (W + pad[1] + pad[3] - pool_size[1]) // strides[1] + 1
    oshape = [0] * 4
    for ch, v in zip("NCHW", (N, C, OH, OW)):
        oshape[layout.index(ch)] = v
    out = np.empty(oshape, dtype=data.dtype)
    L = ctypes.c_int64
    lib().tko_pool2d(_p(data), DT[str(data.dtype)], _p(out), int(is_avg), int(count_include_pad), L(N), L(C),
                     L(H), L(W), L(pool_size[0]), L(pool_size[1]), L(strides[0]), L(strides[1]), L(pad[0]),
                     L(pad[1]), L(pad[2]), L(pad[3]), _i64(_strides(data.shape, layout, "NCHW")),
                     _i64(_strides(oshape, layout, "NCHW")))
    return out


def dnnl_float_epilogue(acc, axis, bias_f, o_scl, act_scl, sum_t, sum_scl, dst_zp, out_dtype):
    acc = np.ascontiguousarray(acc, dtype=np.int32)
    outer = int(np.prod(acc.shape[:axis], dtype=np.int64))
    C = acc.shape[axis]
    inner = int(np.prod(acc.shape[axis + 1:], dtype=np.int64))
    out = np.empty(acc.shape, dtype=out_dtype)
    o_scl = np.ascontiguousarray(np.asarray(o_scl, dtype=np.float32).reshape(-1))
    bias_p = None
    if bias_f is not None:
        bias_f = np.ascontiguousarray(np.broadcast_to(np.asarray(bias_f, dtype=np.float32).reshape(-1), (C,)))
        bias_p = _p(bias_f)
    sum_p, sum_dt = None, 0
    if sum_t is not None:
        sum_t = np.ascontiguousarray(sum_t)
        sum_p, sum_dt = _p(sum_t), DT[str(sum_t.dtype)]
    L = ctypes.c_int64
    lib().tko_dnnl_float_epilogue(_p(acc), L(outer), L(C), L(inner), bias_p, _p(o_scl), int(o_scl.size),
                                  ctypes.c_float(float(act_scl)), sum_p, sum_dt, ctypes.c_float(float(sum_scl)),
                                  ctypes.c_float(float(dst_zp)), DT[out_dtype], _p(out))
    return out


# ----------------------------------------------------------------------------------------------
# primitive-op dispatch shared by the expression and JSON interpreters
# ----------------------------------------------------------------------------------------------

def _wrap_add_i32(a, b):
    return (a.astype(np.int64) + b.astype(np.int64)).astype(np.int32) if a.dtype == np.int32 else a + b


def eval_op(op, attrs, args):
    """Evaluate one Relay operator on numpy inputs (constants included in ``args``)."""
    # float32 half of the tachikoma op list (oracle/float_ops.py); integer tensors never take these branches
    if not op.startswith("qnn.") and args and all(np.asarray(a).dtype == np.float32 for a in args):
        from . import float_ops as F
        if op in F.FLOAT_OPS:
            return F.eval_float_op(op, attrs, args)
        if op == "add":
            return F.float_add(args[0], args[1])
        if op == "nn.bias_add":
            shp = [1] * np.asarray(args[0]).ndim
            shp[attrs.get("axis", 1)] = -1
            return F.float_add(args[0], np.asarray(args[1]).reshape(shp))
        if op in ("nn.max_pool2d", "nn.avg_pool2d", "nn.global_avg_pool2d"):
            return F.float_pool(op, attrs, args[0])
    if op == "qnn.conv2d":
        return conv2d(args[0], args[1], args[2], args[3], attrs["strides"], attrs["padding"], attrs["dilation"],
                      attrs["groups"], attrs["data_layout"], attrs["kernel_layout"], attrs.get("out_layout", ""))
    if op == "qnn.dense":
        return dense(args[0], args[1], args[2], args[3])
    if op == "qnn.batch_matmul":
        return batch_matmul(args[0], args[1], args[2], args[3])
    if op == "qnn.requantize":
        return requantize(args[0], args[1], args[2], args[3], int(np.asarray(args[4]).reshape(-1)[0]),
                          attrs.get("axis", -1), attrs.get("rounding", "None"), attrs.get("compute_dtype", "None"),
                          attrs["out_dtype"])
    if op == "qnn.add":
        return qnn_add(*args[:8])
    if op == "qnn.mul":
        return qnn_mul(*args[:8])
    if op == "qnn.subtract":
        return qnn_subtract(*args[:8])
    if op == "qnn.concatenate":
        n = int(attrs["num_inputs"])
        return qnn_concatenate(args[:n], args[n:2 * n], args[2 * n:3 * n], args[3 * n], args[3 * n + 1], int(attrs["axis"]))
    if op == "qnn.quantize":
        return quantize(args[0], args[1], args[2], int(attrs.get("axis", -1)), attrs["out_dtype"])
    if op == "qnn.dequantize":
        return dequantize(args[0], args[1], args[2], int(attrs.get("axis", -1)))
    if op == "qnn.leaky_relu":
        return qnn_leaky_relu(args[0], float(attrs["alpha"]), *args[1:5])
    if op in _unary_funcs():
        return qnn_unary(op, *args[:5])
    if op == "reinterpret":
        return np.ascontiguousarray(args[0]).view(attrs["dtype"])
    if op == "take":  # 1-D table, axis 0, mode "fast" (indices are in range by construction)
        return np.ascontiguousarray(np.asarray(args[0])[np.asarray(args[1]).astype(np.int64)])
    if op == "nn.bias_add":
        ax = attrs.get("axis", 1)
        shp = [1] * args[0].ndim
        shp[ax] = -1
        return _wrap_add_i32(args[0], np.asarray(args[1]).reshape(shp).astype(args[0].dtype))
    if op == "add":
        b = np.broadcast_to(np.asarray(args[1]), np.broadcast_shapes(args[0].shape, np.asarray(args[1]).shape))
        return _wrap_add_i32(np.broadcast_to(args[0], b.shape), b.astype(args[0].dtype))
    if op == "clip":
        x = args[0]
        if x.dtype.kind in "iu":  # relay clip on integers compares in the tensor dtype
            lo = max(int(attrs["a_min"]), int(np.iinfo(x.dtype).min))
            hi = min(int(attrs["a_max"]), int(np.iinfo(x.dtype).max))
            return np.clip(x, lo, hi).astype(x.dtype)
        return np.clip(x, attrs["a_min"], attrs["a_max"]).astype(x.dtype)
    if op == "cast":
        return args[0].astype(attrs["dtype"])  # truncating two's-complement cast, as LLVM trunc
    if op in ("nn.max_pool2d", "nn.avg_pool2d"):
        return pool2d(args[0], attrs["pool_size"], attrs["strides"], attrs["padding"], attrs["layout"],
                      op == "nn.avg_pool2d", attrs.get("count_include_pad", False))
    if op == "nn.global_avg_pool2d":
        lay = attrs["layout"]
        hw = [args[0].shape[lay.index("H")], args[0].shape[lay.index("W")]]
        return pool2d(args[0], hw, [1, 1], [0, 0, 0, 0], lay, 2, False)  # 2: adaptive form, divisor in the tensor dtype
    if op == "nn.pad":
        pv = np.asarray(args[1]).reshape(-1)[0]
        return np.pad(args[0], attrs["pad_width"], mode="constant", constant_values=args[0].dtype.type(pv))
    if op == "reshape":
        return args[0].reshape(attrs["newshape"])
    if op == "nn.batch_flatten":
        return args[0].reshape(args[0].shape[0], -1)
    if op == "transpose":
        return np.ascontiguousarray(args[0].transpose(attrs["axes"]))
    raise NotImplementedError(op)


def run_expr(func, inputs):
    """Evaluate an un-partitioned relay Function; returns (output, {node: value})."""
    from tachikoma_b200 import relay as R

    env = {}
    for p in func.params:
        env[id(p)] = np.ascontiguousarray(inputs[p.name]).astype(p.dtype, copy=False).reshape(p.shape)
    for e in R.post_order(func.body):
        if isinstance(e, R.Var):
            assert id(e) in env, f"unbound var {e.name}"
        elif isinstance(e, R.Constant):
            env[id(e)] = e.data
        elif e.op == "tuple":  # several outputs: the function returns the list of its fields
            env[id(e)] = [env[id(a)] for a in e.args]
        else:
            env[id(e)] = eval_op(e.op, e.attrs, [env[id(a)] for a in e.args])
            assert tuple(env[id(e)].shape) == tuple(e.shape), (e.op, env[id(e)].shape, e.shape)
            assert str(env[id(e)].dtype) == e.dtype, (e.op, env[id(e)].dtype, e.dtype)
    return env[id(func.body)], env


# ----------------------------------------------------------------------------------------------
# tachikoma JSON graph interpreter (composites evaluated as their primitive chain) + trace writer
# ----------------------------------------------------------------------------------------------

def _attr(node, key, default=None):
    v = node.get("attrs", {}).get(key)
    return default if v is None else v[0]


def _ints(node, key):
    return [int(x) for x in _attr(node, key, [])]


def _conv_attrs(node):
    return dict(strides=_ints(node, "strides"), padding=_ints(node, "padding"), dilation=_ints(node, "dilation"),
                groups=int(_attr(node, "groups", ["1"])[0]), data_layout=_attr(node, "data_layout", ["NCHW"])[0],
                kernel_layout=_attr(node, "kernel_layout", ["OIHW"])[0], out_layout=_attr(node, "out_layout", [""])[0])


def _eval_composite(node, ins):
    """tachikoma.qnn.conv2d / tachikoma.qnn.dense: run the fused chain op by op."""
    is_conv = node["name"].endswith("conv2d")

    def arg(key):
        v = _attr(node, key)
        return None if v is None else ins[int(v[0])]

    if _attr(node, "rq_in_scl_idx") is not None:  # integer mode (path A)
        if is_conv:
            acc = eval_op("qnn.conv2d", _conv_attrs(node), [ins[0], ins[1], arg("src_zp_idx"), arg("wgh_zp_idx")])
            ch_axis = (_conv_attrs(node)["out_layout"] or _conv_attrs(node)["data_layout"]).index("C")
        else:
            acc = dense(ins[0], ins[1], arg("src_zp_idx"), arg("wgh_zp_idx"))
            ch_axis = acc.ndim - 1
        bias = arg("bias_idx")
        if bias is not None:
            shp = [1] * acc.ndim
            shp[ch_axis] = -1
            acc = _wrap_add_i32(acc, np.asarray(bias).reshape(shp).astype(np.int32))
        y = requantize(acc, arg("rq_in_scl_idx"), arg("rq_in_zp_idx"), arg("rq_out_scl_idx"),
                       int(np.asarray(arg("rq_out_zp_idx")).reshape(-1)[0]), int(_attr(node, "rq_axis", ["-1"])[0]),
                       _attr(node, "rounding", ["None"])[0], _attr(node, "compute_dtype", ["None"])[0],
                       _attr(node, "rq_out_dtype", ["int32"])[0])
        y = eval_op("clip", dict(a_min=float(_attr(node, "clip_a_min")[0]), a_max=float(_attr(node, "clip_a_max")[0])), [y])
        y = y.astype(_attr(node, "cast_dtype")[0])
        s = arg("sum_idx")
        if s is not None:
            y = qnn_add(y, s, arg("sum_lhs_scl_idx"), arg("sum_lhs_zp_idx"), arg("sum_rhs_scl_idx"), arg("sum_rhs_zp_idx"),
                        arg("sum_out_scl_idx"), arg("sum_out_zp_idx"))
            y = eval_op("clip", dict(a_min=float(_attr(node, "sum_clip_a_min")[0]),
                                     a_max=float(_attr(node, "sum_clip_a_max")[0])), [y])
        return y
    # float mode (path B): zero zero-points, float post-ops
    if is_conv:
        acc = eval_op("qnn.conv2d", _conv_attrs(node), [ins[0], ins[1], np.int32(0), np.int32(0)])
        ch_axis = (_conv_attrs(node)["out_layout"] or _conv_attrs(node)["data_layout"]).index("C")
    else:
        acc = dense(ins[0], ins[1], 0, 0)
        ch_axis = acc.ndim - 1
    act = _attr(node, "activation")
    act_scl = float(np.asarray(ins[int(act[1])]).reshape(-1)[0]) if act else 1.0
    sum_scl = arg("sum_scl_idx")
    dst_zp = arg("dst_zp_idx")
    out_dtype = node["attrs"]["dtype"][0][0]
    return dnnl_float_epilogue(acc, ch_axis, arg("bias_idx"), arg("o_scl_idx"), act_scl, arg("sum_idx"),
                               0.0 if sum_scl is None else float(np.asarray(sum_scl).reshape(-1)[0]),
                               0.0 if dst_zp is None else float(np.asarray(dst_zp).reshape(-1)[0]), out_dtype)


def _plain_attrs(node):
    a = {}
    for k, v in node.get("attrs", {}).items():
        if k in ("shape", "dtype", "num_inputs", "num_outputs"):
            continue
        a[k] = v[0]
    return a


_FLOAT_ATTR_LISTS = ("strides", "padding", "dilation", "pool_size", "output_padding", "kernel_size")
_FLOAT_ATTR_INTS = ("groups", "axis", "channels", "units", "transpose_a", "transpose_b", "count_include_pad", "ceil_mode",
                    "center", "scale")
_FLOAT_ATTR_FLOATS = ("epsilon", "alpha", "a_min", "a_max")


def _float_attrs(node):
    out = {}
    for k, v in _plain_attrs(node).items():
        if k in _FLOAT_ATTR_LISTS:
            out[k] = [int(x) for x in v]
        elif k in _FLOAT_ATTR_INTS:
            out[k] = int(v[0])
        elif k in _FLOAT_ATTR_FLOATS:
            out[k] = float(v[0])
        elif len(v) == 1:
            out[k] = v[0]
    return out


def _eval_json_node(node, ins):
    op = node["name"]
    if op.startswith("tachikoma.qnn."):
        return _eval_composite(node, ins)
    if ins and all(np.asarray(a).dtype == np.float32 for a in ins) and not op.startswith("qnn.") and op not in ("clip", "cast", "reshape", "transpose", "nn.pad", "nn.batch_flatten"):
        from . import float_ops as F
        if op.startswith("tachikoma."):
            return F.eval_float_composite(op, _float_attrs(node), ins)
        return eval_op(op, _float_attrs(node), ins)
    a = _plain_attrs(node)
    g = lambda k, d=None: a[k][0] if k in a else d
    if op == "qnn.conv2d":
        return eval_op(op, _conv_attrs(node), ins)
    if op == "qnn.requantize":
        return eval_op(op, dict(axis=int(g("axis", -1)), rounding=g("rounding", "None"),
                                compute_dtype=g("compute_dtype", "None"), out_dtype=g("out_dtype")), ins)
    if op == "clip":
        return eval_op(op, dict(a_min=float(g("a_min")), a_max=float(g("a_max"))), ins)
    if op == "cast":
        # the cast's own "dtype" attribute is shadowed by the node's type list (codegen_json.h:228-231)
        return eval_op(op, dict(dtype=node["attrs"]["dtype"][0][0]), ins)
    if op == "nn.bias_add":
        return eval_op(op, dict(axis=int(g("axis", 1))), ins)
    if op in ("nn.max_pool2d", "nn.avg_pool2d"):
        return eval_op(op, dict(pool_size=[int(x) for x in a["pool_size"]], strides=[int(x) for x in a["strides"]],
                                padding=[int(x) for x in a["padding"]], layout=g("layout", "NCHW"),
                                count_include_pad=bool(int(g("count_include_pad", 0)))), ins)
    if op == "nn.global_avg_pool2d":
        return eval_op(op, dict(layout=g("layout", "NCHW")), ins)
    if op == "nn.pad":
        pw = [int(x) for x in a["pad_width"]]
        return eval_op(op, dict(pad_width=[[pw[2 * i], pw[2 * i + 1]] for i in range(len(pw) // 2)]), ins)
    if op == "reshape":
        return ins[0].reshape(node["attrs"]["shape"][0][0])
    if op == "transpose":
        return eval_op(op, dict(axes=[int(x) for x in a["axes"]]), ins)
    if op == "qnn.concatenate":
        return eval_op(op, dict(axis=int(g("axis")), num_inputs=(len(ins) - 2) // 3), ins)
    if op == "qnn.quantize":
        return eval_op(op, dict(axis=int(g("axis", -1)), out_dtype=node["attrs"]["dtype"][0][0]), ins)
    if op == "qnn.dequantize":
        return eval_op(op, dict(axis=int(g("axis", -1))), ins)
    if op == "qnn.leaky_relu":
        return eval_op(op, dict(alpha=float(g("alpha"))), ins)
    if op == "reinterpret":
        return eval_op(op, dict(dtype=node["attrs"]["dtype"][0][0]), ins)
    return eval_op(op, {}, ins)  # qnn.dense, qnn.batch_matmul, qnn.add, qnn.mul, add, nn.batch_flatten


def run_json(graph_json, consts, inputs, want_trace=False):
    """Evaluate a tachikoma JSON graph.  ``consts``: arrays in const order; ``inputs``: arrays in graph
    input order.  Returns (outputs, trace) with trace = ordered list of (name, array) records in the
    runtime's naming (``{op}____topo-index:{nid}____input-num:{i}`` / ``____output-num:0``)."""
    g = json.loads(graph_json)
    nodes = g["nodes"]
    vals = {}
    ci = ii = 0
    for nid in g["arg_nodes"]:
        n = nodes[nid]
        shape, dtype = n["attrs"]["shape"][0][0], n["attrs"]["dtype"][0][0]
        if n["op"] == "const":
            vals[(nid, 0)] = np.ascontiguousarray(consts[ci]).astype(dtype).reshape(shape)
            ci += 1
        else:
            vals[(nid, 0)] = np.ascontiguousarray(inputs[ii]).astype(dtype).reshape(shape)
            ii += 1
    trace = []
    for nid, n in enumerate(nodes):
        if n["op"] != "kernel":
            continue
        ins = [vals[(e[0], e[1])] for e in n["inputs"]]
        out = _eval_json_node(n, ins)
        shape, dtype = n["attrs"]["shape"][0][0], n["attrs"]["dtype"][0][0]
        assert list(out.shape) == list(shape) and str(out.dtype) == dtype, (n["name"], out.shape, shape, out.dtype, dtype)
        vals[(nid, 0)] = np.ascontiguousarray(out)
        if want_trace:
            for i, a in enumerate(ins):
                trace.append((f"{n['name']}____topo-index:{nid}____input-num:{i}", a))
            trace.append((f"{n['name']}____topo-index:{nid}____output-num:0", vals[(nid, 0)]))
    outs = [vals[(h[0], h[1])] for h in g["heads"]]
    return outs, trace


_DL = {"int8": (0, 8), "uint8": (1, 8), "int16": (0, 16), "int32": (0, 32), "int64": (0, 64), "float32": (2, 32)}


def save_ndarray_list(records):
    """TVM NDArray-list blob (src/runtime/file_utils.cc:210-229 SaveParams; per-array layout
    include/tvm/runtime/ndarray.h:449-494 SaveDLTensor), written independently of the CUDA library."""
    import struct

    out = bytearray()
    out += struct.pack("<QQ", 0xF7E58D4F05049CB7, 0)
    out += struct.pack("<Q", len(records))
    for name, _ in records:
        b = name.encode()
        out += struct.pack("<Q", len(b)) + b
    out += struct.pack("<Q", len(records))
    for _, arr in records:
        arr = np.asarray(arr)
        shape = arr.shape
        arr = np.ascontiguousarray(arr)
        code, bits = _DL[str(arr.dtype)]
        out += struct.pack("<QQ", 0xDD5E40F096B4A13F, 0)
        out += struct.pack("<iii", 1, 0, len(shape))
        out += struct.pack("<BBH", code, bits, 1)
        for d in shape:
            out += struct.pack("<q", d)
        out += struct.pack("<q", arr.nbytes)
        out += arr.tobytes()
    return bytes(out)


def load_ndarray_list(blob):
    """Inverse of save_ndarray_list (src/runtime/file_utils.cc:184-208 LoadParams)."""
    import struct

    pos = 0
    magic, _ = struct.unpack_from("<QQ", blob, pos); pos += 16
    assert magic == 0xF7E58D4F05049CB7
    (n,) = struct.unpack_from("<Q", blob, pos); pos += 8
    names = []
    for _ in range(n):
        (ln,) = struct.unpack_from("<Q", blob, pos); pos += 8
        names.append(blob[pos:pos + ln].decode()); pos += ln
    (m,) = struct.unpack_from("<Q", blob, pos); pos += 8
    assert m == n
    inv = {v: k for k, v in _DL.items()}
    recs = []
    for i in range(n):
        magic, _ = struct.unpack_from("<QQ", blob, pos); pos += 16
        assert magic == 0xDD5E40F096B4A13F
        _dev, _id, ndim = struct.unpack_from("<iii", blob, pos); pos += 12
        code, bits, _lanes = struct.unpack_from("<BBH", blob, pos); pos += 4
        shape = struct.unpack_from("<" + "q" * ndim, blob, pos); pos += 8 * ndim
        (nbytes,) = struct.unpack_from("<q", blob, pos); pos += 8
        arr = np.frombuffer(blob, dtype=inv[(code, bits)], count=nbytes // (bits // 8), offset=pos).reshape(shape)
        pos += nbytes
        recs.append((names[i], arr))
    return recs
