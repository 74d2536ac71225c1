"""Synthetic integer-only QNN models for the BASELINE.json configs (random-init quantized weights).

The graphs have exactly the operator chain the reference's PyTorch frontend emits for prequantized
models (python/tvm/relay/frontend/qnn_torch.py:649-682): qnn.conv2d / qnn.dense (uint8 activations,
int8 weights, per-output-channel weight scales) -> nn.bias_add (int32) -> qnn.requantize(axis=1,
int32) -> clip -> cast(uint8); residual joins are qnn.add -> clip (quantized::add_relu); average
pooling is done on int32 (apply_with_upcast, qnn_torch.py:587-590).  Explicit zero-point padding is
expressed through the convolution's padding attribute, which is what the reference's SimplifyExpr
(FoldExplicitPadding) leaves after partition_for_tachikoma (tests/python/contrib/test_tachikoma.py:88).

Output scales are calibrated once on one seeded sample with a float simulation (numpy float32 convolution used
as a calculator only) and rounded to 4 significant digits, so that activations keep spanning the
uint8 range through the whole depth instead of collapsing to the zero point.
"""
from __future__ import annotations

import numpy as np

from . import relay as R


def _round_sig(x, sig=4):
    x = float(x)
    if x == 0:
        return 0.0
    from math import floor, log10
    return round(x, sig - int(floor(log10(abs(x)))) - 1)


def _std(a):
    """Sample standard deviation (n - 1) in float64 of a float32 array."""
    return float(np.std(a.astype(np.float64), ddof=1))


def _windows(x, k, stride, pad, fill):
    """[N, C, OH, OW, k, k] strided view of the (padded) NCHW array."""
    if pad:
        x = np.pad(x, ((0, 0), (0, 0), (pad, pad), (pad, pad)), constant_values=fill)
    v = np.lib.stride_tricks.sliding_window_view(x, (k, k), axis=(2, 3))
    return v[:, :, ::stride, ::stride]


def _conv2d_f32(x, w, stride, pad, groups):
    """float32 NCHW / OIHW convolution for calibration (numpy only: the package uses torch for buffers, never for math)."""
    n, c = x.shape[:2]
    o, cg, kh, kw = w.shape
    v = _windows(x, kh, stride, pad, 0.0)                       # [N, C, OH, OW, kh, kw]
    oh, ow = v.shape[2:4]
    if groups == c and cg == 1:                                  # depthwise (multiplier o // c)
        m = o // c
        out = np.einsum("nchwrs,cmrs->ncmhw", v, w.reshape(c, m, kh, kw), optimize=True)
        return out.reshape(n, o, oh, ow).astype(np.float32)
    og = o // groups
    out = np.empty((n, o, oh, ow), np.float32)
    for g in range(groups):
        cols = v[:, g * cg:(g + 1) * cg].transpose(0, 2, 3, 1, 4, 5).reshape(n * oh * ow, cg * kh * kw)
        res = cols @ w[g * og:(g + 1) * og].reshape(og, -1).T     # [N*OH*OW, og]
        out[:, g * og:(g + 1) * og] = res.reshape(n, oh, ow, og).transpose(0, 3, 1, 2)
    return out


def _pool2d_f32(x, k, stride, pad, kind):
    v = _windows(x, k, stride, pad, -np.inf if kind == "max" else 0.0)
    return (v.max(axis=(4, 5)) if kind == "max" else v.mean(axis=(4, 5), dtype=np.float32)).astype(np.float32)


class _Builder:
    """Tracks (expr, scale, zero_point) and a float shadow activation used for calibration."""

    def __init__(self, seed, calib_input, in_scale, in_zp):
        self.rng = np.random.default_rng(seed)
        self.n_layers = 0
        self.shadow = calib_input.astype(np.float32)  # quantized values as floats

    def weights(self, shape, k):
        w = self.rng.integers(-127, 128, size=shape).astype(np.int8)
        s_w = (self.rng.uniform(0.5, 1.5, size=shape[0]) / (127.0 * np.sqrt(k))).astype(np.float32)
        b = self.rng.integers(-(1 << 12), 1 << 12, size=shape[0]).astype(np.int32)
        return w, s_w, b

    def _calibrate(self, acc, s_x, s_w, relu, ch_axis):
        """acc: float tensor of int32 accumulators (+bias).  Returns (s_out, zp_out, next shadow)."""
        shp = [1] * acc.ndim
        shp[ch_axis] = -1
        real = acc * (s_x * s_w).astype(np.float32).reshape(shp)
        std = _std(real) + 1e-12
        zp_out = 0 if relu else 128
        s_out = _round_sig(std / (50.0 if relu else 40.0))
        q = np.floor(real / np.float32(s_out) + np.float32(0.5)) + np.float32(zp_out)
        q = np.clip(q, float(zp_out if relu else 0), 255.0).astype(np.float32)
        return np.float32(s_out), zp_out, q

    def conv(self, x, s_x, zp_x, oc, ksize, stride=1, pad=0, groups=1, relu=True, residual=None):
        """x: expr NCHW uint8.  residual: (expr, scale, zp, shadow) joined with qnn.add + relu."""
        ic = x.shape[1]
        w, s_w, b = self.weights((oc, ic // groups, ksize, ksize), ksize * ksize * ic // groups)
        acc = _conv2d_f32(self.shadow - np.float32(zp_x), w.astype(np.float32), stride, pad, groups) \
            + b.astype(np.float32).reshape(1, -1, 1, 1)
        s_out, zp_out, q = self._calibrate(acc, s_x, s_w, relu and residual is None, 1)
        e = R.qnn_conv2d(x, R.const(w), zp_x, 0, np.float32(s_x), R.const(s_w), (ksize, ksize), oc,
                         strides=(stride, stride), padding=(pad, pad), groups=groups)
        e = R.bias_add(e, R.const(b))
        e = R.qnn_requantize(e, R.const((np.float32(s_x) * s_w).astype(np.float32)), 0, s_out, zp_out, axis=1,
                             out_dtype="int32")
        e = R.cast(R.clip(e, float(zp_out if (relu and residual is None) else 0), 255.0), "uint8")
        self.shadow = q
        if residual is not None:
            r_e, r_s, r_zp, r_shadow = residual
            real = (q - np.float32(zp_out)) * np.float32(s_out) + (r_shadow - np.float32(r_zp)) * np.float32(r_s)
            so = np.float32(_round_sig(_std(real) / 50.0))
            zo = 0
            e = R.qnn_add(e, r_e, s_out, zp_out, np.float32(r_s), r_zp, so, zo)
            e = R.clip(e, float(zo), 255.0)
            self.shadow = np.clip(np.floor(real / so + np.float32(0.5)) + zo, float(zo), 255.0).astype(np.float32)
            s_out, zp_out = so, zo
        self.n_layers += 1
        return e, s_out, zp_out

    def dense(self, x, s_x, zp_x, units, relu=True):
        k = x.shape[-1]
        w, s_w, b = self.weights((units, k), k)
        acc = (self.shadow - np.float32(zp_x)) @ w.astype(np.float32).T + b.astype(np.float32)
        s_out, zp_out, q = self._calibrate(acc, s_x, s_w, relu, acc.ndim - 1)
        e = R.qnn_dense(x, R.const(w), zp_x, 0, np.float32(s_x), R.const(s_w), units)
        e = R.bias_add(e, R.const(b), axis=1 if len(x.shape) == 2 else -1)
        e = R.qnn_requantize(e, R.const((np.float32(s_x) * s_w).astype(np.float32)), 0, s_out, zp_out, axis=-1,
                             out_dtype="int32")
        e = R.cast(R.clip(e, float(zp_out if relu else 0), 255.0), "uint8")
        self.shadow = q
        self.n_layers += 1
        return e, s_out, zp_out

    def max_pool(self, x, k, stride, pad):
        self.shadow = _pool2d_f32(self.shadow, k, stride, pad, "max")
        return R.max_pool2d(x, (k, k), (stride, stride), (pad, pad))

    def avg_pool_upcast(self, x, k=None, stride=None):
        """apply_with_upcast: cast int32 -> avg pool (truncating) -> cast uint8."""
        e = R.cast(x, "int32")
        if k is None:
            e = R.global_avg_pool2d(e)
            self.shadow = np.trunc(self.shadow.mean(axis=(2, 3), keepdims=True, dtype=np.float32))
        else:
            e = R.avg_pool2d(e, (k, k), (stride, stride))
            self.shadow = np.trunc(_pool2d_f32(self.shadow, k, stride, 0, "avg"))
        return R.cast(e, "uint8")


def _input(shape, seed=1):
    return np.random.default_rng(seed).integers(0, 256, size=shape).astype(np.uint8)


def synthetic_input(name, batch, seed=1):
    shape = {"mlp": (batch, 784), "lenet5": (batch, 1, 28, 28), "resnet18": (batch, 3, 224, 224),
             "mobilenetv2": (batch, 3, 224, 224), "bert_encoder": (batch * 128, 768), "bert_base": (batch * 128, 768)}[name]
    return _input(shape, seed)


IN_SCALE, IN_ZP = np.float32(1.0 / 64), 128


def mlp(batch=256, seed=0):
    """C1: int8 MLP 784-128-10 (qnn.dense + requantize)."""
    x = R.var("data", (batch, 784), "uint8")
    b = _Builder(seed, _input((4, 784), 7), IN_SCALE, IN_ZP)
    e, s, zp = b.dense(x, IN_SCALE, IN_ZP, 128, relu=True)
    e, s, zp = b.dense(e, s, zp, 10, relu=False)
    return R.Function([x], e)


def lenet5(batch=256, seed=0):
    """C2: LeNet-5: conv5x5x6 -> pool -> conv5x5x16 -> pool -> 120 -> 84 -> 10."""
    x = R.var("data", (batch, 1, 28, 28), "uint8")
    b = _Builder(seed, _input((4, 1, 28, 28), 7), IN_SCALE, IN_ZP)
    e, s, zp = b.conv(x, IN_SCALE, IN_ZP, 6, 5, pad=2)
    e = b.max_pool(e, 2, 2, 0)
    e, s, zp = b.conv(e, s, zp, 16, 5)
    e = b.max_pool(e, 2, 2, 0)
    e = R.batch_flatten(e)
    b.shadow = b.shadow.reshape(b.shadow.shape[0], -1)
    e, s, zp = b.dense(e, s, zp, 120)
    e, s, zp = b.dense(e, s, zp, 84)
    e, s, zp = b.dense(e, s, zp, 10, relu=False)
    return R.Function([x], e)


def resnet18(batch=64, seed=0, image=224, num_classes=1000):
    """C3: ResNet-18 (torchvision topology, quantized::conv2d_relu / add_relu chain)."""
    x = R.var("data", (batch, 3, image, image), "uint8")
    b = _Builder(seed, _input((1, 3, image, image), 7), IN_SCALE, IN_ZP)
    e, s, zp = b.conv(x, IN_SCALE, IN_ZP, 64, 7, stride=2, pad=3)
    e = b.max_pool(e, 3, 2, 1)
    for stage, (oc, stride) in enumerate([(64, 1), (128, 2), (256, 2), (512, 2)]):
        for blk in range(2):
            st = stride if blk == 0 else 1
            ident, ident_s, ident_zp, ident_shadow = e, s, zp, b.shadow
            y, ys, yzp = b.conv(e, s, zp, oc, 3, stride=st, pad=1)
            main_shadow = b.shadow
            if st != 1 or ident.shape[1] != oc:
                b.shadow = ident_shadow
                ident, ident_s, ident_zp = b.conv(ident, ident_s, ident_zp, oc, 1, stride=st, relu=False)
                ident_shadow = b.shadow
            b.shadow = main_shadow
            e, s, zp = b.conv(y, ys, yzp, oc, 3, pad=1, relu=True, residual=(ident, ident_s, ident_zp, ident_shadow))
    e = b.avg_pool_upcast(e)
    e = R.batch_flatten(e)
    b.shadow = b.shadow.reshape(b.shadow.shape[0], -1)
    e, s, zp = b.dense(e, s, zp, num_classes, relu=False)
    return R.Function([x], e)


def mobilenetv2(batch=128, seed=0, image=224, num_classes=1000):
    """C4: MobileNetV2 (depthwise 3x3 + pointwise 1x1 qnn.conv2d, ReLU6 folded into the clip)."""
    x = R.var("data", (batch, 3, image, image), "uint8")
    b = _Builder(seed, _input((1, 3, image, image), 7), IN_SCALE, IN_ZP)
    e, s, zp = b.conv(x, IN_SCALE, IN_ZP, 32, 3, stride=2, pad=1)
    cfg = [(1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1)]
    ic = 32
    for t, c, n, st0 in cfg:
        for i in range(n):
            st = st0 if i == 0 else 1
            ident, ident_s, ident_zp, ident_shadow = e, s, zp, b.shadow
            y, ys, yzp = e, s, zp
            if t != 1:
                y, ys, yzp = b.conv(y, ys, yzp, ic * t, 1)
            y, ys, yzp = b.conv(y, ys, yzp, ic * t, 3, stride=st, pad=1, groups=ic * t)
            if st == 1 and ic == c:
                # linear bottleneck joined with quantized::add (no relu): qnn.add -> clip(0, 255)
                y, ys, yzp = b.conv(y, ys, yzp, c, 1, relu=False)
                real = (b.shadow - np.float32(yzp)) * np.float32(ys) + (ident_shadow - np.float32(ident_zp)) * np.float32(ident_s)
                so, zo = np.float32(_round_sig(_std(real) / 40.0)), 128
                y = R.qnn_add(y, ident, ys, yzp, np.float32(ident_s), ident_zp, so, zo)
                b.shadow = np.clip(np.floor(real / so + np.float32(0.5)) + zo, 0.0, 255.0).astype(np.float32)
                ys, yzp = so, zo
            else:
                y, ys, yzp = b.conv(y, ys, yzp, c, 1, relu=False)
            e, s, zp, ic = y, ys, yzp, c
    e, s, zp = b.conv(e, s, zp, 1280, 1)
    e = b.avg_pool_upcast(e)
    e = R.batch_flatten(e)
    b.shadow = b.shadow.reshape(b.shadow.shape[0], -1)
    e, s, zp = b.dense(e, s, zp, num_classes, relu=False)
    return R.Function([x], e)


def bert_encoder(batch=8, seed=0, seq=128, hidden=768, heads=12, ffn=3072, layers=1):
    """C5: the integer chain of a BERT-base encoder layer (seq 128, hidden 768, 12 heads, FFN 3072): Q/K/V
    qnn.dense -> requantize to int8 (symmetric) -> head split (reshape / transpose) -> qnn.batch_matmul Q.K^T ->
    requantize -> uint8 "P" (zero point 0) -> qnn.batch_matmul P.V^T-form -> requantize -> head merge -> output
    qnn.dense -> qnn.add residual -> FFN qnn.dense x2 -> qnn.add residual.  Softmax and LayerNorm are float
    operators outside the integer path and are not modelled: P is the
    requantized, clipped score matrix, which keeps every integer operator at its BERT-base shape."""
    rng = np.random.default_rng(seed)
    d = hidden // heads
    x = R.var("data", (batch * seq, hidden), "uint8")
    shadow = _input((2 * seq, hidden), 7).astype(np.float32)  # calibration sample: 2 sequences
    cb = 2

    def calib(real, kind):
        std = _std(real) + 1e-12
        if kind == "i8":
            s_out = np.float32(_round_sig(std / 40.0))
            return s_out, 0, np.clip(np.floor(real / s_out + np.float32(0.5)), -128.0, 127.0).astype(np.float32)
        zp = 0 if kind == "relu" else 128
        s_out = np.float32(_round_sig(std / (50.0 if kind == "relu" else 40.0)))
        return s_out, zp, np.clip(np.floor(real / s_out + np.float32(0.5)) + zp, 0.0, 255.0).astype(np.float32)

    def dense(e, sh, s_x, zp_x, units, kind):
        k = e.shape[-1]
        w = rng.integers(-127, 128, size=(units, k)).astype(np.int8)
        s_w = (rng.uniform(0.5, 1.5, size=units) / (127.0 * np.sqrt(k))).astype(np.float32)
        b = rng.integers(-(1 << 12), 1 << 12, size=units).astype(np.int32)
        acc = (sh - np.float32(zp_x)) @ w.astype(np.float32).T + b.astype(np.float32)
        real = acc * (np.float32(s_x) * s_w).astype(np.float32)
        s_out, zp_out, q = calib(real, kind)
        y = R.qnn_dense(e, R.const(w), zp_x, 0, np.float32(s_x), R.const(s_w), units)
        y = R.bias_add(y, R.const(b), axis=1)
        y = R.qnn_requantize(y, R.const((np.float32(s_x) * s_w).astype(np.float32)), 0, s_out, zp_out, axis=-1,
                             out_dtype="int32")
        if kind == "i8":
            y = R.cast(R.clip(y, -128.0, 127.0), "int8")
        else:
            y = R.cast(R.clip(y, float(zp_out if kind == "relu" else 0), 255.0), "uint8")
        return y, s_out, zp_out, q

    def split_heads(e, sh, nb):  # [nb*seq, hidden] -> [nb*heads, seq, d]
        e = R.reshape(R.transpose(R.reshape(e, (nb, seq, heads, d)), (0, 2, 1, 3)), (nb * heads, seq, d))
        sh = sh.reshape(cb, seq, heads, d).transpose(0, 2, 1, 3).reshape(cb * heads, seq, d)
        return e, sh

    def bmm(a, a_sh, s_a, b_, b_sh, s_b, kind):
        real = (a_sh @ b_sh.transpose(0, 2, 1)) * np.float32(s_a) * np.float32(s_b)
        s_out, zp_out, q = calib(real, kind)
        y = R.qnn_batch_matmul(a, b_, 0, 0, np.float32(s_a), np.float32(s_b))
        y = R.qnn_requantize(y, np.float32(np.float32(s_a) * np.float32(s_b)), 0, s_out, zp_out, axis=-1, out_dtype="int32")
        if kind == "i8":
            y = R.cast(R.clip(y, -128.0, 127.0), "int8")
        else:
            y = R.cast(R.clip(y, 0.0, 255.0), "uint8")
        return y, s_out, zp_out, q

    def residual(a, a_sh, s_a, zp_a, b_, b_sh, s_b, zp_b):
        real = (a_sh - np.float32(zp_a)) * np.float32(s_a) + (b_sh - np.float32(zp_b)) * np.float32(s_b)
        so, zo = np.float32(_round_sig(_std(real) / 40.0)), 128
        y = R.clip(R.qnn_add(a, b_, np.float32(s_a), zp_a, np.float32(s_b), zp_b, so, zo), 0.0, 255.0)
        return y, so, zo, np.clip(np.floor(real / so + np.float32(0.5)) + zo, 0.0, 255.0).astype(np.float32)

    e, s, zp = x, IN_SCALE, IN_ZP
    for _ in range(layers):
        q, s_q, _, q_sh = dense(e, shadow, s, zp, hidden, "i8")
        k, s_k, _, k_sh = dense(e, shadow, s, zp, hidden, "i8")
        v, s_v, _, v_sh = dense(e, shadow, s, zp, hidden, "i8")
        q, q_sh = split_heads(q, q_sh, batch)
        k, k_sh = split_heads(k, k_sh, batch)
        v, v_sh = split_heads(v, v_sh, batch)
        p, s_p, _, p_sh = bmm(q, q_sh, s_q, k, k_sh, s_k, "relu")            # [B*h, S, S] uint8, zero point 0
        vt, vt_sh = R.transpose(v, (0, 2, 1)), v_sh.transpose(0, 2, 1)        # [B*h, d, S]
        c, s_c, _, c_sh = bmm(p, p_sh, s_p, vt, vt_sh, s_v, "i8")            # [B*h, S, d]
        c = R.reshape(R.transpose(R.reshape(c, (batch, heads, seq, d)), (0, 2, 1, 3)), (batch * seq, hidden))
        c_sh = c_sh.reshape(cb, heads, seq, d).transpose(0, 2, 1, 3).reshape(cb * seq, hidden)
        o, s_o, zp_o, o_sh = dense(c, c_sh, s_c, 0, hidden, "u8")
        e, s, zp, shadow = residual(o, o_sh, s_o, zp_o, e, shadow, s, zp)
        f, s_f, zp_f, f_sh = dense(e, shadow, s, zp, ffn, "relu")
        g, s_g, zp_g, g_sh = dense(f, f_sh, s_f, zp_f, hidden, "u8")
        e, s, zp, shadow = residual(g, g_sh, s_g, zp_g, e, shadow, s, zp)
        # LayerNorm stand-in between stacked layers: the quantized values are kept, the scale they are read with is reset
        # (without a normalisation the real-valued scales grow by orders of magnitude per layer and leave float32 after
        # four layers; the integer operators and their shapes are unaffected, a one-layer graph is unchanged)
        s = IN_SCALE
    return R.Function([x], e)


def bert_base(batch=8, seed=0):
    """C5 at full depth: all 12 encoder layers of BERT-base (bert_encoder with layers=12)."""
    return bert_encoder(batch=batch, seed=seed, layers=12)


MODELS = {"mlp": mlp, "lenet5": lenet5, "resnet18": resnet18, "mobilenetv2": mobilenetv2, "bert_encoder": bert_encoder,
          "bert_base": bert_base}
