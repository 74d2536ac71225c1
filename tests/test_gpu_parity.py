"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C ABI
(libtkb200.so), against the CPU oracle on the same inputs.  Bit-exact for the integer path
(np.testing.assert_equal); the oneDNN-style float epilogue mode is judged at +-1 like the
reference's own test (tests/python/contrib/test_tachikoma.py:1615-1616)."""
import numpy as np
import pytest

from golden_util import load_cases
from oracle import interp
from tachikoma_b200 import models, relay as R, runtime as RT, tachikoma as T

pytestmark = pytest.mark.gpu
CASES = load_cases()


def run_gpu(func, inputs, mode="int", conv_impl="auto", trace=False):
    part = T.partition_for_tachikoma(func, mode=mode)
    sym, js, names, consts = T.serialize(part)
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    mod.set_option("conv_impl", conv_impl)
    mod.init(consts)
    if trace:
        mod.trace_enable(2)
    ins = [np.asarray(np.asarray(inputs[p.name]).astype(p.dtype).reshape(p.shape), order="C") for p in part.params]
    out = np.empty(part.body.shape, dtype=part.body.dtype)
    mod.run(*ins, out)
    return out, mod, (js, consts, ins)


@pytest.mark.parametrize("impl", ["auto", "tc"])
@pytest.mark.parametrize("case", CASES, ids=[c["id"] for c in CASES])
def test_reference_goldens_on_gpu(case, impl):
    out, mod, _ = run_gpu(case["func"], case["inputs"], conv_impl=impl)
    np.testing.assert_equal(out, case["golden"])


# ---- tests/python/contrib/test_tachikoma.py:1422-1522 QNN conv profiles, replayed -------------------
def _conv_profile(shape=(1, 8, 5, 5), ker=(3, 3), pad=(1, 1), dil=(1, 1), oc=16, gr=1, nhwc=False, bias=True,
                  with_sum=False, asym=False):
    np.random.seed(0)
    n, ic, ih, iw = shape
    d_zp, rq_zp, rq_scl, o_zp = (3, 10, 0.1, 4) if asym else (0, 30, 0.2, 5)
    oh = (ih + 2 * pad[0] - (ker[0] - 1) * dil[0] - 1) + 1
    ow = (iw + 2 * pad[1] - (ker[1] - 1) * dil[1] - 1) + 1
    dl, kl = ("NHWC", "HWIO") if nhwc else ("NCHW", "OIHW")
    d_shape = (n, ih, iw, ic) if nhwc else shape
    w_shape = (ker[0], ker[1], ic // gr, oc) if nhwc else (oc, ic // gr, ker[0], ker[1])
    s_shape = (n, oh, ow, oc) if nhwc else (n, oc, oh, ow)
    b_shape = (oc,) if nhwc else (oc, 1, 1)
    data = R.var("in_0", d_shape, "uint8")
    args = {"in_0": np.random.uniform(0, 20, d_shape).astype("uint8")}
    wgh = R.const(np.random.uniform(-20, 20, w_shape).astype("int8"))
    op = R.qnn_conv2d(data, wgh, d_zp, 0, 0.2, 0.1, ker, oc, padding=pad, dilation=dil, groups=gr, data_layout=dl,
                      kernel_layout=kl)
    if bias:
        op = R.add(op, R.const(np.random.uniform(-50, 50, b_shape).astype("int32")))
    op = R.qnn_requantize(op, np.float32(0.2 * 0.1), 0, rq_scl, rq_zp, out_dtype="int32")
    op = R.cast(R.clip(op, 0.0, 255.0), "uint8")
    params = [data]
    if with_sum:
        s = R.var("in_1", s_shape, "uint8")
        args["in_1"] = np.random.uniform(0, 10, s_shape).astype("uint8")
        params.append(s)
        op = R.clip(R.qnn_add(op, s, rq_scl, rq_zp, 0.3, 15, 0.2, o_zp), 0.0, 255.0)
    return R.Function(params, op), args


CONV_PROFILES = {
    "Base": dict(), "NHWC": dict(nhwc=True), "Group": dict(pad=(0, 0), gr=2, asym=True),
    "DW": dict(shape=(1, 16, 5, 5), pad=(0, 0), gr=16, asym=True), "NoBias": dict(bias=False),
    "AsymmetricInput": dict(pad=(0, 0), asym=True), "AsymmetricInput_NHWC": dict(pad=(0, 0), asym=True, nhwc=True),
    "WithSum": dict(pad=(0, 0), asym=True, with_sum=True), "WithSum_NHWC": dict(pad=(0, 0), asym=True, with_sum=True, nhwc=True),
    "WithSum_NoBias": dict(pad=(0, 0), asym=True, with_sum=True, bias=False),
    "Dilated": dict(pad=(2, 2), dil=(2, 2)), "AsymmetricPadded": dict(asym=True),
}


@pytest.mark.parametrize("impl", ["simt", "tc"])
@pytest.mark.parametrize("name", list(CONV_PROFILES))
def test_qnn_conv2d_profiles(name, impl):
    func, args = _conv_profile(**CONV_PROFILES[name])
    ref, _ = interp.run_expr(func, args)
    out, mod, _ = run_gpu(func, args, mode="int", conv_impl=impl)
    assert "tachikoma.qnn.conv2d" in mod.plan_summary()
    np.testing.assert_equal(out, ref)
    if name != "AsymmetricPadded":  # "Input ZP is not compatible with padding" in the float legalization
        outf, _, _ = run_gpu(func, args, mode="dnnl_float", conv_impl=impl)
        # atol=1: int values should match with +-1 quantum value tolerance (test_tachikoma.py:1615)
        assert np.abs(outf.astype(np.int32) - ref.astype(np.int32)).max() <= 1


@pytest.mark.parametrize("impl", ["simt", "tc"])
@pytest.mark.parametrize("profile", ["Base", "AsymmetricInput", "AsymmetricInput_Sum"])
def test_qnn_dense_profiles(profile, impl):
    # tests/python/contrib/test_tachikoma.py:1690-1762: N=2, IC=10, OC=16 (IC=16 added for the TMA path)
    for ic in (10, 16):
        np.random.seed(0)
        asym = profile != "Base"
        d_zp, rq_zp, rq_scl, o_zp = (3, 10, 0.1, 4) if asym else (0, 30, 0.2, 5)
        data = R.var("in_0", (2, ic), "uint8")
        args = {"in_0": np.random.uniform(0, 20, (2, ic)).astype("uint8")}
        op = R.qnn_dense(data, R.const(np.random.uniform(-20, 20, (16, ic)).astype("int8")), d_zp, 0, 0.2, 0.1, 16)
        op = R.add(op, R.const(np.random.uniform(-50, 50, (16,)).astype("int32")))
        op = R.qnn_requantize(op, np.float32(0.02), 0, rq_scl, rq_zp, out_dtype="int32")
        op = R.cast(R.clip(op, 0.0, 255.0), "uint8")
        params = [data]
        if profile.endswith("Sum"):
            s = R.var("in_1", (2, 16), "uint8")
            args["in_1"] = np.random.uniform(0, 10, (2, 16)).astype("uint8")
            params.append(s)
            op = R.clip(R.qnn_add(op, s, rq_scl, rq_zp, 0.3, 15, 0.2, o_zp), 0.0, 255.0)
        func = R.Function(params, op)
        ref, _ = interp.run_expr(func, args)
        out, _, _ = run_gpu(func, args, conv_impl=impl)
        np.testing.assert_equal(out, ref)
        outf, _, _ = run_gpu(func, args, mode="dnnl_float", conv_impl=impl)
        assert np.abs(outf.astype(np.int32) - ref.astype(np.int32)).max() <= 1


GEMM_SHAPES = [(1, 8, 16), (128, 64, 128), (130, 72, 144), (257, 10, 784), (300, 300, 1024), (64, 1000, 512),
               (1024, 128, 2304), (77, 200, 48)]


@pytest.mark.parametrize("mnk", GEMM_SHAPES, ids=[f"{m}x{n}x{k}" for m, n, k in GEMM_SHAPES])
@pytest.mark.parametrize("xdt", ["uint8", "int8"])
def test_tcgen05_gemm_vs_oracle(mnk, xdt):
    m, n, k = mnk
    rng = np.random.default_rng(m * 7 + n * 3 + k)
    lo, hi = (0, 256) if xdt == "uint8" else (-128, 128)
    x = rng.integers(lo, hi, (m, k)).astype(xdt)
    w = rng.integers(-128, 128, (n, k)).astype("int8")
    zp = 7 if xdt == "uint8" else -3
    f = R.Function([R.var("x", (m, k), xdt)], R.qnn_dense(R.var("x", (m, k), xdt), R.const(w), zp, 0, 1.0, 1.0, n))
    xv = R.var("x", (m, k), xdt)
    f = R.Function([xv], R.qnn_dense(xv, R.const(w), zp, 0, 1.0, 1.0, n))
    ref, _ = interp.run_expr(f, {"x": x})
    out, mod, _ = run_gpu(f, {"x": x}, conv_impl="tc")
    assert "tcgen05" in mod.plan_summary()
    np.testing.assert_equal(out, ref)


def test_tcgen05_batch_matmul_vs_oracle():
    rng = np.random.default_rng(5)
    for (b, m, n, k, yb) in [(3, 128, 64, 64, 3), (4, 50, 130, 32, 1), (12, 128, 128, 64, 12)]:
        x = rng.integers(-128, 128, (b, m, k)).astype("int8")
        y = rng.integers(-128, 128, (yb, n, k)).astype("int8")
        xv, yv = R.var("x", x.shape, "int8"), R.var("y", y.shape, "int8")
        f = R.Function([xv, yv], R.qnn_batch_matmul(xv, yv, 0, 0, 1.0, 1.0))
        ref, _ = interp.run_expr(f, {"x": x, "y": y})
        for impl in ("tc", "simt"):
            out, _, _ = run_gpu(f, {"x": x, "y": y}, conv_impl=impl)
            np.testing.assert_equal(out, ref)


# ---- implicit-GEMM convolution (TMA im2col mode) and the folded epilogue ----------------------------
def _chain_conv(n, ic, h, w, oc, k, stride, pad, dil, zp_a, per_channel, relu, with_sum, seed, nhwc=False,
                out_scale=None, raw=False, rounding="None", xdt="uint8", sum_scales=(0.3, 15, 0.2, 4), zp_relu=0,
                out_dt="uint8", clip=None, groups=1):
    rng = np.random.default_rng(seed)
    dl, kl = ("NHWC", "HWIO") if nhwc else ("NCHW", "OIHW")
    d_shape = (n, h, w, ic) if nhwc else (n, ic, h, w)
    w_shape = (k, k, ic // groups, oc) if nhwc else (oc, ic // groups, k, k)
    if groups > 1 and nhwc:
        kl = "HWOI" if groups == ic else kl
        w_shape = (k, k, oc, 1) if groups == ic else w_shape
    lo, hi = (0, 256) if xdt == "uint8" else (-128, 128)
    x = rng.integers(lo, hi, d_shape).astype(xdt)
    wt = rng.integers(-127, 128, w_shape).astype("int8")
    s_x = np.float32(0.02)
    s_w = (rng.uniform(0.5, 1.5, oc) / (127.0 * np.sqrt(k * k * ic))).astype("float32") if per_channel \
        else np.float32(1.0 / (127.0 * np.sqrt(k * k * ic)))
    data = R.var("data", d_shape, xdt)
    e = R.qnn_conv2d(data, R.const(wt), zp_a, 0, s_x, R.const(s_w) if per_channel else s_w, (k, k), oc,
                     strides=(stride, stride), padding=(pad, pad), dilation=(dil, dil), data_layout=dl, kernel_layout=kl,
                     groups=groups)
    args = {"data": x}
    params = [data]
    if not raw:
        b = rng.integers(-(1 << 12), 1 << 12, oc).astype("int32")
        e = R.bias_add(e, R.const(b), axis=3 if nhwc else 1)
        zp_out = zp_relu if relu else (128 if out_dt == "uint8" else 0)
        so = np.float32(out_scale if out_scale is not None else 0.02 * 1.2)
        e = R.qnn_requantize(e, R.const((s_x * s_w).astype("float32")) if per_channel else np.float32(s_x * s_w), 0, so,
                             zp_out, axis=3 if nhwc else 1, out_dtype="int32", rounding=rounding)
        dlo, dhi = (0.0, 255.0) if out_dt == "uint8" else (-128.0, 127.0)
        c_lo, c_hi = clip if clip is not None else (float(zp_out) if relu else dlo, dhi)
        e = R.cast(R.clip(e, c_lo, c_hi), out_dt)
        if with_sum:
            sv = R.var("res", e.shape, "uint8")
            args["res"] = rng.integers(0, 256, e.shape).astype("uint8")
            params.append(sv)
            rs, rz, os_, oz = sum_scales
            e = R.clip(R.qnn_add(e, sv, so, zp_out, rs, rz, os_, oz), 0.0, 255.0)
    return R.Function(params, e), args


IMPLICIT_CASES = {
    # name: (n, ic, h, w, oc, k, stride, pad, dil, zp_a, per_channel, relu, with_sum)
    "c64_3x3_p1": (2, 64, 12, 12, 64, 3, 1, 1, 1, 0, True, True, False),
    "c64_3x3_p1_zp": (2, 64, 12, 12, 64, 3, 1, 1, 1, 128, True, True, False),
    "c64_3x3_p1_zp_sum": (2, 64, 9, 11, 64, 3, 1, 1, 1, 7, True, False, True),
    "c128_3x3_s2": (3, 128, 14, 14, 128, 3, 2, 1, 1, 5, True, True, False),
    "c64_1x1_s2": (2, 64, 16, 16, 128, 1, 2, 0, 1, 9, True, False, False),
    "c32_3x3_p1_pt": (1, 32, 7, 7, 48, 3, 1, 1, 1, 3, False, True, True),
    "c96_3x3_dil2": (1, 96, 13, 13, 40, 3, 1, 2, 2, 11, True, True, False),
    "c256_3x3_n512": (1, 256, 7, 7, 512, 3, 1, 1, 1, 0, True, True, True),
    "c32_7x7_s2_p3": (2, 32, 30, 30, 64, 7, 2, 3, 1, 128, True, True, False),
    "c64_5x5_p2_big": (4, 64, 31, 29, 72, 5, 1, 2, 1, 1, True, True, False),
    "c512_1x1": (2, 512, 7, 7, 1000, 1, 1, 0, 1, 4, True, False, False),
    "c64_3x3_p0_s8": (2, 64, 12, 12, 64, 3, 1, 0, 1, 128, True, True, False),
    "c24_1x1_rowpad": (2, 24, 9, 9, 40, 1, 1, 0, 1, 6, True, True, False),   # K = 24 padded to 32 by pad_rows8_kernel
    "c40_1x1_rowpad_sum": (1, 40, 7, 5, 24, 1, 1, 0, 1, 200, False, False, True),
}


@pytest.mark.parametrize("nhwc", [False, True], ids=["nchw", "nhwc"])
@pytest.mark.parametrize("name", list(IMPLICIT_CASES))
def test_implicit_conv_vs_oracle(name, nhwc):
    cfg = IMPLICIT_CASES[name]
    func, args = _chain_conv(*cfg, seed=len(name) * 13 + cfg[1], nhwc=nhwc, xdt="int8" if name.endswith("s8") else "uint8")
    ref, _ = interp.run_expr(func, args)
    out, mod, _ = run_gpu(func, args, conv_impl="tc")
    plan = mod.plan_summary()
    assert "tcgen05" in plan and "epi=fast" in plan, plan
    if cfg[5] > 1 or cfg[6] > 1:
        assert "implicit-GEMM" in plan or "halo-reuse" in plan, plan
    np.testing.assert_equal(out, ref)
    # the TMA-im2col route (halo kernel off) and the explicit-im2col route agree bit for bit
    out2, mod2, _ = run_gpu(func, args, conv_impl="tc_im2col")
    assert "halo-reuse" not in mod2.plan_summary()
    np.testing.assert_equal(out2, ref)
    out3, _, _ = run_gpu(func, args, conv_impl="tc_explicit")
    np.testing.assert_equal(out3, ref)


HALO_CASES = {
    # stride-1 convolutions that take the halo-reuse kernel (patch loads, taps as shifted descriptors, resident weights)
    "h_c64_56": (2, 64, 56, 56, 64, 3, 1, 1, 1, 3, True, True, False),            # ResNet-18 layer1 geometry
    "h_c128_28_sum": (3, 128, 28, 28, 128, 3, 1, 1, 1, 128, True, False, True),  # layer2 geometry + residual
    "h_c64_ragged_n48": (2, 64, 13, 20, 48, 3, 1, 1, 1, 5, True, True, True),    # last tile of an image is partial
    "h_c64_5x5": (1, 64, 17, 17, 32, 5, 1, 2, 1, 9, False, True, False),
    "h_c192_3chunks": (2, 192, 14, 14, 64, 3, 1, 1, 1, 77, True, True, False),   # 3 channel chunks of 64
    "h_c256_2chunks": (1, 256, 14, 14, 64, 3, 1, 1, 1, 200, True, False, True),  # 2 channel chunks of 128
    "h_c64_p0_s8": (5, 64, 30, 30, 96, 3, 1, 0, 1, 0, True, True, False),        # no padding, 150 tiles > 148 SMs
}


@pytest.mark.parametrize("nhwc", [False, True], ids=["nchw", "nhwc"])
@pytest.mark.parametrize("name", list(HALO_CASES))
def test_halo_conv_vs_oracle(name, nhwc):
    cfg = HALO_CASES[name]
    func, args = _chain_conv(*cfg, seed=len(name) * 7 + cfg[2], nhwc=nhwc, xdt="int8" if name.endswith("s8") else "uint8")
    ref, _ = interp.run_expr(func, args)
    out, mod, _ = run_gpu(func, args, conv_impl="tc")
    assert "halo-reuse" in mod.plan_summary(), mod.plan_summary()
    np.testing.assert_equal(out, ref)


CLAMP_CASES = {
    # (oc, relu, zp_relu, out_dt, clip): which epilogue clamp the tile takes
    "u8_full_range_sat": (64, True, 0, "uint8", None),
    "u8_relu_zp10_explicit": (64, True, 10, "uint8", None),
    "u8_clip_20_200_explicit": (64, False, 0, "uint8", (20.0, 200.0)),
    "s8_full_range_sat": (64, False, 0, "int8", None),
    "s8_clip_explicit": (48, False, 0, "int8", (-100.0, 90.0)),
    "u8_n128_full_range": (128, True, 0, "uint8", None),
    "s8_n128_full_range": (128, False, 0, "int8", None),
}


@pytest.mark.parametrize("stride", [1, 2], ids=["halo", "im2col"])
@pytest.mark.parametrize("name", list(CLAMP_CASES))
def test_fast_epilogue_clamp_variants(name, stride):
    """Saturating-pack epilogue (clip == dtype range) and explicit-clamp epilogue, uint8 and int8 outputs, on the
    halo kernel (stride 1) and the implicit-GEMM kernel (stride 2); accumulators span far beyond the clamp."""
    oc, relu, zp_relu, out_dt, clip = CLAMP_CASES[name]
    func, args = _chain_conv(2, 64, 14, 14, oc, 3, stride, 1, 1, 9, True, relu, False, seed=len(name), zp_relu=zp_relu,
                             out_dt=out_dt, clip=clip, out_scale=0.004)
    ref, _ = interp.run_expr(func, args)
    assert ref.min() == (clip[0] if clip else (zp_relu if relu else (0 if out_dt == "uint8" else -128)))
    out, mod, _ = run_gpu(func, args, conv_impl="tc")
    assert "epi=fast" in mod.plan_summary(), mod.plan_summary()
    np.testing.assert_equal(out, ref)


DEPTHWISE_CASES = {
    # name: (n, c, h, w, k, stride, pad, zp_a, per_channel, relu, xdt, raw)
    "dw_c32_3x3_s1": (2, 32, 14, 14, 3, 1, 1, 0, True, True, "uint8", False),
    "dw_c96_3x3_s2_zp": (2, 96, 15, 13, 3, 2, 1, 128, True, True, "uint8", False),
    "dw_c16_5x5_p2": (1, 16, 9, 9, 5, 1, 2, 7, True, False, "uint8", False),
    "dw_c48_3x3_s8": (3, 48, 10, 10, 3, 1, 1, 0, False, True, "int8", False),
    "dw_c64_3x3_raw": (2, 64, 8, 8, 3, 1, 1, 77, True, True, "uint8", True),
    "dw_c144_3x3_p0": (1, 144, 12, 12, 3, 2, 0, 5, True, True, "uint8", False),
}


@pytest.mark.parametrize("name", list(DEPTHWISE_CASES))
def test_depthwise_conv_vs_oracle(name):
    """Depthwise qnn.conv2d (groups == channels) through the dedicated channels-last kernel, the generic CUDA-core
    kernel ("simt") and the oracle: identical bits."""
    n, c, h, w, k, stride, pad, zp_a, per_channel, relu, xdt, raw = DEPTHWISE_CASES[name]
    func, args = _chain_conv(n, c, h, w, c, k, stride, pad, 1, zp_a, per_channel, relu, False, seed=len(name) + c, xdt=xdt,
                             raw=raw, groups=c, out_scale=0.003)
    ref, _ = interp.run_expr(func, args)
    out, mod, _ = run_gpu(func, args)
    assert "depthwise conv" in mod.plan_summary(), mod.plan_summary()
    np.testing.assert_equal(out, ref)
    out2, mod2, _ = run_gpu(func, args, conv_impl="simt")
    assert "depthwise conv" not in mod2.plan_summary()
    np.testing.assert_equal(out2, ref)


QADD_CASES = [
    # (dtype, lhs_scale, lhs_zp, rhs_scale, rhs_zp, out_scale, out_zp)
    ("uint8", 0.02, 128, 0.03, 127, 0.05, 128), ("uint8", 0.5, 3, 0.5, 3, 0.5, 3), ("uint8", 0.25, 0, 0.5, 10, 0.125, 20),
    ("int8", 0.02, -5, 0.013, 7, 0.031, -3), ("int8", 1.0, 0, 0.5, 0, 0.25, 0), ("uint8", 0.001, 255, 7.5, 0, 0.9, 1),
]


@pytest.mark.parametrize("cfg", QADD_CASES, ids=[str(i) for i in range(len(QADD_CASES))])
def test_qnn_add_standalone_vs_oracle(cfg):
    """Un-fused qnn.add on 8-bit tensors (table-lookup kernel for >= 4096 elements, generic kernel below): every byte
    pair of the full 256 x 256 input space appears."""
    dt, ls, lz, rs, rz, os_, oz = cfg
    lo = 0 if dt == "uint8" else -128
    a = (np.arange(65536) % 256 + lo).astype(dt).reshape(4, 64, 256)
    b = (np.arange(65536) // 256 + lo).astype(dt).reshape(4, 64, 256)
    va, vb = R.var("a", a.shape, dt), R.var("b", b.shape, dt)
    func = R.Function([va, vb], R.qnn_add(va, vb, ls, lz, rs, rz, os_, oz))
    ref, _ = interp.run_expr(func, {"a": a, "b": b})
    out, mod, _ = run_gpu(func, {"a": a, "b": b})
    np.testing.assert_equal(out, ref)
    wa, wb = R.var("a", (3, 5), dt), R.var("b", (3, 5), dt)
    small = R.Function([wa, wb], R.qnn_add(wa, wb, ls, lz, rs, rz, os_, oz))
    sa, sb = a.reshape(-1)[1000:1015].reshape(3, 5), b.reshape(-1)[5000:5015].reshape(3, 5)
    ref2, _ = interp.run_expr(small, {"a": sa, "b": sb})
    out2, _, _ = run_gpu(small, {"a": sa, "b": sb})
    np.testing.assert_equal(out2, ref2)


def test_implicit_conv_raw_int32_and_generic_epilogue():
    # plain qnn.conv2d (int32 accumulators out) through the implicit path with border correction
    func, args = _chain_conv(2, 64, 10, 10, 64, 3, 1, 1, 1, 77, True, True, False, seed=3, raw=True)
    ref, _ = interp.run_expr(func, args)
    out, mod, _ = run_gpu(func, args, conv_impl="tc")
    assert "implicit-GEMM" in mod.plan_summary() and "epi=raw" in mod.plan_summary()
    np.testing.assert_equal(out, ref)
    # TONEAREST rounding and multipliers >= 0.5 are outside the folded epilogue: generic path, same bits
    for kw in (dict(rounding="TONEAREST"), dict(out_scale=0.02 / (127.0 * 24) * 1.1)):
        func, args = _chain_conv(2, 64, 10, 10, 64, 3, 1, 1, 1, 77, True, True, True, seed=4, **kw)
        ref, _ = interp.run_expr(func, args)
        out, mod, _ = run_gpu(func, args, conv_impl="tc")
        assert "epi=int" in mod.plan_summary(), mod.plan_summary()
        np.testing.assert_equal(out, ref)


@pytest.mark.parametrize("sum_scales", [(0.3, 15, 0.2, 4), (0.024, 0, 0.024, 0), (0.024, 3, 0.024, 9), (0.012, 0, 0.024, 0),
                                        (0.05, 200, 0.0125, 30), (0.024, 0, 0.04, 0)])
def test_fast_epilogue_residual_variants(sum_scales):
    """qnn.add sides: generic multiplier, identity, equal scales with different zero points, power-of-two
    ratios (the intrin_rule.cc:221-239 shortcut), ratios > 1 (left shift)."""
    func, args = _chain_conv(2, 64, 8, 8, 64, 3, 1, 1, 1, 0, True, True, True, seed=11, sum_scales=sum_scales)
    ref, _ = interp.run_expr(func, args)
    out, mod, _ = run_gpu(func, args, conv_impl="tc")
    assert "epi=fast" in mod.plan_summary(), mod.plan_summary()
    np.testing.assert_equal(out, ref)
    out2, _, _ = run_gpu(func, args, conv_impl="simt")
    np.testing.assert_equal(out2, ref)


def test_fast_epilogue_saturating_accumulators():
    """All-255 activations x all-(+/-127) weights: the largest accumulators the shapes allow."""
    for sign in (1, -1):
        n, ic, h, w, oc, k = 1, 512, 6, 6, 64, 3
        x = np.full((n, ic, h, w), 255, "uint8")
        wt = np.full((oc, ic, k, k), 127 * sign, "int8")
        data = R.var("data", x.shape, "uint8")
        e = R.qnn_conv2d(data, R.const(wt), 0, 0, 0.02, 0.001, (k, k), oc, padding=(1, 1))
        e = R.bias_add(e, R.const(np.full(oc, sign * (1 << 20), "int32")))
        e = R.qnn_requantize(e, np.float32(2e-5), 0, np.float32(2.0), 128, axis=1, out_dtype="int32")
        e = R.cast(R.clip(e, 0.0, 255.0), "uint8")
        func = R.Function([data], e)
        ref, _ = interp.run_expr(func, {"data": x})
        out, mod, _ = run_gpu(func, {"data": x}, conv_impl="tc")
        np.testing.assert_equal(out, ref)


MODEL_CASES = [("mlp", dict(batch=64)), ("lenet5", dict(batch=8)), ("resnet18", dict(batch=2, image=64, num_classes=100)),
               ("mobilenetv2", dict(batch=2, image=64, num_classes=100)),
               ("bert_encoder", dict(batch=2, seq=32, hidden=128, heads=4, ffn=256))]


@pytest.mark.parametrize("impl", ["simt", "tc"])
@pytest.mark.parametrize("name,kw", MODEL_CASES, ids=[c[0] for c in MODEL_CASES])
def test_models_bit_exact_with_trace(name, kw, impl):
    func = models.MODELS[name](**kw)
    x = models.synthetic_input(name, kw["batch"])
    if "image" in kw:
        x = np.ascontiguousarray(x[:, :, :kw["image"], :kw["image"]])
    if name == "bert_encoder":
        x = np.random.default_rng(1).integers(0, 256, (kw["batch"] * kw["seq"], kw["hidden"])).astype("uint8")
    ref, _ = interp.run_expr(func, {"data": x})  # un-partitioned, op by op
    out, mod, (js, consts, ins) = run_gpu(func, {"data": x}, conv_impl=impl, trace=True)
    np.testing.assert_equal(out, ref)
    # serialized trace bytes: GPU-produced blob == oracle-produced blob
    _, recs = interp.run_json(js, consts, ins, want_trace=True)
    want = interp.save_ndarray_list(recs)
    got = mod.trace_blob()
    assert len(got) == len(want)
    assert got == want
    assert mod.launch_count() > 0


@pytest.mark.parametrize("impl", ["tc_chain_all", "tc_chain_siblings", "tc_chain_split"])
@pytest.mark.parametrize("name,kw", [MODEL_CASES[2], MODEL_CASES[4], ("resnet18", dict(batch=4, image=224))], ids=["resnet18-64px", "bert", "resnet18-224px"])
def test_chained_launches_bit_exact_with_trace(name, kw, impl):
    """Chained persistent tcgen05 launches (csrc/kernels_tc_chain.cu): layers merged into one kernel with per-image
    completion counters, dynamic work claiming, TMA-store epilogue.  An opt-in experiment (slower than the single-layer
    kernels on B200), kept bit-exact: output and serialized trace equal the oracle's."""
    func = models.MODELS[name](**kw)
    x = models.synthetic_input(name, kw["batch"])
    if "image" in kw:
        x = np.ascontiguousarray(x[:, :, :kw["image"], :kw["image"]])
    if name == "bert_encoder":
        x = np.random.default_rng(1).integers(0, 256, (kw["batch"] * kw["seq"], kw["hidden"])).astype("uint8")
    out, mod, (js, consts, ins) = run_gpu(func, {"data": x}, conv_impl=impl, trace=True)
    if name == "resnet18":
        assert "chained tcgen05 launch" in mod.plan_summary() or "chain 0" in mod.plan_summary()
    outs, recs = interp.run_json(js, consts, ins, want_trace=True)
    np.testing.assert_equal(out, outs[0])
    assert mod.trace_blob() == interp.save_ndarray_list(recs)
    for _ in range(3):  # counters are re-zeroed every run
        out2 = np.zeros_like(out)
        mod.run(*ins, out2)
        np.testing.assert_equal(out2, outs[0])


def test_gpu_trace_passes_record_by_record_replay():
    """the GPU-produced trace is a valid witness -- every record re-computed on the oracle
    from the recorded inputs alone, chain and constants checked (oracle/replay.py)."""
    from oracle import replay
    func = models.resnet18(batch=2, image=64, num_classes=100)
    x = np.ascontiguousarray(models.synthetic_input("resnet18", 2)[:, :, :64, :64])
    _, mod, (js, consts, _ins) = run_gpu(func, {"data": x}, conv_impl="tc", trace=True)
    assert replay.verify_trace(js, mod.trace_blob(), consts) == []


def test_resnet18_full_size_properties():
    """Full BASELINE size (224x224, batch 8 here to bound oracle time): GPU tc == GPU simt bit for bit, and the
    first image equals the oracle."""
    func = models.resnet18(batch=8)
    x = models.synthetic_input("resnet18", 8)
    a, _, _ = run_gpu(func, {"data": x}, conv_impl="tc")
    b, _, _ = run_gpu(func, {"data": x}, conv_impl="simt")
    np.testing.assert_equal(a, b)
    f1 = models.resnet18(batch=1)
    ref, _ = interp.run_expr(f1, {"data": x[:1]})
    np.testing.assert_equal(a[:1], ref)


def test_bert_base_layer_full_size():
    """C5 at BASELINE size (seq 128, hidden 768, 12 heads, FFN 3072; 2 sequences to bound oracle time): dense and
    batch_matmul run on tcgen05 and the output and the serialized trace equal the oracle's byte for byte."""
    func = models.bert_encoder(batch=2)
    x = models.synthetic_input("bert_encoder", 2)
    ref, _ = interp.run_expr(func, {"data": x})
    out, mod, (js, consts, ins) = run_gpu(func, {"data": x}, conv_impl="tc", trace=True)
    plan = mod.plan_summary()
    assert plan.count("tcgen05") >= 8, plan  # 6 dense + 2 batch_matmul
    np.testing.assert_equal(out, ref)
    _, recs = interp.run_json(js, consts, ins, want_trace=True)
    assert mod.trace_blob() == interp.save_ndarray_list(recs)
    out2, _, _ = run_gpu(func, {"data": x}, conv_impl="simt")
    np.testing.assert_equal(out2, ref)


def test_bert_base_all_twelve_layers():
    """The whole BERT-base encoder stack (12 layers at seq 128, hidden 768, 12 heads, FFN 3072; two sequences): 72 dense and
    24 batch_matmul launches on tcgen05, output and trace byte for byte the oracle's (bench.py times ONE layer and says
    so in its workload string)."""
    func = models.bert_encoder(batch=2, layers=12)
    x = models.synthetic_input("bert_encoder", 2)
    out, mod, (js, consts, ins) = run_gpu(func, {"data": x}, conv_impl="tc", trace=True)
    assert mod.plan_summary().count("tcgen05") >= 96
    want, recs = interp.run_json(js, consts, ins, want_trace=True)
    np.testing.assert_equal(out, want[0])
    assert mod.trace_blob() == interp.save_ndarray_list(recs)


POOL_CASES = [
    # (n, h, w, c, k, stride, pad, dtype)
    (2, 16, 16, 64, 3, 2, 1, "uint8"), (3, 13, 9, 32, 3, 2, 1, "uint8"), (1, 7, 7, 16, 3, 2, 1, "int8"),
    (2, 15, 14, 48, 3, 2, 0, "int8"), (1, 2, 2, 16, 3, 2, 1, "uint8"), (2, 12, 12, 32, 3, 2, 2, "uint8"),
    (2, 12, 10, 32, 2, 2, 0, "uint8"), (1, 9, 9, 16, 3, 1, 1, "int8"),
]


@pytest.mark.parametrize("cfg", POOL_CASES, ids=["x".join(str(v) for v in c) for c in POOL_CASES])
def test_max_pool_nhwc_8bit_vs_oracle(cfg):
    n, h, w, c, k, st, pad, dt = cfg
    rng = np.random.default_rng(h * 31 + w)
    lo, hi = (0, 256) if dt == "uint8" else (-128, 128)
    x = rng.integers(lo, hi, (n, h, w, c)).astype(dt)
    x[0, 0, 0, :] = lo  # minimum of the dtype in a corner window: padding must never win
    data = R.var("data", x.shape, dt)
    func = R.Function([data], R.max_pool2d(data, (k, k), (st, st), (pad, pad), layout="NHWC"))
    ref, _ = interp.run_expr(func, {"data": x})
    out, _, _ = run_gpu(func, {"data": x})
    np.testing.assert_equal(out, ref)


AVG_POOL_CASES = [
    # (n, c, h, w, k, stride, pad, count_include_pad, dtype, layout)
    (2, 16, 12, 12, 2, 2, 0, False, "int32", "NCHW"), (1, 8, 9, 7, 3, 2, 1, False, "int32", "NCHW"),
    (1, 8, 9, 7, 3, 2, 1, True, "int32", "NCHW"), (2, 32, 8, 8, 3, 1, 1, False, "int32", "NHWC"),
    (1, 4, 6, 6, 2, 2, 0, False, "uint8", "NCHW"), (1, 4, 6, 6, 3, 3, 0, False, "int8", "NHWC"),
    (1, 4, 5, 5, 5, 1, 0, False, "int16", "NCHW"),
]


@pytest.mark.parametrize("cfg", AVG_POOL_CASES, ids=["x".join(str(v) for v in c) for c in AVG_POOL_CASES])
def test_windowed_avg_pool_vs_oracle(cfg):
    """nn.avg_pool2d (pool2d_kernel): truncating division, padding excluded / included, and the window sum wrapping in
    the tensor's own dtype for 8 / 16-bit tensors (include/tvm/topi/nn/pooling.h:599)."""
    n, c, h, w, k, st, pad, cip, dt, lay = cfg
    rng = np.random.default_rng(h * 17 + w + k)
    lo, hi = {"int32": (-100000, 100000), "uint8": (150, 256), "int8": (-128, -60), "int16": (20000, 32767)}[dt]
    shape = (n, c, h, w) if lay == "NCHW" else (n, h, w, c)
    x = rng.integers(lo, hi, shape).astype(dt)
    data = R.var("data", shape, dt)
    func = R.Function([data], R.avg_pool2d(data, (k, k), (st, st), (pad, pad), layout=lay, count_include_pad=cip))
    ref, _ = interp.run_expr(func, {"data": x})
    out, mod, _ = run_gpu(func, {"data": x})
    assert "pooling kernel" in mod.plan_summary()
    np.testing.assert_equal(out, ref)


def test_global_avg_pool_narrow_dtype_and_rejection():
    """Adaptive form: the divisor is multiplied up in the tensor dtype (pooling.h:384-389).  uint8 20x20: divisor 400
    wraps to 144; 16x16 wraps to zero, which is a division by zero in the reference: rejected at init."""
    x = np.random.default_rng(3).integers(0, 256, (2, 8, 20, 20)).astype("uint8")
    data = R.var("data", x.shape, "uint8")
    func = R.Function([data], R.global_avg_pool2d(data))
    ref, _ = interp.run_expr(func, {"data": x})
    out, _, _ = run_gpu(func, {"data": x})
    np.testing.assert_equal(out, ref)
    d2 = R.var("data", (1, 8, 16, 16), "uint8")
    with pytest.raises(RT.TachikomaError, match="Unsupported op: nn.global_avg_pool2d"):
        run_gpu(R.Function([d2], R.global_avg_pool2d(d2)), {"data": np.zeros((1, 8, 16, 16), "uint8")})


PAD_CASES = [((2, 3, 5, 7), [[0, 0], [0, 0], [1, 2], [3, 0]], 7, "uint8"), ((4, 6), [[1, 1], [0, 3]], -5, "int32"),
             ((2, 5, 4), [[0, 1], [2, 0], [0, 0]], 0, "int8"), ((1, 8, 6, 6), [[0, 0], [0, 0], [3, 3], [3, 3]], 128, "uint8")]


@pytest.mark.parametrize("cfg", PAD_CASES, ids=[str(c[0]) for c in PAD_CASES])
def test_nn_pad_vs_oracle(cfg):
    """nn.pad with the input zero point as pad value: the explicit padding the PyTorch frontend emits in front of
    qnn.conv2d (python/tvm/relay/frontend/qnn_torch.py:736-744)."""
    shape, pw, value, dt = cfg
    rng = np.random.default_rng(len(shape))
    x = rng.integers(0, 100, shape).astype(dt)
    data = R.var("data", shape, dt)
    func = R.Function([data], R.pad(data, pw, float(value)))
    ref, _ = interp.run_expr(func, {"data": x})
    out, mod, _ = run_gpu(func, {"data": x})
    assert "pad kernel" in mod.plan_summary()
    np.testing.assert_equal(out, ref)
    assert out.shape == tuple(s + a + b for s, (a, b) in zip(shape, pw))


def _module(func, impl="tc"):
    part = T.partition_for_tachikoma(func)
    sym, js, names, consts = T.serialize(part)
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    mod.set_option("conv_impl", impl)
    mod.init(consts)
    return part, mod, js, consts


def test_packed_run_entry_point():
    """tkb_packed_run: the TVMBackendPackedCFunc-shaped shim (include/tvm/runtime/c_backend_api.h:49-51) a graph executor
    would call: args[0] = module handle, then DLTensor handles; wrong argument counts / type codes fail with the
    reference's messages."""
    func = models.lenet5(batch=4)
    x = models.synthetic_input("lenet5", 4)
    ref, _ = interp.run_expr(func, {"data": x})
    part, mod, js, consts = _module(func)
    out = np.zeros(part.body.shape, dtype=part.body.dtype)
    mod.packed_run(x, out)
    np.testing.assert_equal(out, ref)
    with pytest.raises(RT.TachikomaError, match="mismatch in the number of provided data entries"):
        mod.packed_run(x)
    with pytest.raises(RT.TachikomaError, match="dtype code"):
        mod.packed_run(x.astype(np.int8), out)


def test_two_threads_two_streams_different_trace_modes():
    """Run(args) const is thread-safe in the reference (tachikoma_json_runtime.cc:76-90): two host threads call tkb_run
    on the same module concurrently, each on its own stream and with its own trace mode; outputs and both traces
    (streamed and deferred read-out) equal the oracle's, iteration after iteration."""
    import threading
    import torch
    func = models.resnet18(batch=2, image=64, num_classes=100)
    part, mod, js, consts = _module(func)
    xs = [np.ascontiguousarray(models.synthetic_input("resnet18", 2, seed=11 + t)[:, :, :64, :64]) for t in range(2)]
    want = []
    for x in xs:
        outs, recs = interp.run_json(js, consts, [x], want_trace=True)
        want.append((outs[0], interp.save_ndarray_list(recs)))
    streams = [torch.cuda.Stream() for _ in range(2)]
    errors = []

    def worker(t):
        try:
            mode = 2 - t  # thread 0 streams the trace to the host, thread 1 keeps it on the device
            for it in range(6):
                out = np.zeros(part.body.shape, dtype=part.body.dtype)
                mod.run(xs[t], out, stream=streams[t].cuda_stream, trace_mode=mode)
                assert (out == want[t][0]).all(), "output differs (thread %d, iteration %d)" % (t, it)
                if it % 2 == 1:
                    assert mod.trace_blob(stream=streams[t].cuda_stream) == want[t][1], "trace differs (thread %d)" % t
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    th = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errors, errors


def test_two_batches_in_flight_from_one_thread():
    """What bench.py's e2e leg does: ONE host thread alternates two streams with host inputs and host outputs.  tkb_run
    returns once the logits are on the host while the trace keeps streaming; before a stream is reused the caller waits
    for that stream's trace (tkb_trace_sync_on).  Every batch's logits and every trace equal the oracle's."""
    import torch
    func = models.resnet18(batch=2, image=64, num_classes=100)
    part, mod, js, consts = _module(func)
    xs = [np.ascontiguousarray(models.synthetic_input("resnet18", 2, seed=21 + t)[:, :, :64, :64]) for t in range(3)]
    want = []
    for x in xs:
        outs, recs = interp.run_json(js, consts, [x], want_trace=True)
        want.append((outs[0], interp.save_ndarray_list(recs)))
    streams = [torch.cuda.Stream() for _ in range(2)]
    outs = [np.zeros(part.body.shape, dtype=part.body.dtype) for _ in range(2)]
    fed = [None, None]
    for i in range(9):
        k, j = i % 2, i % 3
        sk = streams[k].cuda_stream
        if fed[k] is not None:  # the trace of the batch this stream ran before: complete, then checked, then overwritten
            mod.trace_sync(stream=sk)
            assert mod.trace_blob(stream=sk) == want[fed[k]][1], "trace of batch %d differs (step %d)" % (fed[k], i)
        mod.run(xs[j], outs[k], stream=sk, trace_mode=2)
        assert (outs[k] == want[j][0]).all(), "logits differ (step %d)" % i
        fed[k] = j
    for k in range(2):
        assert mod.trace_blob(stream=streams[k].cuda_stream) == want[fed[k]][1]


def test_device_resident_trace_deferred_readout():
    """Trace mode 1 (what bench.py's `value` runs): nothing crosses PCIe during the run; tkb_trace_blob afterwards reads
    the ring / arena out and the bytes equal the streamed (mode 2) trace and the oracle's."""
    import torch
    func = models.resnet18(batch=2, image=64, num_classes=100)
    x = np.ascontiguousarray(models.synthetic_input("resnet18", 2)[:, :, :64, :64])
    part, mod, js, consts = _module(func)
    _, recs = interp.run_json(js, consts, [x], want_trace=True)
    want = interp.save_ndarray_list(recs)
    xd = torch.from_numpy(x).cuda()
    od = torch.empty(tuple(part.body.shape), dtype=torch.uint8, device="cuda")
    for mode in (1, 2, 1):
        mod.run(xd, od, trace_mode=mode)
        assert mod.trace_blob() == want
    mod.run(xd, od, trace_mode=0)
    with pytest.raises(RT.TachikomaError, match="no trace captured"):
        mod.trace_blob()


def test_debug_executor_compatible_dump(tmp_path):
    """output_tensors.params + _tvmdbg_graph_dump.json in the debug executor's format
    (python/tvm/contrib/debugger/debug_result.py:25, :114-149, :191-202); the pair is loaded back by the replay
    verifier's reader and every kernel node's output equals the oracle's."""
    import json
    from oracle import replay
    from tachikoma_b200 import trace as TR
    func = models.resnet18(batch=2, image=64, num_classes=100)
    x = np.ascontiguousarray(models.synthetic_input("resnet18", 2)[:, :, :64, :64])
    out, mod, (js, consts, ins) = run_gpu(func, {"data": x}, conv_impl="tc", trace=True)
    mod.trace_dump_debug(tmp_path)
    graph, tensors = replay.load_debug_dump(tmp_path)
    g = json.loads(js)
    assert [n["name"] for n in graph["nodes"]] == [n["name"] for n in g["nodes"]]
    assert len(tensors) == graph["node_row_ptr"][-1]
    _, recs = interp.run_json(js, consts, ins, want_trace=True)
    want = {n: a for n, a in recs if "____output-num:" in n}
    for nid, n in enumerate(graph["nodes"]):
        key = "%s____topo-index:%d____output-num:0" % (n["name"], nid)
        assert key in tensors
        assert n["shape"] == list(tensors[key].shape)
        if n["op"] != "param":
            np.testing.assert_equal(tensors[key], want[key])
            assert n["inputs"] == [g["nodes"][i[0]]["name"] for i in g["nodes"][nid]["inputs"]]
    assert replay.verify_debug_dump(js, tmp_path, consts) == []


# ---- BASELINE.json configurations at their stated sizes ("the benchmarked configuration has never had
# its trace compared") ---------------------------------------------------------------------------------------------
def _full_size_trace_check(name, batch, impl="tc"):
    func = models.MODELS[name](batch=batch)
    x = models.synthetic_input(name, batch)
    out, mod, (js, consts, ins) = run_gpu(func, {"data": x}, conv_impl=impl, trace=True)
    outs, recs = interp.run_json(js, consts, ins, want_trace=True)
    np.testing.assert_equal(out, outs[0])
    want = interp.save_ndarray_list(recs)
    got = mod.trace_blob()
    assert len(got) == len(want)
    assert got == want
    return mod


def test_resnet18_b64_224_trace_equals_oracle():
    """C3 exactly as bench.py times it: batch 64, 224 x 224, full per-op trace; logits and every serialized trace
    byte (437 MB) equal the oracle's."""
    mod = _full_size_trace_check("resnet18", 64)
    assert mod.plan_summary().count("tcgen05") >= 21


def test_mobilenetv2_b8_224_trace_equals_oracle():
    """C4 at 224 x 224 (batch 8 bounds the oracle's time; per-image kernels do not depend on the batch beyond grid
    sizes, which test_mobilenetv2_b128_properties covers)."""
    _full_size_trace_check("mobilenetv2", 8)


def test_mobilenetv2_b128_properties():
    """C4 at BASELINE batch 128: tc == simt bit for bit on the whole batch (two independent implementations of every
    conv), and images [0:4] and [124:128] equal the oracle run on those images alone (inferences are independent)."""
    func = models.mobilenetv2(batch=128)
    x = models.synthetic_input("mobilenetv2", 128)
    a, _, _ = run_gpu(func, {"data": x}, conv_impl="tc")
    b, _, _ = run_gpu(func, {"data": x}, conv_impl="simt")
    np.testing.assert_equal(a, b)
    f4 = models.mobilenetv2(batch=4)
    for lo in (0, 124):
        ref, _ = interp.run_expr(f4, {"data": x[lo:lo + 4]})
        np.testing.assert_equal(a[lo:lo + 4], ref)


@pytest.mark.parametrize("name", ["mlp", "lenet5"])
def test_c1_c2_batch_256_trace_equals_oracle(name):
    _full_size_trace_check(name, 256)


# ---- qnn.mul / qnn.concatenate / qnn.quantize / qnn.dequantize beyond the reference's literals
@pytest.mark.parametrize("dt", ["uint8", "int8"])
@pytest.mark.parametrize("shape", [(2, 16, 14, 14), (3, 5, 7), (4096,)])
def test_qnn_mul_vs_oracle(shape, dt):
    rng = np.random.default_rng(len(shape))
    lo, hi = (0, 256) if dt == "uint8" else (-128, 128)
    a, b = rng.integers(lo, hi, shape).astype(dt), rng.integers(lo, hi, shape).astype(dt)
    va, vb = R.var("a", shape, dt), R.var("b", shape, dt)
    for (sa, za, sb, zb, so, zo) in [(0.0156863, 127 + lo, 0.0117647, 85 + lo, 0.0235294, 128 + lo), (0.5, lo, 0.25, lo, 0.125, lo),
                                     (0.004, 3 + lo, 0.004, 250 + lo, 0.00001, 7 + lo)]:  # the last one saturates
        func = R.Function([va, vb], R.qnn_mul(va, vb, np.float32(sa), za, np.float32(sb), zb, np.float32(so), zo))
        ref, _ = interp.run_expr(func, {"a": a, "b": b})
        out, mod, _ = run_gpu(func, {"a": a, "b": b})
        assert "qnn.mul" in mod.plan_summary()
        np.testing.assert_equal(out, ref)


@pytest.mark.parametrize("axis", [0, 1, -1])
def test_qnn_concatenate_vs_oracle(axis):
    rng = np.random.default_rng(5 + axis)
    shapes = [(2, 16, 6, 4), (2, 16, 6, 4), (2, 16, 6, 4)]
    xs = [rng.integers(0, 256, sh).astype("uint8") for sh in shapes]
    vs = [R.var("x%d" % i, sh, "uint8") for i, sh in enumerate(shapes)]
    # one input shares the output qparams (plain copy), two are requantized
    e = R.qnn_concatenate(vs, [np.float32(0.5), np.float32(0.25), np.float32(0.7)], [3, 9, 128], np.float32(0.25), 9, axis)
    func = R.Function(vs, e)
    ins = {"x%d" % i: x for i, x in enumerate(xs)}
    ref, _ = interp.run_expr(func, ins)
    out, mod, _ = run_gpu(func, ins)
    assert "copy into the output slice" in mod.plan_summary() and "requantize into the output slice" in mod.plan_summary()
    np.testing.assert_equal(out, ref)


@pytest.mark.parametrize("out_dtype", ["uint8", "int8", "int32"])
def test_quantize_dequantize_vs_oracle(out_dtype):
    """Float edges of a graph: ties (x / scale exactly .5: llvm.round goes away from zero), saturation, per-channel."""
    rng = np.random.default_rng(11)
    x = (rng.standard_normal((2, 6, 5, 8)) * 40).astype("float32")
    x.reshape(-1)[:8] = [0.25, -0.25, 0.75, -0.75, 1.25, -1.25, 1e9, -1e9]  # ties at scale 0.5, and saturation
    d = R.var("d", x.shape, "float32")
    zp = 0 if out_dtype != "uint8" else 120
    for scale, z, axis in [(np.float32(0.5), zp, -1),
                           (np.array([0.5, 0.25, 0.125, 1.0, 2.0, 0.3], "float32"), np.array([zp] * 6, "int32") + np.arange(6, dtype="int32"), 1)]:
        q = R.qnn_quantize(d, scale, z, axis=axis, out_dtype=out_dtype)
        func = R.Function([d], q)
        ref, _ = interp.run_expr(func, {"d": x})
        out, _, _ = run_gpu(func, {"d": x})
        np.testing.assert_equal(out, ref)
        qv = R.var("q", x.shape, out_dtype)
        f2 = R.Function([qv], R.qnn_dequantize(qv, scale, z, axis=axis))
        ref2, _ = interp.run_expr(f2, {"q": ref})
        out2, _, _ = run_gpu(f2, {"q": ref})
        assert out2.dtype == np.float32
        np.testing.assert_equal(out2, ref2)  # float32 bit patterns: the reference's single multiply, no fused ops


def test_float_edges_around_the_integer_chain():
    """quantize -> qnn.conv2d composite -> dequantize: a graph with float edges is ingested whole and its trace holds the
    float32 records too."""
    rng = np.random.default_rng(2)
    x = rng.standard_normal((2, 16, 12, 12)).astype("float32")
    d = R.var("d", x.shape, "float32")
    q = R.qnn_quantize(d, np.float32(0.05), 128, out_dtype="uint8")
    w = rng.integers(-127, 128, (32, 16, 3, 3)).astype("int8")
    e = R.qnn_conv2d(q, R.const(w), 128, 0, np.float32(0.05), np.float32(0.01), (3, 3), 32, padding=(1, 1))
    e = R.bias_add(e, R.const(rng.integers(-1000, 1000, 32).astype("int32")))
    e = R.qnn_requantize(e, np.float32(0.0005), 0, np.float32(0.1), 0, axis=1, out_dtype="int32")
    e = R.cast(R.clip(e, 0.0, 255.0), "uint8")
    e = R.qnn_dequantize(e, np.float32(0.1), 0)
    func = R.Function([d], e)
    ref, _ = interp.run_expr(func, {"d": x})
    out, mod, (js, consts, ins) = run_gpu(func, {"d": x}, conv_impl="tc", trace=True)
    np.testing.assert_equal(out, ref)
    _, recs = interp.run_json(js, consts, ins, want_trace=True)
    assert mod.trace_blob() == interp.save_ndarray_list(recs)


def test_int8_peak_probe_runs():
    tops = RT.probe_int8_peak(2000)
    assert tops > 100.0, tops


def test_unsupported_op_fails_at_init_not_run():
    # an operator outside the reference's dispatch (tachikoma_json_runtime.cc:255-286): same message, at init
    xv = R.var("x", (1, 4), "float32")
    part = R.Function([xv], R.Call("nn.log_softmax", [xv], {}, (1, 4), "float32"))
    sym, js, names, consts = T.serialize(part)
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    with pytest.raises(RT.TachikomaError, match="Unsupported op: nn.log_softmax"):
        mod.init(consts)
    # a float operator on integer tensors is not the float path either
    xi = R.var("x", (1, 4), "int32")
    part = R.Function([xi], R.Call("nn.softmax", [xi], {"axis": -1}, (1, 4), "int32"))
    sym, js, names, consts = T.serialize(part)
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    with pytest.raises(RT.TachikomaError, match="Unsupported op: nn.softmax"):
        mod.init(consts)


# ---- N3, second part: qnn.leaky_relu and the table-lookup unary operators ------------------------------------------
@pytest.mark.parametrize("dtype", ["uint8", "int8"])
def test_table_lookup_unary_ops_on_gpu(dtype):
    """qnn.{sqrt,rsqrt,exp,erf,sigmoid,hardswish,tanh,log,abs} -> take(table, reinterpret(x)) -> lut8_kernel: every bit
    pattern, ragged and 16-byte-multiple sizes, equal to the oracle's element-by-element evaluation."""
    rng = np.random.default_rng(5)
    for shape in [(4, 64), (3, 5, 7), (2, 16, 12, 12)]:
        x = rng.integers(0, 256, shape).astype(np.uint8).view(dtype)
        x.reshape(-1)[:256] = np.arange(256, dtype=np.uint8).view(dtype)[:x.size]
        for op in R.QNN_UNARY_OPS:
            v = R.var("in_0", shape, dtype)
            f = R.Function([v], R.qnn_unary(op, v, 0.05, 3, 0.02, 7 if dtype == "uint8" else -4))
            want, _ = interp.run_expr(f, {"in_0": x})
            out, mod, _ = run_gpu(f, {"in_0": x}, trace=True)
            np.testing.assert_equal(out, want)
            assert "table lookup" in mod.plan_summary()


def test_unary_lut_inside_a_convolution_chain_on_gpu():
    """conv -> requantize -> clip -> cast -> qnn.tanh -> conv: the lookup runs on the channels-last tensor between two
    tensor-core convolutions; logits and trace equal the oracle's."""
    rng = np.random.default_rng(6)
    x = rng.integers(0, 256, (2, 16, 12, 12)).astype(np.uint8)
    v = R.var("in_0", x.shape, "uint8")

    def conv(e, oc, zp_in, s_in, s_out, zp_out):
        w = R.const(rng.integers(-20, 20, (oc, e.shape[1], 3, 3)).astype("int8"))
        c = R.qnn_conv2d(e, w, zp_in, 0, s_in, 0.01, (3, 3), oc, padding=(1, 1))
        c = R.qnn_requantize(c, np.float32(s_in * 0.01), 0, s_out, zp_out, axis=1, out_dtype="int32")
        return R.cast(R.clip(c, 0.0, 255.0), "uint8")
    e = conv(v, 32, 128, 0.02, 0.05, 128)
    e = R.qnn_unary("qnn.tanh", e, 0.05, 128, 1.0 / 128, 128)
    e = conv(e, 16, 128, 1.0 / 128, 0.04, 0)
    f = R.Function([v], e)
    want, _ = interp.run_expr(f, {"in_0": x})
    for impl in ("auto", "tc"):
        out, mod, (js, consts, ins) = run_gpu(f, {"in_0": x}, conv_impl=impl, trace=True)
        np.testing.assert_equal(out, want)
        _, recs = interp.run_json(js, consts, ins, want_trace=True)
        assert mod.trace_blob() == interp.save_ndarray_list(recs)


@pytest.mark.parametrize("dtype", ["uint8", "int8"])
def test_qnn_leaky_relu_on_gpu(dtype):
    rng = np.random.default_rng(7)
    for shape in [(1, 256), (3, 5, 7), (2, 16, 12, 12)]:
        x = rng.integers(0, 256, shape).astype(np.uint8).view(dtype)
        x.reshape(-1)[:256] = np.arange(256, dtype=np.uint8).view(dtype)[:x.size]
        zi, zo = (60, 17) if dtype == "uint8" else (-20, 9)
        for alpha, si, so, z_in, z_out in [(0.9, 0.125, 0.6, zi, zo), (0.1, 0.05, 0.05, zi, zi), (0.5, 0.2, 0.1, zi, zo),
                                           (0.25, 0.25, 0.125, zi, zo)]:
            v = R.var("in_0", shape, dtype)
            f = R.Function([v], R.qnn_leaky_relu(v, alpha, si, z_in, so, z_out))
            want, _ = interp.run_expr(f, {"in_0": x})
            out, _, _ = run_gpu(f, {"in_0": x}, trace=True)
            np.testing.assert_equal(out, want)


# ---- randomized sweep: conv / dense chains with every attribute drawn at random -----------------------------------
def _random_chain(seed):
    """One random fused chain: qnn.conv2d | qnn.dense (+ bias) -> requantize -> clip -> cast [-> qnn.add -> clip]."""
    rng = np.random.default_rng(1000 + seed)
    pick = lambda xs: xs[int(rng.integers(0, len(xs)))]
    xdt, odt = pick(["uint8", "uint8", "int8"]), pick(["uint8", "uint8", "int8"])
    xi, oi = np.iinfo(xdt), np.iinfo(odt)
    zp_x = int(rng.integers(xi.min, xi.max + 1)) if rng.random() < 0.6 else 0
    per_channel = rng.random() < 0.6
    if seed % 4 == 3:  # dense
        m, k, n = int(rng.integers(1, 70)), pick([1, 7, 16, 48, 100, 256]), pick([1, 5, 16, 64, 100])
        x = rng.integers(xi.min, xi.max + 1, (m, k)).astype(xdt)
        w = rng.integers(-127, 128, (n, k)).astype("int8")
        v = R.var("in_0", x.shape, xdt)
        sw = (rng.uniform(0.5, 1.5, n) / (127 * np.sqrt(k))).astype("float32") if per_channel else np.float32(0.7 / (127 * np.sqrt(k)))
        e = R.qnn_dense(v, R.const(w), zp_x, 0, np.float32(0.05), R.const(sw) if per_channel else sw, n)
        axis, oc, bshape = -1, n, (n,)
    else:
        kh, kw = pick([1, 1, 2, 3, 3, 5]), pick([1, 1, 2, 3, 3, 5])
        dil, st = (pick([1, 1, 2]), pick([1, 1, 2])), (pick([1, 1, 2, 3]), pick([1, 1, 2, 3]))
        pad = (int(rng.integers(0, 3)), int(rng.integers(0, 3)), int(rng.integers(0, 3)), int(rng.integers(0, 3)))
        c = pick([1, 3, 4, 8, 16, 24, 32, 64, 96, 128])
        groups = pick([1, 1, 1, c]) if c > 1 else 1
        oc = pick([1, 4, 16, 32, 64, 100, 128, 192]) if groups == 1 else c * pick([1, 2])
        h = int(rng.integers((kh - 1) * dil[0] + 1, (kh - 1) * dil[0] + 18))
        wd = int(rng.integers((kw - 1) * dil[1] + 1, (kw - 1) * dil[1] + 18))
        nb = int(rng.integers(1, 5))
        x = rng.integers(xi.min, xi.max + 1, (nb, c, h, wd)).astype(xdt)
        w = rng.integers(-127, 128, (oc, c // groups, kh, kw)).astype("int8")
        kk = kh * kw * c // groups
        sw = (rng.uniform(0.5, 1.5, oc) / (127 * np.sqrt(kk))).astype("float32") if per_channel else np.float32(0.7 / (127 * np.sqrt(kk)))
        v = R.var("in_0", x.shape, xdt)
        e = R.qnn_conv2d(v, R.const(w), zp_x, 0, np.float32(0.05), R.const(sw) if per_channel else sw, (kh, kw), oc,
                         strides=st, padding=pad, dilation=dil, groups=groups)
        axis, bshape = 1, (oc,)
    if rng.random() < 0.7:
        e = R.bias_add(e, R.const(rng.integers(-3000, 3000, bshape).astype("int32")), axis=axis)
    rq_in = (np.float32(0.05) * np.asarray(sw, dtype=np.float32)).astype("float32")
    zp_o = int(rng.integers(oi.min, oi.max + 1)) if rng.random() < 0.5 else 0
    e = R.qnn_requantize(e, R.const(rq_in) if per_channel else np.float32(rq_in), 0, np.float32(pick([0.02, 0.05, 0.11])), zp_o,
                         axis=axis, out_dtype="int32")
    lo = float(pick([oi.min, oi.min, max(oi.min, zp_o)]))
    e = R.cast(R.clip(e, lo, float(oi.max)), odt)
    ins = {"in_0": x}
    params = [v]
    if rng.random() < 0.4:
        r = R.var("in_1", e.shape, odt)
        ins["in_1"] = rng.integers(oi.min, oi.max + 1, e.shape).astype(odt)
        params.append(r)
        e = R.qnn_add(e, r, np.float32(pick([0.02, 0.05])), zp_o, np.float32(pick([0.02, 0.03, 0.05])),
                      int(rng.integers(oi.min, oi.max + 1)), np.float32(pick([0.02, 0.06])), int(rng.integers(oi.min, oi.max + 1)))
        e = R.clip(e, float(oi.min), float(oi.max))
    return R.Function(params, e), ins


@pytest.mark.parametrize("seed", range(96))
def test_randomized_chain_sweep(seed):
    func, ins = _random_chain(seed)
    want, _ = interp.run_expr(func, ins)
    for impl in ("auto", "tc", "simt"):
        out, mod, (js, consts, xs) = run_gpu(func, ins, conv_impl=impl, trace=(impl == "tc"))
        np.testing.assert_equal(out, want, err_msg=f"seed {seed} impl {impl}: {mod.plan_summary()}")
        if impl == "tc":
            _, recs = interp.run_json(js, consts, xs, want_trace=True)
            assert mod.trace_blob() == interp.save_ndarray_list(recs)


# ---- cta_group::2 pair kernel (kernels_tc_pair.cu), opt-in through conv_impl = tc_pair ----------------------------
PAIR_CASES = {
    # name: (N, C, H, O, k, stride, pad, zp_in, with_sum)
    "l3_3x3": (2, 256, 14, 256, 3, 1, 1, 0, False),
    "l3_3x3_sum": (3, 256, 14, 256, 3, 1, 1, 0, True),
    "l2_3x3_bn128": (4, 128, 28, 128, 3, 1, 1, 0, True),
    "l4_3x3_partial_pair": (2, 512, 7, 512, 3, 1, 1, 0, False),
    "l3_1x1_gemm": (2, 256, 14, 512, 1, 1, 0, 0, False),
    "l3_3x3_s2": (2, 128, 28, 256, 3, 2, 1, 0, False),
    "l3_1x1_s2": (2, 128, 28, 256, 1, 2, 0, 0, True),
    "zp_border": (2, 128, 12, 128, 3, 1, 1, 7, False),
    "many_tiles": (16, 128, 28, 256, 3, 1, 1, 0, True),
}


@pytest.mark.parametrize("name", list(PAIR_CASES))
def test_pair_kernel_vs_oracle(name):
    n, c, h, o, k, st, pad, zp, with_sum = PAIR_CASES[name]
    rng = np.random.default_rng(11)
    x = rng.integers(0, 256, (n, c, h, h)).astype("uint8")
    w = rng.integers(-127, 128, (o, c, k, k)).astype("int8")
    sw = (rng.uniform(0.5, 1.5, o) / (127 * np.sqrt(k * k * c))).astype("float32")
    v = R.var("in_0", x.shape, "uint8")
    # a leading max pool keeps the convolution's input channels-last inside the runtime (no boundary conversion in between)
    e = R.max_pool2d(v, (1, 1), (1, 1), (0, 0))
    e = R.qnn_conv2d(e, R.const(w), zp, 0, np.float32(0.02), R.const(sw), (k, k), o, strides=(st, st), padding=(pad, pad))
    e = R.bias_add(e, R.const(rng.integers(-4000, 4000, o).astype("int32")))
    e = R.qnn_requantize(e, R.const((np.float32(0.02) * sw).astype("float32")), 0, np.float32(0.03), 0, axis=1, out_dtype="int32")
    e = R.cast(R.clip(e, 0.0, 255.0), "uint8")
    ins, params = {"in_0": x}, [v]
    if with_sum:
        r = R.var("in_1", e.shape, "uint8")
        ins["in_1"] = rng.integers(0, 256, e.shape).astype("uint8")
        params.append(r)
        e = R.clip(R.qnn_add(e, r, np.float32(0.03), 0, np.float32(0.05), 3, np.float32(0.04), 0), 0.0, 255.0)
    f = R.Function(params, e)
    want, _ = interp.run_expr(f, ins)
    out, mod, _ = run_gpu(f, ins, conv_impl="tc_pair", trace=True)
    assert "cta_group::2" in mod.plan_summary(), mod.plan_summary()
    np.testing.assert_equal(out, want)


@pytest.mark.parametrize("dt", ["uint8", "int8"])
def test_qnn_subtract_vs_oracle(dt):
    """qnn.subtract (subtract.cc:40-93) over every pair of byte values, equal and different quantization parameters."""
    info = np.iinfo(dt)
    a = np.repeat(np.arange(info.min, info.max + 1), 256).astype(dt).reshape(256, 256)
    b = np.tile(np.arange(info.min, info.max + 1), 256).astype(dt).reshape(256, 256)
    for qp in [(0.5, 3, 0.25, 7, 0.3, 5), (0.2, 0, 0.2, 0, 0.2, 0), (0.1, 4, 0.3, 9, 0.1, 4), (0.37, -2 if dt == "int8" else 2, 0.11, 1, 0.5, 0)]:
        x, y = R.var("in_0", a.shape, dt), R.var("in_1", b.shape, dt)
        f = R.Function([x, y], R.qnn_subtract(x, y, *qp))
        want, _ = interp.run_expr(f, {"in_0": a, "in_1": b})
        out, mod, _ = run_gpu(f, {"in_0": a, "in_1": b}, trace=True)
        np.testing.assert_equal(out, want)
        assert "qnn.subtract" in mod.plan_summary()


@pytest.mark.parametrize("oc", [8, 16, 24, 40, 48, 72, 144, 200])
@pytest.mark.parametrize("with_sum", [False, True])
def test_ragged_output_channels_on_the_vector_epilogue(oc, with_sum):
    """Output-channel counts that are not multiples of 32 (MobileNetV2 projections: 16, 24, 144 ...): the last chunk of a
    tile is stored as whole 16-byte (8-byte) vectors; 1x1 and 3x3 convolutions, uint8 and int8 outputs, many rows so that
    every alignment of a 24-byte row occurs; also a channel count that still needs the element path (200 = 192 + 8 with a
    residual)."""
    rng = np.random.default_rng(oc)
    for k, odt in ((1, "uint8"), (3, "int8")):
        c = 32
        x = rng.integers(0, 256, (3, c, 9, 11)).astype("uint8")
        w = rng.integers(-127, 128, (oc, c, k, k)).astype("int8")
        sw = (rng.uniform(0.5, 1.5, oc) / (127 * np.sqrt(k * k * c))).astype("float32")
        v = R.var("in_0", x.shape, "uint8")
        e = R.max_pool2d(v, (1, 1), (1, 1), (0, 0))
        e = R.qnn_conv2d(e, R.const(w), 0, 0, np.float32(0.02), R.const(sw), (k, k), oc, padding=(k // 2, k // 2))
        e = R.bias_add(e, R.const(rng.integers(-4000, 4000, oc).astype("int32")))
        zo = 0 if odt == "uint8" else -3
        e = R.qnn_requantize(e, R.const((np.float32(0.02) * sw).astype("float32")), 0, np.float32(0.03), zo, axis=1, out_dtype="int32")
        info = np.iinfo(odt)
        e = R.cast(R.clip(e, float(info.min), float(info.max)), odt)
        ins, params = {"in_0": x}, [v]
        if with_sum:
            r = R.var("in_1", e.shape, odt)
            ins["in_1"] = rng.integers(info.min, info.max + 1, e.shape).astype(odt)
            params.append(r)
            e = R.clip(R.qnn_add(e, r, np.float32(0.03), zo, np.float32(0.05), 3, np.float32(0.04), 1), float(info.min), float(info.max))
        f = R.Function(params, e)
        want, _ = interp.run_expr(f, ins)
        for impl in ("tc", "tc_im2col"):
            out, mod, _ = run_gpu(f, ins, conv_impl=impl, trace=True)
            np.testing.assert_equal(out, want, err_msg=f"oc={oc} k={k} {odt} sum={with_sum} {impl}")


# ---- pointwise convolutions with 8 .. 32 input channels: CUDA-core dp4a kernel in the default plan -----------------------------
PW_CASES = {
    # name: (n, ic, h, w, oc, zp_a, per_channel, relu, xdt, out_dt, clip)
    "mbv2_32_16": (2, 32, 20, 20, 16, 0, True, False, "uint8", "uint8", None),
    "mbv2_16_96": (2, 16, 20, 20, 96, 0, True, True, "uint8", "uint8", None),
    "mbv2_24_144_zp": (1, 24, 17, 13, 144, 128, True, True, "uint8", "uint8", (0.0, 255.0)),
    "k8_n8_ragged_pixels": (3, 8, 13, 17, 8, 3, False, True, "uint8", "uint8", None),     # 663 pixels: a partial pixel group
    "k32_n24_s8": (3, 32, 9, 11, 24, 0, True, False, "int8", "int8", None),
    "k16_n40_explicit_clamp": (1, 16, 12, 12, 40, 7, True, True, "uint8", "uint8", (20.0, 200.0)),
    "k24_n1024": (1, 24, 6, 6, 1024, 0, True, False, "uint8", "int8", None),
}


@pytest.mark.parametrize("nhwc", [False, True], ids=["nchw", "nhwc"])
@pytest.mark.parametrize("name", list(PW_CASES))
def test_pointwise_few_channel_conv_vs_oracle(name, nhwc, monkeypatch):
    monkeypatch.setenv("TKB_PW_MIN_M", "1")  # the default plan takes this kernel from 2^18 pixels up (read at plan time)
    n, ic, h, w, oc, zp_a, pc, relu, xdt, odt, clip = PW_CASES[name]
    func, args = _chain_conv(n, ic, h, w, oc, 1, 1, 0, 1, zp_a, pc, relu, False, seed=len(name) * 13 + oc, nhwc=nhwc, xdt=xdt,
                             out_dt=odt, clip=clip)
    ref, _ = interp.run_expr(func, args)
    out, mod, (js, consts, ins) = run_gpu(func, args, conv_impl="auto", trace=True)
    np.testing.assert_equal(out, ref)
    want, recs = interp.run_json(js, consts, ins, want_trace=True)
    assert mod.trace_blob() == interp.save_ndarray_list(recs)
    out_tc, mod_tc, _ = run_gpu(func, args, conv_impl="tc")
    np.testing.assert_equal(out_tc, ref)
    assert "tcgen05" in mod_tc.plan_summary()
    if nhwc:  # a graph input in logical NCHW form is gathered from NCHW instead (explicit im2col path)
        assert "CUDA-core pointwise conv" in mod.plan_summary(), mod.plan_summary()


def test_mobilenet_default_plan_uses_the_pointwise_kernel_and_matches_tc_plan(monkeypatch):
    monkeypatch.setenv("TKB_PW_MIN_M", "1")
    from tachikoma_b200 import models
    func = models.mobilenetv2(batch=2, image=64, num_classes=100)
    x = np.random.default_rng(3).integers(0, 256, (2, 3, 64, 64)).astype("uint8")
    a, mod_a, _ = run_gpu(func, {"data": x}, conv_impl="auto", trace=True)
    b, mod_b, _ = run_gpu(func, {"data": x}, conv_impl="tc", trace=True)
    np.testing.assert_equal(a, b)
    assert mod_a.trace_blob() == mod_b.trace_blob()
    assert mod_a.plan_summary().count("CUDA-core pointwise conv") >= 3


@pytest.mark.parametrize("dtype", ["uint8", "int8"])
@pytest.mark.parametrize("layout", ["NCHW", "NHWC"])
def test_upcast_global_average_pool_tail_in_one_launch(dtype, layout):
    """cast -> nn.global_avg_pool2d -> cast (qnn_torch.py:587-590) fused by the planner; not fused when the int32 tensor has a
    second consumer.  Sums above 8 bits, negative means (truncation toward zero), traces."""
    rng = np.random.default_rng(11)
    shape = (3, 20, 7, 5) if layout == "NCHW" else (3, 7, 5, 20)
    lo, hi = (0, 256) if dtype == "uint8" else (-128, 128)
    x = rng.integers(lo, hi, shape).astype(dtype)
    v = R.var("x", shape, dtype)
    up = R.cast(v, "int32")
    tail = R.cast(R.global_avg_pool2d(up, layout=layout), dtype)
    f = R.Function([v], tail)
    ref, _ = interp.run_expr(f, {"x": x})
    out, mod, (js, consts, ins) = run_gpu(f, {"x": x}, trace=True)
    np.testing.assert_equal(out, ref)
    plan = mod.plan_summary()
    assert "in one launch" in plan and "pooling kernel" not in plan and "cast -> elementwise" not in plan
    _, recs = interp.run_json(js, consts, ins, want_trace=True)
    assert mod.trace_blob() == interp.save_ndarray_list(recs)
    # untraced run of the same module: same output (the int32 copy is not written then)
    out2 = np.empty_like(out)
    mod.run(x, out2, trace_mode=0)
    np.testing.assert_equal(out2, ref)
    # two consumers of the int32 tensor: three separate kernels
    g = R.Function([v], R.Tuple([tail, up]))
    part = T.partition_for_tachikoma(g)
    sym, js2, names, consts2 = T.serialize(part)
    m2 = RT.TachikomaJSONRuntime(sym, js2, names)
    m2.init(consts2)
    o0, o1 = np.empty(tail.shape, dtype), np.empty(shape, "int32")
    m2.run(x, o0, o1)
    np.testing.assert_equal(o0, ref)
    np.testing.assert_equal(o1, x.astype("int32"))
    assert "in one launch" not in m2.plan_summary()
