"""Relay text format (tachikoma_b200/relay_text.py): print -> parse round trips reproduce the
graph exactly (same tachikoma JSON, same constants, same oracle result), and text in the shape TVM's RelayTextPrinter
emits (src/printer/relay_text_printer.cc, scalar suffixes of src/support/scalars.cc:95-124, `/* ty=... */` comments)
parses into the expected chain."""
import numpy as np
import pytest

from oracle import interp
from tachikoma_b200 import models, relay as R, relay_text as RT_, tachikoma as T


def _same_graph(f, g):
    a = T.serialize(T.partition_for_tachikoma(f))
    b = T.serialize(T.partition_for_tachikoma(g))
    assert a[1] == b[1] and a[2] == b[2]
    assert len(a[3]) == len(b[3]) and all(x.dtype == y.dtype and np.array_equal(x, y) for x, y in zip(a[3], b[3]))


@pytest.mark.parametrize("name,kw", [("mlp", dict(batch=2)), ("lenet5", dict(batch=2)), ("resnet18", dict(batch=1, image=32, num_classes=10)),
                                     ("mobilenetv2", dict(batch=1, image=32, num_classes=10)), ("bert_encoder", dict(batch=1, seq=16, hidden=64, heads=4, ffn=128))])
def test_models_round_trip_through_relay_text(name, kw):
    f = models.MODELS[name](**kw)
    text, consts = RT_.astext(f)
    assert text.startswith('#[version = "0.0.5"]\ndef @main(%') and "meta[relay.Constant][0]" in text
    g = RT_.fromtext(text, consts)
    _same_graph(f, g)
    text2, consts2 = RT_.astext(g)
    assert text2 == text and len(consts2) == len(consts)


def test_every_integer_operator_round_trips():
    rng = np.random.default_rng(0)
    x = R.var("x", (2, 8, 6, 6), "uint8")
    y = R.var("y", (2, 8, 6, 6), "uint8")
    e = R.pad(x, ((0, 0), (0, 0), (1, 1), (1, 1)), pad_value=3.0)
    e = R.qnn_conv2d(e, R.const(rng.integers(-9, 9, (8, 8, 3, 3)).astype("int8")), 3, 0, 0.0123456789, 0.5, (3, 3), 8)
    e = R.cast(R.clip(R.qnn_requantize(e, np.float32(0.0061728), 0, 0.05, 1, axis=1, out_dtype="int32"), 0.0, 255.0), "uint8")
    e = R.qnn_add(e, y, 0.05, 1, 0.1, 2, 0.07, 3)
    e = R.qnn_mul(e, y, 0.07, 3, 0.1, 2, 0.2, 0)
    e = R.qnn_concatenate([e, y], [0.2, 0.1], [0, 2], 0.15, 1, 1)
    e = R.qnn_leaky_relu(e, 0.1, 0.15, 1, 0.15, 1)
    e = R.qnn_unary("qnn.tanh", e, 0.15, 1, 1.0 / 128, 128)
    e = R.max_pool2d(e, (2, 2), (2, 2))
    e = R.cast(R.avg_pool2d(R.cast(e, "int32"), (3, 3), (1, 1)), "uint8")
    e = R.transpose(e, (0, 2, 3, 1))
    e = R.reshape(e, (2, -1))
    e = R.qnn_dense(e, R.const(rng.integers(-9, 9, (5, 16)).astype("int8")), 128, 0, 1.0 / 128, 0.3, 5)
    e = R.qnn_dequantize(R.cast(R.clip(R.qnn_requantize(e, np.float32(0.3 / 128), 0, 0.2, 0, out_dtype="int32"), -128.0, 127.0), "int8"), 0.2, 0)
    f = R.Function([x, y], e)
    text, consts = RT_.astext(f)
    g = RT_.fromtext(text, consts)
    ins = {"x": rng.integers(0, 256, (2, 8, 6, 6)).astype("uint8"), "y": rng.integers(0, 256, (2, 8, 6, 6)).astype("uint8")}
    a, _ = interp.run_expr(f, ins)
    b, _ = interp.run_expr(g, ins)
    np.testing.assert_array_equal(a, b)
    assert RT_.astext(g)[0] == text
    for frag in ("qnn.concatenate((%", "pad_width=[[0, 0], [0, 0], [1, 1], [1, 1]]", 'out_dtype="int32"', "a_max=255f", "0.0123456791f"):
        assert frag in text, frag


TVM_STYLE = """#[version = "0.0.5"]
def @main(%data: Tensor[(1, 4, 5, 5), uint8] /* ty=Tensor[(1, 4, 5, 5), uint8] */) -> Tensor[(1, 6, 5, 5), uint8] {
  %0 = nn.pad(%data, 128f /* ty=float32 */, pad_width=[[0, 0], [0, 0], [1, 1], [1, 1]]) /* ty=Tensor[(1, 4, 7, 7), uint8] */;
  %1 = qnn.conv2d(%0, meta[relay.Constant][0] /* ty=Tensor[(6, 4, 3, 3), int8] */, 128 /* ty=int32 */, 0 /* ty=int32 */, 0.015625f /* ty=float32 */, meta[relay.Constant][1] /* ty=Tensor[(6), float32] */, padding=[0, 0, 0, 0], channels=6, kernel_size=[3, 3], out_dtype="int32") /* ty=Tensor[(1, 6, 5, 5), int32] */;
  %2 = nn.bias_add(%1, meta[relay.Constant][2] /* ty=Tensor[(6), int32] */) /* ty=Tensor[(1, 6, 5, 5), int32] */;
  %3 = qnn.requantize(%2, meta[relay.Constant][3] /* ty=Tensor[(6), float32] */, 0 /* ty=int32 */, 0.05f /* ty=float32 */, 0 /* ty=int32 */, axis=1, out_dtype="int32") /* ty=Tensor[(1, 6, 5, 5), int32] */;
  %4 = clip(%3, a_min=0f, a_max=255f) /* ty=Tensor[(1, 6, 5, 5), int32] */;
  cast(%4, dtype="uint8") /* ty=Tensor[(1, 6, 5, 5), uint8] */
}
"""


def test_text_in_the_reference_printers_shape_parses_to_the_fused_chain():
    rng = np.random.default_rng(1)
    w = rng.integers(-20, 20, (6, 4, 3, 3)).astype("int8")
    sw = rng.uniform(0.001, 0.002, 6).astype("float32")
    b = rng.integers(-100, 100, 6).astype("int32")
    f = RT_.fromtext(TVM_STYLE, [w, sw, b, (np.float32(0.015625) * sw).astype("float32")])
    assert [e.op for e in R.post_order(f.body) if isinstance(e, R.Call)] == ["nn.pad", "qnn.conv2d", "nn.bias_add", "qnn.requantize", "clip", "cast"]
    assert f.params[0].shape == (1, 4, 5, 5) and f.body.shape == [1, 6, 5, 5] or tuple(f.body.shape) == (1, 6, 5, 5)
    part = T.partition_for_tachikoma(f)
    import json
    kernels = [n["name"] for n in json.loads(T.serialize(part)[1])["nodes"] if n["op"] == "kernel"]
    assert kernels == ["tachikoma.qnn.conv2d"]  # explicit zero-point pad folded, chain fused
    x = rng.integers(0, 256, (1, 4, 5, 5)).astype("uint8")
    ref = R.qnn_conv2d(R.var("data", x.shape, "uint8"), R.const(w), 128, 0, 0.015625, R.const(sw), (3, 3), 6, padding=(1, 1))
    got, _ = interp.run_expr(f, {"data": x})
    assert got.shape == (1, 6, 5, 5) and got.dtype == np.uint8
    # the same chain built directly
    v = R.var("data", x.shape, "uint8")
    e = R.qnn_conv2d(v, R.const(w), 128, 0, np.float32(0.015625), R.const(sw), (3, 3), 6, padding=(1, 1))
    e = R.bias_add(e, R.const(b))
    e = R.cast(R.clip(R.qnn_requantize(e, R.const((np.float32(0.015625) * sw).astype("float32")), 0, np.float32(0.05), 0, axis=1, out_dtype="int32"), 0.0, 255.0), "uint8")
    want, _ = interp.run_expr(R.Function([v], e), {"data": x})
    np.testing.assert_array_equal(got, want)


def test_parser_errors():
    with pytest.raises(ValueError, match="Unsupported op: nn.log_softmax"):
        RT_.fromtext("fn (%x: Tensor[(2, 3), float32]) { nn.log_softmax(%x, axis=1) }")
    assert RT_.fromtext("fn (%x: Tensor[(2, 3), float32]) { nn.softmax(%x, axis=1) }").body.op == "nn.softmax"
    with pytest.raises(ValueError, match="only 0 constants given"):
        RT_.fromtext("fn (%x: Tensor[(2, 3), int32]) { add(%x, meta[relay.Constant][0]) }")
    with pytest.raises(ValueError, match="used before its definition"):
        RT_.fromtext("fn (%x: Tensor[(2, 3), int32]) { add(%x, %7) }")


def test_float_graphs_round_trip_through_relay_text():
    """Every float test graph (tests/float_graphs.py: convolutions 1-3D and transposed, dense, batch matmul, norms, softmax,
    pools, activations incl. the gelu / swish / mish compositions, an -inf pad value) prints, parses back to the same text and
    evaluates to the same values."""
    import float_graphs as FG
    from oracle import interp
    for name, f, _ in FG.all_cases():
        text, consts = RT_.astext(f)
        g = RT_.fromtext(text, consts)
        assert RT_.astext(g)[0] == text, name
        ins = FG.inputs_for(f)
        a, _ = interp.run_expr(f, ins)
        b, _ = interp.run_expr(g, ins)
        np.testing.assert_array_equal(a, b, err_msg=name)
