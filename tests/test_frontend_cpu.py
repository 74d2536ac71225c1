"""Importer for prequantized TorchScript models (tachikoma_b200/frontend.py): the emitted
chains are the reference frontend's (python/tvm/relay/frontend/qnn_torch.py:649-682, :685-766), explicit zero-point
padding folds into the convolution, the imported graph partitions fully, and its oracle result agrees with torch's
own quantized CPU kernels to within one output quantum on almost every element (fbgemm requantizes with a float
multiplier, the reference lowering with the fixed-point one: the reference's own frontend tests compare the same
way, tests/python/frontend/pytorch/qnn_test.py)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

import qtorch_models as QM  # noqa: E402
from oracle import interp  # noqa: E402
from tachikoma_b200 import frontend as F, relay as R, tachikoma as T  # noqa: E402


def _ops(func):
    return [e.op for e in R.post_order(func.body) if isinstance(e, R.Call)]


@pytest.fixture(scope="module")
def resnet():
    shape = (2, 3, 64, 64)
    ts, eager = QM.build(QM.TinyResNet, shape)
    return shape, ts, eager, F.from_torchscript(ts, [shape], ["data"])


def test_import_emits_the_reference_chain(resnet):
    shape, ts, eager, func = resnet
    ops = _ops(func)
    # first convolution: quantize -> explicit pad with the input zero point -> conv(padding 0) -> bias_add ->
    # requantize(axis=1, int32) -> clip -> cast
    assert ops[:7] == ["qnn.quantize", "nn.pad", "qnn.conv2d", "nn.bias_add", "qnn.requantize", "clip", "cast"]
    calls = [e for e in R.post_order(func.body) if isinstance(e, R.Call)]
    q, pad, conv, bias, rq, clip, cast = calls[:7]
    zp_in = int(np.asarray(q.args[2].data).reshape(-1)[0])
    assert q.attrs["out_dtype"] == "uint8" and q.attrs["axis"] == 1
    assert float(np.asarray(pad.args[1].data).reshape(-1)[0]) == float(zp_in) and pad.attrs["pad_width"] == [[0, 0], [0, 0], [3, 3], [3, 3]]
    assert conv.attrs["padding"] == [0, 0, 0, 0] and conv.attrs["strides"] == [2, 2] and conv.dtype == "int32"
    assert int(np.asarray(conv.args[2].data).reshape(-1)[0]) == zp_in and conv.args[1].data.dtype == np.int8
    assert bias.args[1].data.dtype == np.int32 and bias.attrs["axis"] == 1
    assert rq.attrs["axis"] == 1 and rq.attrs["out_dtype"] == "int32" and rq.args[1].data.shape == (16,)  # per channel
    np.testing.assert_array_equal(rq.args[1].data, (np.float32(np.asarray(conv.args[4].data).reshape(-1)[0]) * conv.args[5].data).astype(np.float32))
    zp_out = int(np.asarray(rq.args[4].data).reshape(-1)[0])
    assert clip.attrs["a_min"] == float(zp_out) and clip.attrs["a_max"] == 255.0 and cast.attrs["dtype"] == "uint8"  # conv2d_relu
    assert ops.count("qnn.conv2d") == 6 and ops.count("qnn.add") == 2 and ops.count("qnn.dense") == 1
    assert ops[-1] == "qnn.dequantize" and "nn.global_avg_pool2d" in ops and "nn.max_pool2d" in ops
    # bias quantization: round(bias / (s_in * s_w)), zero point 0 (qnn_torch.py:157-164, :205-210)
    _, bf = torch.ops.quantized.conv2d_unpack(eager.c._packed_params)
    want = np.clip(np.round(bf.detach().numpy().astype(np.float64) / rq.args[1].data.astype(np.float64)), -2**31, 2**31 - 1)
    np.testing.assert_array_equal(bias.args[1].data, want.astype(np.int32))


def test_explicit_padding_folds_and_graph_partitions_fully(resnet):
    shape, ts, eager, func = resnet
    folded = F.fold_explicit_padding(func)
    assert "nn.pad" not in _ops(folded)
    pads = [e.attrs["padding"] for e in R.post_order(folded.body) if isinstance(e, R.Call) and e.op == "qnn.conv2d"]
    assert pads[0] == [3, 3, 3, 3] and [1, 1, 1, 1] in pads and [0, 0, 0, 0] in pads
    x = QM.sample(shape)
    a, _ = interp.run_expr(func, {"data": x})
    b, _ = interp.run_expr(folded, {"data": x})
    np.testing.assert_array_equal(a, b)  # pad-with-zero-point == padding attribute, bit for bit
    part = T.partition_for_tachikoma(func)
    sym, js, names, consts = T.serialize(part)
    kernels = [n["name"] for n in __import__("json").loads(js)["nodes"] if n["op"] == "kernel"]
    assert kernels.count("tachikoma.qnn.conv2d") == 6 and kernels.count("tachikoma.qnn.dense") == 1
    assert "qnn.add" not in kernels  # both residual adds fused into the producing convolution's composite
    outs, _ = interp.run_json(js, consts, [x])
    np.testing.assert_array_equal(outs[0], a)


def _compare_with_torch(ts, func, shape, name):
    x = QM.sample(shape)
    got, _ = interp.run_expr(func, {name: x})
    with torch.no_grad():
        want = ts(torch.from_numpy(x)).numpy()
    # output quantum = scale of the last quantized tensor
    last = [e for e in R.post_order(func.body) if isinstance(e, R.Call)][-1]
    scale = float(np.asarray(last.args[1].data).reshape(-1)[0])
    d = np.abs(got - want) / scale
    return d


def test_imported_resnet_agrees_with_torch_quantized_kernels(resnet):
    shape, ts, eager, func = resnet
    d = _compare_with_torch(ts, func, shape, "data")
    assert d.max() <= 3.0 + 1e-3, d.max()          # quanta; rounding differences of 9 stacked requantizations
    assert (d <= 1.0 + 1e-3).mean() >= 0.9, (d <= 1.0 + 1e-3).mean()


@pytest.mark.parametrize("per_channel", [True, False])
def test_cat_mul_avgpool_model(per_channel):
    shape = (2, 4, 8, 8)
    ts, eager = QM.build(QM.CatMulNet, shape, seed=3, per_channel=per_channel)
    func = F.from_torchscript(ts, [shape], ["x"])
    ops = _ops(func)
    assert "qnn.concatenate" in ops and "qnn.mul" in ops and "nn.avg_pool2d" in ops and ops[-1] == "qnn.dequantize"
    # quantized::conv2d without relu clips at 0, with relu at the output zero point (qnn_torch.py:677-680)
    clips = [e for e in R.post_order(func.body) if isinstance(e, R.Call) and e.op == "clip"]
    assert clips[0].attrs["a_min"] == 0.0
    groups = [e.attrs["groups"] for e in R.post_order(func.body) if isinstance(e, R.Call) and e.op == "qnn.conv2d"]
    assert groups == [1, 1, 8]
    part = T.partition_for_tachikoma(func)
    sym, js, names, consts = T.serialize(part)
    x = QM.sample(shape)
    a, _ = interp.run_expr(func, {"x": x})
    outs, _ = interp.run_json(js, consts, [x])
    np.testing.assert_array_equal(outs[0], a)
    # torch multiplies and concatenates in float and rounds its average pool, the reference lowering stays in integers
    # and truncates the pool (apply_with_upcast): up to two quanta apart, one on almost every element
    d = _compare_with_torch(ts, func, shape, "x")
    assert d.max() <= 2.0 + 1e-3 and (d <= 1.0 + 1e-3).mean() >= 0.9, (d.max(), (d <= 1.0 + 1e-3).mean())


def test_linear_relu_model():
    shape = (5, 32)
    ts, eager = QM.build(QM.MLP, shape, seed=5)
    func = F.from_torchscript(ts, [shape], ["x"])
    assert _ops(func) == ["qnn.quantize", "qnn.dense", "nn.bias_add", "qnn.requantize", "clip", "cast",
                          "qnn.dense", "nn.bias_add", "qnn.requantize", "clip", "cast", "qnn.dequantize"]
    clips = [e.attrs["a_min"] for e in R.post_order(func.body) if isinstance(e, R.Call) and e.op == "clip"]
    rqs = [e for e in R.post_order(func.body) if isinstance(e, R.Call) and e.op == "qnn.requantize"]
    assert clips == [float(np.asarray(rqs[0].args[4].data).reshape(-1)[0]), 0.0]  # linear_relu, then linear
    d = _compare_with_torch(ts, func, shape, "x")
    assert d.max() <= 1.0 + 1e-3, d.max()


def test_activation_model_leaky_relu_hardswish_sigmoid_relu6():
    shape = (2, 3, 10, 10)
    ts, eager = QM.build(QM.ActNet, shape, seed=2)
    func = F.from_torchscript(ts, [shape], ["x"])
    ops = _ops(func)
    for op in ("qnn.leaky_relu", "qnn.hardswish", "qnn.sigmoid"):
        assert op in ops
    sig = [e for e in R.post_order(func.body) if isinstance(e, R.Call) and e.op == "qnn.sigmoid"][0]
    assert float(np.asarray(sig.args[3].data).reshape(-1)[0]) == 1.0 / 256 and int(np.asarray(sig.args[4].data).reshape(-1)[0]) == 0
    part = T.partition_for_tachikoma(func)
    sym, js, names, consts = T.serialize(part)
    kernels = [n["name"] for n in __import__("json").loads(js)["nodes"] if n["op"] == "kernel"]
    assert kernels.count("take") == 2 and kernels.count("reinterpret") == 2 and "qnn.leaky_relu" in kernels
    x = QM.sample(shape)
    a, _ = interp.run_expr(func, {"x": x})
    outs, _ = interp.run_json(js, consts, [x])
    np.testing.assert_array_equal(outs[0], a)
    d = _compare_with_torch(ts, func, shape, "x")
    assert d.max() <= 2.0 + 1e-3 and (d <= 1.0 + 1e-3).mean() >= 0.95, (d.max(), (d <= 1.0 + 1e-3).mean())


def test_importer_rejects_unknown_ops():
    class M(torch.nn.Module):
        def forward(self, x):
            return torch.sin(x)
    ts = torch.jit.trace(M().eval(), torch.zeros(2, 3))
    with pytest.raises(ValueError, match="Unsupported op: aten::sin"):
        F.from_torchscript(ts, [(2, 3)])
