"""Imported TorchScript-quantized models (tachikoma_b200/frontend.py) on the CUDA path: float32 input -> qnn.quantize ->
the integer chain -> qnn.dequantize, through the C ABI, bit-exact with the CPU oracle (float outputs included: the same
IEEE operations on the same integers) and with a full trace."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

import qtorch_models as QM  # noqa: E402
from oracle import interp  # noqa: E402
from tachikoma_b200 import frontend as F, runtime as RT, tachikoma as T  # noqa: E402

pytestmark = pytest.mark.gpu


def _run(func, name, x, conv_impl, trace=True):
    part = T.partition_for_tachikoma(func)
    sym, js, names, consts = T.serialize(part)
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    mod.set_option("conv_impl", conv_impl)
    mod.init(consts)
    out = np.empty(part.body.shape, dtype=part.body.dtype)
    mod.run(x, out, trace_mode=2 if trace else 0)
    want, recs = interp.run_json(js, consts, [x], want_trace=True)
    np.testing.assert_array_equal(out, want[0])
    if trace:
        assert mod.trace_blob() == interp.save_ndarray_list(recs)
    return out


@pytest.mark.parametrize("conv_impl", ["auto", "tc", "simt"])
def test_imported_resnet_on_gpu(conv_impl):
    shape = (8, 3, 64, 64)
    ts, _ = QM.build(QM.TinyResNet, shape)
    func = F.from_torchscript(ts, [shape], ["data"])
    x = QM.sample(shape)
    out = _run(func, "data", x, conv_impl)
    with torch.no_grad():
        want = ts(torch.from_numpy(x)).numpy()
    assert np.corrcoef(out.ravel(), want.ravel())[0, 1] > 0.999


@pytest.mark.parametrize("per_channel", [True, False])
def test_imported_cat_mul_model_on_gpu(per_channel):
    shape = (2, 4, 8, 8)
    ts, _ = QM.build(QM.CatMulNet, shape, seed=3, per_channel=per_channel)
    func = F.from_torchscript(ts, [shape], ["x"])
    _run(func, "x", QM.sample(shape), "auto")


def test_imported_mlp_on_gpu():
    shape = (5, 32)
    ts, _ = QM.build(QM.MLP, shape, seed=5)
    func = F.from_torchscript(ts, [shape], ["x"])
    _run(func, "x", QM.sample(shape), "auto")


def test_imported_activation_model_on_gpu():
    shape = (2, 3, 10, 10)
    ts, _ = QM.build(QM.ActNet, shape, seed=2)
    func = F.from_torchscript(ts, [shape], ["x"])
    for impl in ("auto", "tc"):
        _run(func, "x", QM.sample(shape), impl)
