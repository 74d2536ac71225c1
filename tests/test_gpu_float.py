"""Float tachikoma patterns on the CUDA path: every graph of tests/float_graphs.py (the
reference's test_tachikoma.py float tests) through the C ABI, against the numpy oracle of the UNPARTITIONED graph, at the
reference's own tolerance rtol = atol = 1e-3 (tests/python/contrib/test_tachikoma.py:155); the trace holds the same tensors."""
import numpy as np
import pytest

import float_graphs as FG
from oracle import interp
from tachikoma_b200 import runtime as RT, tachikoma as T

pytestmark = pytest.mark.gpu

CASES = FG.all_cases()


def _run(func, trace_mode=0):
    part = T.partition_for_tachikoma(func)
    sym, js, names, consts = T.serialize(part)
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    mod.init(consts)
    ins = FG.inputs_for(func)
    xs = [np.ascontiguousarray(ins[p.name]) for p in func.params]
    out = np.empty(part.body.shape, dtype=part.body.dtype)
    mod.run(*xs, out, trace_mode=trace_mode)
    return mod, js, consts, xs, ins, out


@pytest.mark.parametrize("name,func,expected", CASES, ids=[c[0] for c in CASES])
def test_float_graph_on_gpu(name, func, expected):
    mod, js, consts, xs, ins, out = _run(func)
    want, _ = interp.run_expr(func, ins)
    assert np.isfinite(out).all()
    np.testing.assert_allclose(out, want, rtol=1e-3, atol=1e-3)
    # much tighter in practice: fp32 FMA chains against float64 sums
    assert np.max(np.abs(out - want) / (1e-4 + 1e-4 * np.abs(want))) < 1.0
    assert all(("fp32" in l) or ("clip" in l) or ("nn.pad" in l) for l in mod.plan_summary().splitlines()[1:] if l.strip().startswith("["))


@pytest.mark.parametrize("name", ["conv2d_bias_sum_relu", "conv2d_bias_bn_relu", "dense_bias_gelu", "deconv2d_bias_tanh", "layer_norm"])
def test_float_trace_records(name):
    """Trace mode 2 on a float graph: same record names, shapes and dtypes as the oracle's trace, values within tolerance
    (float kernels are not bit-exact with numpy; the integer path's traces are)."""
    func = next(c[1] for c in CASES if c[0] == name)
    mod, js, consts, xs, ins, out = _run(func, trace_mode=2)
    _, recs = interp.run_json(js, consts, xs, want_trace=True)
    got = interp.load_ndarray_list(mod.trace_blob())
    assert [r[0] for r in got] == [r[0] for r in recs]
    for (gn, ga), (wn, wa) in zip(got, recs):
        assert ga.shape == wa.shape and ga.dtype == wa.dtype, gn
        np.testing.assert_allclose(ga, wa, rtol=1e-3, atol=1e-3, err_msg=gn)


def test_float_and_integer_operators_in_one_graph():
    """qnn.dequantize -> float softmax: the float operators compose with the integer chain."""
    from tachikoma_b200 import relay as R
    x = R.var("x", (4, 10), "uint8")
    f = R.Function([x], R.softmax(R.qnn_dequantize(x, np.float32(0.05), 128), axis=1))
    part = T.partition_for_tachikoma(f)
    sym, js, names, consts = T.serialize(part)
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    mod.init(consts)
    xv = np.random.default_rng(0).integers(0, 256, (4, 10)).astype("uint8")
    out = np.empty((4, 10), "float32")
    mod.run(xv, out)
    want, _ = interp.run_json(js, consts, [xv])
    np.testing.assert_allclose(out, want[0], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(out.sum(axis=1), 1.0, rtol=1e-5)


def test_unimplemented_float_cases_fail_at_init_with_the_reference_message():
    from tachikoma_b200 import relay as R
    x = R.var("x", (1, 4, 8))
    f = R.Function([x], R.layer_norm(x, np.ones(4, "float32"), np.zeros(4, "float32"), axis=1))
    sym, js, names, consts = T.serialize(T.partition_for_tachikoma(f))
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    with pytest.raises(RT.TachikomaError, match="Unimplemented LayerNorm case"):   # tachikoma_json_runtime.cc:611
        mod.init(consts)


def test_rewritten_float_graphs_on_gpu():
    """rewrite_resnetv1 / rewrite_dense_bias_gelu_reshape_last (tachikoma_b200/rewrites.py), then partition and run: the GPU
    result of the REWRITTEN graph equals the oracle's evaluation of the ORIGINAL one."""
    import test_rewrites_cpu as TR
    rng = np.random.default_rng(7)
    f = TR._resnetv1_block(rng)
    r = T.rewrite_resnetv1(f)
    assert r is not f
    mod, js, consts, xs, ins, out = _run(r)
    want, _ = interp.run_expr(f, ins)
    np.testing.assert_allclose(out, want, rtol=1e-3, atol=1e-3)
    assert "tachikoma.conv2d_bias_sum_relu" in mod.plan_summary() or "conv2d_bias" in mod.plan_summary()
    g = TR._dense_reshape_graph(True, rng)
    r = T.rewrite_dense_bias_gelu_reshape_last(g)
    mod, js, consts, xs, ins, out = _run(r)
    want, _ = interp.run_expr(g, ins)
    np.testing.assert_allclose(out, want, rtol=1e-3, atol=1e-3)
    assert "tachikoma.dense_bias_gelu" in mod.plan_summary()


def test_multiple_outputs_on_gpu():
    """test_tachikoma.py:640-650 (test_multiple_outputs) plus an integer graph with two heads."""
    from tachikoma_b200 import relay as R
    x, y = R.var("x", (1, 3)), R.var("y", (1, 3))
    z = R.add(x, y)
    f = R.Function([x, y], R.Tuple([z, R.add(z, y)]))
    sym, js, names, consts = T.serialize(T.partition_for_tachikoma(f))
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    mod.init(consts)
    xv, yv = np.float32([[1, 2, 3]]), np.float32([[10, 20, 30]])
    o0, o1 = np.empty((1, 3), "float32"), np.empty((1, 3), "float32")
    mod.run(xv, yv, o0, o1, trace_mode=2)
    np.testing.assert_array_equal(o0, [[11, 22, 33]])
    np.testing.assert_array_equal(o1, [[21, 42, 63]])
    _, recs = interp.run_json(js, consts, [xv, yv], want_trace=True)
    assert mod.trace_blob() == interp.save_ndarray_list(recs)
    d = R.var("d", (2, 8, 10, 10), "uint8")
    p = R.max_pool2d(d, (2, 2), (2, 2))
    f = R.Function([d], R.Tuple([p, R.cast(p, "int32")]))
    sym, js, names, consts = T.serialize(T.partition_for_tachikoma(f))
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    mod.init(consts)
    dv = np.random.default_rng(0).integers(0, 256, (2, 8, 10, 10)).astype("uint8")
    o0, o1 = np.empty((2, 8, 5, 5), "uint8"), np.empty((2, 8, 5, 5), "int32")
    mod.run(dv, o0, o1)
    want, _ = interp.run_json(js, consts, [dv])
    np.testing.assert_array_equal(o0, want[0])
    np.testing.assert_array_equal(o1, want[1])


def test_float_json_fixture_graph_on_gpu():
    """The graph behind tests/golden/json_fixture_float.json (conv + bias + sum + relu, conv + clip, nn.batch_norm with its
    three declared outputs, softmax) runs from exactly that JSON text."""
    import os
    from test_float_cpu import float_fixture_graph
    f = float_fixture_graph()
    sym, js, names, consts = T.serialize(T.partition_for_tachikoma(f))
    assert js == open(os.path.join(os.path.dirname(__file__), "golden", "json_fixture_float.json")).read()
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    mod.init(consts)
    ins = FG.inputs_for(f)
    out = np.empty((1, 4, 4, 4), "float32")
    mod.run(ins["x"], ins["y"], out, trace_mode=2)
    want, _ = interp.run_expr(f, ins)
    np.testing.assert_allclose(out, want, rtol=1e-3, atol=1e-3)
    got = interp.load_ndarray_list(mod.trace_blob())
    _, recs = interp.run_json(js, consts, [ins["x"], ins["y"]], want_trace=True)
    assert [r[0] for r in got] == [r[0] for r in recs]
