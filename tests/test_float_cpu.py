"""Float tachikoma patterns on the host side: the partitioner fuses the reference's float
patterns into composites with the reference's names (python/tvm/relay/op/contrib/tachikoma.py:131-353, :418-458), the JSON
carries them, and the oracle evaluates a composite exactly as its primitive chain.  The graphs mirror the reference's
tests/python/contrib/test_tachikoma.py float tests (tests/float_graphs.py)."""
import json

import numpy as np
import pytest

import float_graphs as FG
from oracle import float_ops as FO, interp
from tachikoma_b200 import relay as R, tachikoma as T

CASES = FG.all_cases()


@pytest.mark.parametrize("name,func,expected", CASES, ids=[c[0] for c in CASES])
def test_float_pattern_partition_and_oracle(name, func, expected):
    part = T.partition_for_tachikoma(func)
    ops = [e.op for e in R.post_order(part.body) if isinstance(e, R.Call)]
    assert ops == expected
    sym, js, names, consts = T.serialize(part)
    ins = FG.inputs_for(func)
    want, _ = interp.run_expr(func, ins)                       # primitive chain
    got, _ = interp.run_json(js, consts, [ins[p.name] for p in func.params])  # composites, as the runtime reads them
    np.testing.assert_allclose(got[0], want, rtol=1e-3, atol=1e-3)  # tests/python/contrib/test_tachikoma.py:155


def test_every_float_composite_name_is_in_the_pattern_table():
    table = set(T.pattern_table())
    seen = set()
    for _, func, _ in CASES:
        part = T.partition_for_tachikoma(func)
        seen |= {e.op for e in R.post_order(part.body) if isinstance(e, R.Call) and e.op.startswith("tachikoma.")}
    assert seen <= table
    assert {"tachikoma.conv2d_bias_sum_relu", "tachikoma.conv2d_bias_gelu", "tachikoma.dense_bias_mish", "tachikoma.deconv3d_tanh",
            "tachikoma.conv1d_bias_swish", "tachikoma.conv3d_bias_sigmoid"} <= seen


def test_batch_norm_node_declares_three_outputs():
    _, func, _ = next(c for c in CASES if c[0] == "conv2d_bias_bn_relu")
    sym, js, names, consts = T.serialize(T.partition_for_tachikoma(func))
    g = json.loads(js)
    bn = next(i for i, n in enumerate(g["nodes"]) if n["name"] == "nn.batch_norm")
    assert g["nodes"][bn]["attrs"]["num_outputs"] == "3"
    assert g["nodes"][bn]["attrs"]["shape"] == [[[1, 16, 6, 6], [16], [16]]]
    assert g["node_row_ptr"][bn + 1] - g["node_row_ptr"][bn] == 3 and g["node_row_ptr"][-1] == len(g["nodes"]) + 2


def test_conv_bias_predicate_quirk_of_the_reference():
    """make_bias_add_pattren_predicate / add_checker (tachikoma.py:267-293) accept a bias add only behind nn.conv2d when the
    pattern ends in the add or in relu(add): conv1d + bias [+ relu] stays conv1d, add [, relu]; with tanh it fuses."""
    x = R.var("x", (1, 3, 32))
    w = R.const(np.ones((4, 3, 3), "float32"))
    b = R.const(np.ones((4, 1), "float32"))
    ops = lambda f: [e.op for e in R.post_order(T.partition_for_tachikoma(f).body) if isinstance(e, R.Call)]
    assert ops(R.Function([x], R.relu(R.add(R.conv1d(x, w), b)))) == ["tachikoma.conv1d", "add", "nn.relu"]
    assert ops(R.Function([x], R.tanh(R.add(R.conv1d(x, w), b)))) == ["tachikoma.conv1d_bias_tanh"]


def test_float_oracle_transposed_convolution_is_the_adjoint_of_convolution():
    """<conv(x, w), y> == <x, conv_transpose(y, w)> for every stride / padding / dilation: pins the scatter form of the
    transposed convolution against the (independently written) gather form of the forward one."""
    rng = np.random.default_rng(0)
    for st, pd, di in ((1, 0, 1), (2, 1, 1), (2, 0, 2), (3, 2, 1)):
        x = rng.standard_normal((2, 4, 9, 10))
        w = rng.standard_normal((6, 4, 3, 3))               # OIHW for the forward op == IOHW (I = 6) for the transposed one
        y = FO.conv_nd(x, w, (st, st), (pd, pd), (di, di))
        g = rng.standard_normal(y.shape)
        full = FO.conv_transpose_nd(g, w, (st, st), (pd, pd), (di, di), (0, 0))
        opad = [x.shape[2 + i] - full.shape[2 + i] for i in range(2)]
        full = FO.conv_transpose_nd(g, w, (st, st), (pd, pd), (di, di), opad)
        assert full.shape == x.shape
        np.testing.assert_allclose(np.sum(y * g), np.sum(x * full), rtol=1e-10)


def test_float_oracle_known_answers():
    # convolution with a delta kernel is a shift; pooling and softmax by hand
    x = np.arange(2 * 1 * 4 * 4, dtype=np.float32).reshape(2, 1, 4, 4)
    w = np.zeros((1, 1, 3, 3), np.float32)
    w[0, 0, 1, 1] = 1
    np.testing.assert_array_equal(FO.eval_float_op("nn.conv2d", dict(padding=[1, 1, 1, 1]), [x, w]), x)
    np.testing.assert_array_equal(FO.float_pool("nn.max_pool2d", dict(pool_size=[2, 2], strides=[2, 2], padding=[0, 0, 0, 0]), x)[0, 0],
                                  [[5, 7], [13, 15]])
    np.testing.assert_allclose(FO.float_pool("nn.avg_pool2d", dict(pool_size=[2, 2], strides=[2, 2], padding=[1, 1, 1, 1]), x)[0, 0, 0],
                               [0, 1.5, 3])                   # count_include_pad False: corner 0/1, edge (1+2)/2, corner 3/1
    np.testing.assert_allclose(FO.float_pool("nn.avg_pool2d", dict(pool_size=[2, 2], strides=[2, 2], padding=[1, 1, 1, 1], count_include_pad=True), x)[0, 0, 0],
                               [0, 0.75, 0.75])
    s = FO.eval_float_op("nn.softmax", dict(axis=-1), [np.log(np.array([[1, 2, 5]], np.float32))])
    np.testing.assert_allclose(s, [[0.125, 0.25, 0.625]], rtol=1e-6)
    np.testing.assert_allclose(FO.activation(np.array([0.0, 1.0]), "gelu"), [0.0, 0.8413447460685429], rtol=1e-12)
    np.testing.assert_allclose(FO.activation(np.array([1.0]), "mish"), [0.8650983882673103], rtol=1e-12)
    np.testing.assert_array_equal(FO.eval_float_op("round", {}, [np.array([0.5, 1.5, 2.5, -0.5], np.float32)]), [0, 2, 2, -0])


def test_multiple_outputs_become_graph_heads():
    """test_tachikoma.py:640-650: out = Tuple((x + y, (x + y) + y)) -> two heads, no node for the tuple."""
    x, y = R.var("x", (1, 3)), R.var("y", (1, 3))
    z = R.add(x, y)
    f = R.Function([x, y], R.Tuple([z, R.add(z, y)]))
    part = T.partition_for_tachikoma(f)
    sym, js, names, consts = T.serialize(part)
    g = json.loads(js)
    assert len(g["heads"]) == 2 and [n["name"] for n in g["nodes"] if n["op"] == "kernel"] == ["add", "add"]
    xv, yv = np.float32([[1, 2, 3]]), np.float32([[10, 20, 30]])
    want, _ = interp.run_expr(f, {"x": xv, "y": yv})
    got, _ = interp.run_json(js, consts, [xv, yv])
    assert len(got) == 2
    np.testing.assert_array_equal(got[0], want[0])
    np.testing.assert_array_equal(got[1], [[21, 42, 63]])
    # an integer graph with two outputs: the convolution composite and its pooled copy
    d = R.var("d", (1, 4, 6, 6), "uint8")
    import pytest
    with pytest.raises(ValueError, match="only as the function body"):
        T.serialize(R.Function([x, y], R.add(R.Tuple([z, z]), y)))


def float_fixture_graph():
    x = R.var("x", (1, 4, 6, 6))
    y = R.var("y", (1, 8, 4, 4))
    w = R.const(np.arange(8 * 4 * 3 * 3, dtype="float32").reshape(8, 4, 3, 3) / 100)
    b = R.const(np.arange(8, dtype="float32").reshape(8, 1, 1))
    c = R.relu(R.add(R.add(R.conv2d(x, w), b), y))
    c = R.clip(R.conv2d(c, R.const(np.ones((4, 8, 1, 1), "float32"))), 0.0, 6.0)
    bn = R.batch_norm(c, np.ones(4, "float32"), np.zeros(4, "float32"), np.zeros(4, "float32"), np.ones(4, "float32"))
    return R.Function([x, y], R.softmax(bn, axis=1))


def test_float_serialization_matches_the_json_fixture():
    """tests/golden/json_fixture_float.json pins the JSON of a float subgraph: composite names, inputs in the composite call's
    argument order (data, weight, bias, sum: codegen.cc:516-524), the root call's attributes plus the clip's a_min / a_max on a
    `_clip` composite (codegen.cc:531-536), the three declared outputs of nn.batch_norm and node_row_ptr."""
    import os
    f = float_fixture_graph()
    _, js, names, consts = T.serialize(T.partition_for_tachikoma(f))
    want = open(os.path.join(os.path.dirname(__file__), "golden", "json_fixture_float.json")).read()
    assert js == want, "serialize() drifted from tests/golden/json_fixture_float.json"
    g = json.loads(js)
    kern = {n["name"]: n for n in g["nodes"] if n["op"] == "kernel"}
    assert list(kern) == ["tachikoma.conv2d_bias_sum_relu", "tachikoma.conv2d_clip", "nn.batch_norm", "nn.softmax"]
    assert kern["tachikoma.conv2d_bias_sum_relu"]["attrs"]["num_inputs"] == "4"
    assert kern["tachikoma.conv2d_clip"]["attrs"]["a_min"] == [["0"]] and kern["tachikoma.conv2d_clip"]["attrs"]["a_max"] == [["6"]]
    assert kern["tachikoma.conv2d_clip"]["attrs"]["kernel_size"] == [["1", "1"]]
    assert kern["nn.batch_norm"]["attrs"]["num_outputs"] == "3"
