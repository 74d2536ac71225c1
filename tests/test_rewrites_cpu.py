"""The reference wrapper's three float-graph rewrites (python/tvm/relay/op/contrib/tachikoma.py:796-1119, SURVEY.md section 8
row B6) on the mini IR: structure of the result as in the reference's own tests (tests/python/contrib/test_tachikoma.py
test_layer_norm :1123, test_rewrite_dense_bias_gelu_reshape_last :1132-1153, test_resnetv1_rewrite :1156-1253) and numerical
equivalence of the rewritten graph on the oracle."""
import numpy as np

from oracle import interp
from tachikoma_b200 import relay as R, tachikoma as T


def _ops(f):
    return [e.op for e in R.post_order(f.body) if isinstance(e, R.Call)]


def _c(v):
    return R.const(np.asarray(v, "float32"))


def _layer_norm_graph(variant, rng):
    x = R.var("x", (1, 49, 64))
    g, b = _c(rng.uniform(0.5, 1.5, 64)), _c(rng.standard_normal(64))
    call = lambda op, args, shape, **attrs: R.Call(op, args, attrs, shape, "float32")
    mu = call("mean", [x], (1, 49, 1), axis=[-1], keepdims=True)
    diff = call("subtract", [x, mu], x.shape)
    if variant == "variance":
        var = call("variance", [x, mu], (1, 49, 1), axis=[-1], keepdims=True)
    else:
        d = call("cast", [diff], x.shape, dtype="float32") if variant == "cast" else diff
        var = call("mean", [call("power", [d, _c(2.0)], x.shape)], (1, 49, 1), axis=[-1], keepdims=True)
    eps = call("add", [var, _c(1e-5)], (1, 49, 1))
    if variant == "rsqrt":
        norm = call("multiply", [diff, call("rsqrt", [eps], (1, 49, 1))], x.shape)
    else:
        norm = call("divide", [diff, call("sqrt", [eps], (1, 49, 1))], x.shape)
    out = call("add", [call("multiply", [norm, g], x.shape), b], x.shape)
    return R.Function([x], R.relu(out)), g, b


def test_rewrite_layer_norm_every_variant():
    rng = np.random.default_rng(0)
    for variant in ("plain", "cast", "variance", "rsqrt"):
        f, g, b = _layer_norm_graph(variant, rng)
        r = T.rewrite_layer_norm(f)
        assert _ops(r) == ["nn.layer_norm", "nn.relu"], variant
        ln = r.body.args[0]
        assert ln.args[1] is g and ln.args[2] is b and ln.attrs["axis"] == -1 and ln.attrs["epsilon"] == 1e-5
        x = rng.standard_normal((1, 49, 64)).astype("float32")
        got, _ = interp.run_expr(r, {"x": x})
        mu = x.astype(np.float64).mean(-1, keepdims=True)
        var = ((x - mu) ** 2).mean(-1, keepdims=True)
        want = np.maximum((x - mu) / np.sqrt(var + 1e-5) * g.data + b.data, 0)
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
        # and the rewritten graph is offloadable as one subgraph
        assert _ops(T.partition_for_tachikoma(r)) == ["nn.layer_norm", "nn.relu"]


def test_rewrite_layer_norm_leaves_other_graphs_alone():
    x = R.var("x", (1, 49, 64))
    f = R.Function([x], R.relu(R.add(R.multiply(x, _c(np.ones(64))), _c(np.zeros(64)))))
    assert T.rewrite_layer_norm(f) is f
    # a different epsilon is not the pattern (is_expr(const(1e-5)) | is_expr(const(1e-6)), tachikoma.py:838)
    g, _, _ = _layer_norm_graph("plain", np.random.default_rng(1))
    for n in R.post_order(g.body):
        if isinstance(n, R.Constant) and n.data.size == 1 and np.isclose(n.data.reshape(-1)[0], 1e-5):
            n.data = np.asarray(1e-3, "float32")
    assert T.rewrite_layer_norm(g) is g


def _dense_reshape_graph(gelu, rng):
    x = R.var("x", (3136, 64))
    w, b = _c(rng.standard_normal((512, 64)) / 8), _c(rng.standard_normal(512))
    d = R.reshape(R.dense(x, w), [1, 3136, 512])
    added = R.Call("add", [b, d], {}, d.shape, "float32")  # bias first, as in the pattern (tachikoma.py:890)
    if not gelu:
        return R.Function([x], added)
    out = R.divide(added, _c(1.41421))
    out = R.add(R.erf(out), _c(1.0))
    out = R.multiply(R.multiply(added, out), _c(0.5))
    return R.Function([x], out)


def test_rewrite_dense_bias_gelu_reshape_last():
    rng = np.random.default_rng(2)
    x = rng.standard_normal((3136, 64)).astype("float32")
    for gelu in (True, False):
        f = _dense_reshape_graph(gelu, rng)
        r = T.rewrite_dense_bias_gelu_reshape_last(f)
        assert r is not f and _ops(r)[-1] == "reshape" and _ops(r)[0] == "nn.dense"
        assert r.body.attrs["newshape"] == [1, 3136, 512]
        want, _ = interp.run_expr(f, {"x": x})
        got, _ = interp.run_expr(r, {"x": x})
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6)
        # the point of the rewrite: dense + bias [+ gelu] now fuse (test_tachikoma.py:1147-1153 expects one subgraph)
        assert _ops(T.partition_for_tachikoma(r)) == ["tachikoma.dense_bias_gelu" if gelu else "tachikoma.dense_bias", "reshape"]
        if gelu:  # before the rewrite the GELU's operators stand alone and are not offloadable
            import pytest
            with pytest.raises(ValueError, match="Unsupported op: divide"):
                T.partition_for_tachikoma(f)
        else:
            assert _ops(T.partition_for_tachikoma(f)) == ["tachikoma.dense", "reshape", "add"]


def _resnetv1_block(rng):
    x = R.var("x", (1, 256, 14, 14))
    w = lambda o, i, k: _c(rng.standard_normal((o, i, k, k)) / np.sqrt(i * k * k))
    b = lambda o: _c(rng.standard_normal((o, 1, 1)) * 0.1)
    cbr = lambda d, o, i, k, s=1, relu=True: (lambda c: R.relu(c) if relu else c)(
        R.add(R.conv2d(d, w(o, i, k), strides=(s, s), padding=(k // 2, k // 2)), b(o)))
    c1 = cbr(x, 64, 256, 1)
    c2 = cbr(c1, 64, 64, 3)
    c3 = R.relu(R.add(cbr(c2, 256, 64, 1, relu=False), x))
    l4 = cbr(c3, 128, 256, 1, s=2)
    l5 = cbr(l4, 128, 128, 3)
    l6 = cbr(l5, 512, 128, 1, relu=False)
    r7 = cbr(c3, 512, 256, 1, s=2, relu=False)
    return R.Function([x], R.relu(R.add(l6, r7)))


def test_rewrite_resnetv1_moves_the_stride_forward_and_keeps_the_result():
    rng = np.random.default_rng(3)
    f = _resnetv1_block(rng)
    r = T.rewrite_resnetv1(f)
    assert r is not f
    convs = [e for e in R.post_order(r.body) if isinstance(e, R.Call) and e.op == "nn.conv2d"]
    assert [c.attrs["strides"] for c in convs] == [[1, 1], [2, 2], [1, 1], [1, 1], [1, 1], [1, 1], [1, 1]]
    pools = [e for e in R.post_order(r.body) if isinstance(e, R.Call) and e.op == "nn.max_pool2d"]
    assert len(pools) == 1 and pools[0].attrs["pool_size"] == [1, 1] and pools[0].attrs["strides"] == [2, 2]
    x = rng.standard_normal((1, 256, 14, 14)).astype("float32")
    want, _ = interp.run_expr(f, {"x": x})
    got, _ = interp.run_expr(r, {"x": x})
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5)
    # fewer multiply-accumulates: the 3x3 convolution and everything after it now run at 7x7
    macs = lambda g: sum(int(np.prod(e.shape)) * int(np.prod(e.args[1].shape[1:])) for e in R.post_order(g.body)
                         if isinstance(e, R.Call) and e.op == "nn.conv2d")
    assert macs(r) < 0.8 * macs(f)
    assert T.rewrite_resnetv1(r) is r  # idempotent: the rewritten block is not the pattern
