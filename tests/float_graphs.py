"""Float32 test graphs that mirror the builders of the reference's tests/python/contrib/test_tachikoma.py (get_conv1d :265,
get_conv2d :322, get_conv2d_transpose :349, get_conv2d_bias :400, get_conv2d_bias_sum_relu :452, get_conv3d :459,
get_conv3d_transpose :486, get_dense :551, get_bmm :568, get_dense_bias :597, add_activation :220-262) on the mini IR.
Every case is (name, Function, expected composite / operator names)."""
import numpy as np

from tachikoma_b200 import relay as R

ACTS = [None, "relu", "tanh", "sigmoid", "clip", "swish", "gelu", "mish"]


def _rng(seed):
    return np.random.default_rng(seed)


def _w(rng, shape, scale=None):
    fan = int(np.prod(shape[1:])) or 1
    return R.const((rng.standard_normal(shape) / np.sqrt(fan) if scale is None else rng.standard_normal(shape) * scale).astype("float32"))


def add_activation(act, out):
    """test_tachikoma.py:220-262."""
    if act is None:
        return out
    if act == "relu":
        return R.relu(out)
    if act == "tanh":
        return R.tanh(out)
    if act == "sigmoid":
        return R.sigmoid(out)
    if act == "clip":
        return R.clip(out, 0.0, 6.0)
    if act == "swish":
        return R.multiply(out, R.sigmoid(out))
    if act == "gelu":
        out2 = R.divide(out, R.const(np.float32(1.41421)))
        out2 = R.erf(out2)
        out2 = R.add(out2, R.const(np.float32(1.0)))
        out2 = R.multiply(out, out2)
        return R.multiply(out2, R.const(np.float32(0.5)))
    if act == "mish":
        e = R.exp(out)
        a = R.add(e, R.const(np.float32(1.0)))
        return R.multiply(out, R.tanh(R.log(a)))
    raise ValueError(act)


def _bias(rng, conv, sp):
    o = conv.shape[1]
    return R.add(conv, R.const(rng.standard_normal([o] + [1] * sp).astype("float32") * 0.5))


def conv_cases():
    rng = _rng(0)
    cases = []
    # test_conv1d / test_conv1d_pattern
    x1 = R.var("x", (1, 3, 224))
    cases.append(("conv1d", R.Function([x1], R.conv1d(x1, _w(rng, (16, 3, 3)), padding=(1, 1))), ["tachikoma.conv1d"]))
    x1b = R.var("x", (1, 32, 224))
    cases.append(("conv1d_32", R.Function([x1b], R.conv1d(x1b, _w(rng, (16, 32, 3)), padding=(1, 1))), ["tachikoma.conv1d"]))
    for act in ACTS:
        x = R.var("x", (1, 3, 224))
        c = R.conv1d(x, _w(rng, (10, 3, 3)), padding=(1, 1))
        if act is not None:
            cases.append((f"conv1d_{act}", R.Function([x], add_activation(act, c)), [f"tachikoma.conv1d_{act}"]))
        x = R.var("x", (1, 3, 224))
        cb = _bias(rng, R.conv1d(x, _w(rng, (10, 3, 3)), padding=(1, 1)), 1)
        # the reference's bias predicate only accepts nn.conv2d when the pattern ends in the add / relu (tachikoma.py:267-293)
        exp = ["tachikoma.conv1d", "add"] if act is None else (["tachikoma.conv1d", "add", "nn.relu"] if act == "relu" else [f"tachikoma.conv1d_bias_{act}"])
        cases.append((f"conv1d_bias_{act}", R.Function([x], add_activation(act, cb)), exp))
    # test_conv2d: groups / padding / strides / dilation sweep (:735-752)
    for k_shape, groups in (((16, 32, 3, 3), 1), ((32, 1, 3, 3), 32)):
        for padding in ((0, 0), (1, 1)):
            for strides in ((1, 1), (2, 2)):
                for dilation in ((1, 1), (2, 2)):
                    x = R.var("x", (1, 32, 8, 8))
                    c = R.conv2d(x, _w(rng, k_shape), strides=strides, padding=padding, dilation=dilation, groups=groups)
                    cases.append((f"conv2d_g{groups}_p{padding[0]}_s{strides[0]}_d{dilation[0]}", R.Function([x], c), ["tachikoma.conv2d"]))
    # test_conv2d_pattern (:771-794)
    for act in ACTS:
        x = R.var("x", (1, 32, 8, 8))
        c = R.conv2d(x, _w(rng, (16, 32, 3, 3)))
        if act is not None:
            cases.append((f"conv2d_{act}", R.Function([x], add_activation(act, c)), [f"tachikoma.conv2d_{act}"]))
        x = R.var("x", (1, 32, 8, 8))
        cb = _bias(rng, R.conv2d(x, _w(rng, (16, 32, 3, 3))), 2)
        cases.append((f"conv2d_bias_{act}", R.Function([x], add_activation(act, cb)), ["tachikoma.conv2d_bias" + (f"_{act}" if act else "")]))
    # nn.bias_add spelling of the bias
    x = R.var("x", (2, 8, 10, 10))
    cb = R.bias_add(R.conv2d(x, _w(rng, (12, 8, 3, 3)), padding=(1, 1)), R.const(rng.standard_normal(12).astype("float32")))
    cases.append(("conv2d_bias_add_relu", R.Function([x], R.relu(cb)), ["tachikoma.conv2d_bias_relu"]))
    # test_conv2d_bias_sum_relu (:797-836)
    for relu in (True, False):
        x = R.var("x", (1, 32, 8, 8))
        y = R.var("y", (1, 16, 6, 6))
        cb = _bias(rng, R.conv2d(x, _w(rng, (16, 32, 3, 3))), 2)
        out = R.add(cb, y)
        cases.append((f"conv2d_bias_sum{'_relu' if relu else ''}", R.Function([x, y], R.relu(out) if relu else out),
                      ["tachikoma.conv2d_bias_sum" + ("_relu" if relu else "")]))
    # test_conv2d_transpose (:839-868)
    for padding in ((0, 0), (1, 1)):
        for strides in ((1, 1), (2, 2)):
            x = R.var("x", (1, 32, 8, 8))
            c = R.conv2d_transpose(x, _w(rng, (32, 16, 3, 3)), strides=strides, padding=padding)
            cases.append((f"deconv2d_p{padding[0]}_s{strides[0]}", R.Function([x], c), ["tachikoma.deconv2d"]))
    x = R.var("x", (1, 32, 8, 8))
    c = R.conv2d_transpose(x, _w(rng, (32, 4, 3, 3)), strides=(2, 2), padding=(1, 1), groups=4, output_padding=(1, 1))
    cases.append(("deconv2d_groups_opad", R.Function([x], c), ["tachikoma.deconv2d"]))
    for act in ("relu", "tanh", "sigmoid"):
        x = R.var("x", (1, 32, 8, 8))
        c = R.conv2d_transpose(x, _w(rng, (32, 16, 3, 3)))
        cases.append((f"deconv2d_{act}", R.Function([x], add_activation(act, c)), [f"tachikoma.deconv2d_{act}"]))
        x = R.var("x", (1, 32, 8, 8))
        cb = _bias(rng, R.conv2d_transpose(x, _w(rng, (32, 16, 3, 3))), 2)
        exp = ["tachikoma.deconv2d", "add", "nn.relu"] if act == "relu" else [f"tachikoma.deconv2d_bias_{act}"]
        cases.append((f"deconv2d_bias_{act}", R.Function([x], add_activation(act, cb)), exp))
    # test_conv3d / test_conv3d_pattern / test_conv3d_transpose (:871-934)
    x = R.var("x", (1, 32, 8, 8, 8))
    cases.append(("conv3d", R.Function([x], R.conv3d(x, _w(rng, (16, 32, 3, 3, 3)))), ["tachikoma.conv3d"]))
    x = R.var("x", (1, 32, 8, 8, 8))
    cases.append(("conv3d_pad", R.Function([x], R.conv3d(x, _w(rng, (16, 32, 3, 3, 3)), padding=(0, 0, 0, 1, 1, 1))), ["tachikoma.conv3d"]))
    x = R.var("x", (1, 32, 8, 8, 8))
    cases.append(("conv3d_groups", R.Function([x], R.conv3d(x, _w(rng, (16, 8, 3, 3, 3)), groups=4, strides=(2, 2, 2), padding=(1, 1, 1))), ["tachikoma.conv3d"]))
    for act in ("relu", "tanh", "sigmoid"):
        x = R.var("x", (1, 32, 8, 8, 8))
        cases.append((f"conv3d_{act}", R.Function([x], add_activation(act, R.conv3d(x, _w(rng, (16, 32, 3, 3, 3))))), [f"tachikoma.conv3d_{act}"]))
        x = R.var("x", (1, 32, 8, 8, 8))
        cb = _bias(rng, R.conv3d(x, _w(rng, (16, 32, 3, 3, 3))), 3)
        exp = ["tachikoma.conv3d", "add", "nn.relu"] if act == "relu" else [f"tachikoma.conv3d_bias_{act}"]
        cases.append((f"conv3d_bias_{act}", R.Function([x], add_activation(act, cb)), exp))
    x = R.var("x", (1, 32, 8, 8, 8))
    cases.append(("deconv3d", R.Function([x], R.conv3d_transpose(x, _w(rng, (32, 16, 3, 3, 3)))), ["tachikoma.deconv3d"]))
    x = R.var("x", (1, 32, 8, 8, 8))
    cases.append(("deconv3d_s2", R.Function([x], R.conv3d_transpose(x, _w(rng, (32, 16, 3, 3, 3)), strides=(2, 2, 2))), ["tachikoma.deconv3d"]))
    x = R.var("x", (1, 32, 8, 8, 8))
    cases.append(("deconv3d_s2_opad", R.Function([x], R.conv3d_transpose(x, _w(rng, (32, 16, 3, 3, 3)), strides=(2, 2, 2), output_padding=(1, 1, 1))), ["tachikoma.deconv3d"]))
    x = R.var("x", (1, 32, 8, 8, 8))
    cases.append(("deconv3d_tanh", R.Function([x], R.tanh(R.conv3d_transpose(x, _w(rng, (32, 16, 3, 3, 3))))), ["tachikoma.deconv3d_tanh"]))
    return cases


def dense_cases():
    rng = _rng(1)
    cases = []
    # test_dense / test_dense_pattern (:937-974), test_bmm (:581-594)
    x = R.var("x", (1, 16))
    cases.append(("dense", R.Function([x], R.dense(x, _w(rng, (32, 16)))), ["tachikoma.dense"]))
    x = R.var("x", (1, 16))
    cases.append(("dense_k1", R.Function([x], R.dense(x, _w(rng, (1, 16)))), ["tachikoma.dense"]))
    x = R.var("x", (130, 70))
    cases.append(("dense_ragged", R.Function([x], R.dense(x, _w(rng, (75, 70)))), ["tachikoma.dense"]))
    for act in ACTS:
        x = R.var("x", (1, 16))
        d = R.dense(x, _w(rng, (32, 16)))
        if act is not None:
            cases.append((f"dense_{act}", R.Function([x], add_activation(act, d)), [f"tachikoma.dense_{act}"]))
        x = R.var("x", (2, 16))
        db = R.add(R.dense(x, _w(rng, (32, 16))), R.const(rng.standard_normal(32).astype("float32")))
        cases.append((f"dense_bias_{act}", R.Function([x], add_activation(act, db)), ["tachikoma.dense_bias" + (f"_{act}" if act else "")]))
    for ta, tb, xs, ys in ((False, True, (1, 16, 8), (1, 4, 8)), (False, False, (3, 16, 8), (3, 8, 4)),
                           (True, False, (2, 8, 16), (2, 8, 4)), (True, True, (2, 8, 70), (2, 65, 8))):
        x, y = R.var("x", xs), R.var("y", ys)
        cases.append((f"bmm_ta{int(ta)}_tb{int(tb)}", R.Function([x, y], R.batch_matmul(x, y, transpose_a=ta, transpose_b=tb)), ["nn.batch_matmul"]))
    return cases


def other_cases():
    rng = _rng(2)
    cases = []
    # test_elementwise / test_clip / test_leaky_relu (:653-689)
    for nm, op in (("abs", R.abs_), ("exp", R.exp), ("log", R.log), ("sqrt", R.sqrt), ("relu", R.relu), ("tanh", R.tanh),
                   ("sigmoid", R.sigmoid), ("round", R.round_)):
        x = R.var("x", (1, 8, 3, 3))
        cases.append((f"elt_{nm}", R.Function([x], op(x)), [op(x).op]))
    x = R.var("x", (1, 8, 3, 3))
    cases.append(("elt_clip", R.Function([x], R.clip(x, -0.2, 0.4)), ["clip"]))
    x = R.var("x", (1, 8, 3, 3))
    cases.append(("elt_leaky_relu", R.Function([x], R.leaky_relu(x, 0.1)), ["nn.leaky_relu"]))
    # test_softmax (:692-702)
    for shape, axis in (((1, 1000), 1), ((1, 1000), -1), ((1, 3, 4), -2), ((1, 3, 4), 1), ((4, 7, 5, 3), 2)):
        x = R.var("x", shape)
        cases.append((f"softmax_{len(shape)}d_ax{axis}", R.Function([x], R.softmax(x, axis)), ["nn.softmax"]))
    # test_pool2d / test_global_avg_pooling2d / test_pool3d (:977-1088)
    for pool, nm in ((R.max_pool2d, "max"), (R.avg_pool2d, "avg")):
        for ps, st, pd in (((2, 2), (2, 2), (0, 0)), ((3, 3), (2, 2), (1, 1)), ((3, 3), (1, 1), (0, 0, 1, 1)), ((2, 2), (1, 1), (0, 0))):
            x = R.var("x", (1, 3, 32, 32))
            cases.append((f"{nm}_pool2d_k{ps[0]}_s{st[0]}_p{'x'.join(map(str, pd))}", R.Function([x], pool(x, ps, st, pd)), [pool(x, ps, st, pd).op]))
    x = R.var("x", (1, 3, 32, 32))
    cases.append(("avg_pool2d_include_pad", R.Function([x], R.avg_pool2d(x, (3, 3), (2, 2), (1, 1), count_include_pad=True)), ["nn.avg_pool2d"]))
    x = R.var("x", (1, 3, 32, 32))
    cases.append(("global_avg_pool2d", R.Function([x], R.global_avg_pool2d(x)), ["nn.global_avg_pool2d"]))
    for ps, st, pd in (((2, 2, 2), (2, 2, 2), (0, 0, 0)), ((3, 3, 3), (1, 1, 1), (0, 0, 0, 1, 1, 1)), ((2, 2, 2), (2, 1, 1), (1, 1, 1))):
        x = R.var("x", (1, 3, 8, 32, 32))
        cases.append((f"max_pool3d_k{ps[0]}_s{st[0]}_p{len(pd)}", R.Function([x], R.max_pool3d(x, ps, st, padding=pd)), ["nn.max_pool3d"]))
        x = R.var("x", (1, 3, 8, 32, 32))
        cases.append((f"avg_pool3d_k{ps[0]}_s{st[0]}_p{len(pd)}", R.Function([x], R.avg_pool3d(x, ps, st, padding=pd)), ["nn.avg_pool3d"]))
    # batch norm (get_conv2d_bias_bn_relu :422-439), layer norm (get_layer_norm :442-449)
    x = R.var("x", (1, 32, 8, 8))
    cb = _bias(rng, R.conv2d(x, _w(rng, (16, 32, 3, 3))), 2)
    bn = R.batch_norm(cb, rng.uniform(0.5, 1.5, 16).astype("float32"), rng.standard_normal(16).astype("float32"),
                      rng.standard_normal(16).astype("float32"), rng.uniform(0.5, 2.0, 16).astype("float32"))
    cases.append(("conv2d_bias_bn_relu", R.Function([x], R.relu(bn)), ["tachikoma.conv2d_bias", "nn.batch_norm", "nn.relu"]))
    x = R.var("x", (1, 49, 64))
    ln = R.layer_norm(x, rng.uniform(0.5, 1.5, 64).astype("float32"), rng.standard_normal(64).astype("float32"))
    cases.append(("layer_norm", R.Function([x], ln), ["nn.layer_norm"]))
    x = R.var("x", (3, 7, 100))
    ln = R.layer_norm(x, rng.uniform(0.5, 1.5, 100).astype("float32"), rng.standard_normal(100).astype("float32"), epsilon=1e-3)
    cases.append(("layer_norm_100", R.Function([x], ln), ["nn.layer_norm"]))
    # test_multiple_outputs-like binary chains and broadcasting (test_binary)
    x, y = R.var("x", (1, 3)), R.var("y", (1, 3))
    cases.append(("add_add", R.Function([x, y], R.add(R.add(x, y), y)), ["add", "add"]))
    x, y = R.var("x", (2, 8, 5, 5)), R.var("y", (8, 1, 1))
    cases.append(("multiply_bcast", R.Function([x, y], R.multiply(x, y)), ["multiply"]))
    x, y = R.var("x", (2, 8, 5, 5)), R.var("y", (2, 8, 5, 5))
    cases.append(("multiply_same", R.Function([x, y], R.multiply(x, y)), ["multiply"]))
    x = R.var("x", (4, 6))
    cases.append(("add_scalar", R.Function([x], R.add(x, R.const(np.float32(1.5)))), ["add"]))
    return cases


def all_cases():
    return conv_cases() + dense_cases() + other_cases() + pad_fold_cases()


def inputs_for(func, seed=3):
    rng = np.random.default_rng(seed)
    ops = {e.op for e in R.post_order(func.body) if isinstance(e, R.Call)}
    out = {}
    for p in func.params:
        v = rng.standard_normal(p.shape).astype("float32")
        if ops & {"log", "sqrt"}:
            v = np.abs(v) + np.float32(0.1)
        out[p.name] = v
    return out


def pad_fold_cases():
    """test_fuse_pad_avg_pool (test_tachikoma.py:1256-1270) and the convolution / max-pool forms of FoldExplicitPadding."""
    rng = _rng(4)
    cases = []
    x = R.var("x", (1, 24, 17, 17))
    p = R.pad(x, [[0, 0], [0, 0], [1, 1], [1, 1]])
    cases.append(("pad_avg_pool", R.Function([x], R.avg_pool2d(p, (3, 3))), ["nn.avg_pool2d"]))
    x = R.var("x", (1, 8, 9, 9))
    p = R.pad(x, [[0, 0], [0, 0], [1, 2], [0, 1]])
    cases.append(("pad_conv2d_relu", R.Function([x], R.relu(R.conv2d(p, _w(rng, (4, 8, 3, 3)), padding=(1, 1)))), ["tachikoma.conv2d_relu"]))
    x = R.var("x", (1, 8, 9, 9))
    p = R.pad(x, [[0, 0], [0, 0], [1, 1], [1, 1]], pad_value=float("-inf"))
    cases.append(("pad_max_pool", R.Function([x], R.max_pool2d(p, (3, 3), (2, 2))), ["nn.max_pool2d"]))
    # a non-zero pad value is not foldable into a convolution: the pad stays an operator
    x = R.var("x", (1, 8, 9, 9))
    p = R.pad(x, [[0, 0], [0, 0], [1, 1], [1, 1]], pad_value=0.5)
    cases.append(("pad_kept_conv2d", R.Function([x], R.conv2d(p, _w(rng, (4, 8, 3, 3)))), ["nn.pad", "tachikoma.conv2d"]))
    return cases
