"""The reference wrapper's three float-graph rewrites (python/tvm/relay/op/contrib/tachikoma.py) on the mini IR:

  rewrite_layer_norm                     :796-857   mean / subtract / power / mean (or variance) / add eps / sqrt (or rsqrt) /
                                                    divide / multiply gamma / add beta      -> nn.layer_norm
  rewrite_dense_bias_gelu_reshape_last   :860-951   nn.dense -> reshape -> add bias [-> erf GELU]   -> the reshape moved last
  rewrite_resnetv1                       :954-1119  ResNet-v1 downsample block: the stride of the two 1x1 stride-2 convolutions
                                                    moved forward into the preceding 3x3 convolution, the shortcut through a
                                                    1x1 stride-2 max pool

Each is a DFPatternCallback in the reference; here a small structural matcher over ``relay.Call`` nodes plays the part of
the dataflow-pattern language (wildcards bind by name, the same name must bind the same node) and the rewrite runs in post
order on already-rewritten arguments, as ``tvm.relay.dataflow_pattern.rewrite`` does.  On a graph without the pattern the
function is returned unchanged (the very same object)."""
from __future__ import annotations

import numpy as np

from . import relay as R


class W:
    """wildcard(); ``const`` restricts it to constants, ``values`` to scalar constants with one of these values."""

    def __init__(self, name, const=False, values=None):
        self.name, self.const, self.values = name, const or values is not None, values


class Op:
    """is_op(name)(*args) [.has_attr(attrs)]"""

    def __init__(self, op, *args, attrs=None):
        self.op, self.args, self.attrs = op, args, attrs or {}


class Alt:
    """p1 | p2 | ..."""

    def __init__(self, *alts):
        self.alts = alts


def _match(p, e, env):
    if isinstance(p, Alt):
        for a in p.alts:
            trial = dict(env)
            if _match(a, e, trial):
                env.clear()
                env.update(trial)
                return True
        return False
    if isinstance(p, W):
        if p.const and not isinstance(e, R.Constant):
            return False
        if p.values is not None:
            if e.data.size != 1 or not any(np.isclose(float(e.data.reshape(-1)[0]), v, rtol=1e-6, atol=0) for v in p.values):
                return False
        if p.name in env:
            return env[p.name] is e
        env[p.name] = e
        return True
    if not (isinstance(e, R.Call) and e.op == p.op and len(e.args) == len(p.args)):
        return False
    for k, v in p.attrs.items():
        if list(e.attrs.get(k, [])) != list(v):
            return False
    trial = dict(env)
    if all(_match(pa, ea, trial) for pa, ea in zip(p.args, e.args)):
        env.update(trial)
        return True
    if p.op in ("add", "multiply") and len(p.args) == 2:  # the dataflow matcher tries commutative operators both ways
        trial = dict(env)
        if _match(p.args[0], e.args[1], trial) and _match(p.args[1], e.args[0], trial):
            env.update(trial)
            return True
    return False


def _rewrite(func, patterns, build):
    """Post-order rewrite: at every call (with rewritten arguments) try the patterns in order; ``build(i, env, node)``
    returns the replacement or None."""
    memo = {}
    changed = [False]

    def visit(e):
        if id(e) in memo:
            return memo[id(e)]
        r = e
        if isinstance(e, R.Call):
            args = [visit(a) for a in e.args]
            if any(a is not b for a, b in zip(args, e.args)):
                r = R.Call(e.op, args, e.attrs, e.shape, e.dtype)
            for i, p in enumerate(patterns):
                env = {}
                if _match(p, r, env):
                    new = build(i, env, r)
                    if new is not None:
                        assert tuple(new.shape) == tuple(r.shape) and new.dtype == r.dtype, (new, r)
                        r = new
                        changed[0] = True
                        break
        memo[id(e)] = r
        return r

    body = visit(func.body)
    return R.Function(func.params, body, func.attrs) if changed[0] else func


# ---- layer norm --------------------------------------------------------------------------------------------------------------
def _layer_norm_pattern():
    data, gamma, beta = W("data"), W("gamma"), W("beta")
    mu = Op("mean", data)
    diff = Op("subtract", data, mu)
    cdiff = Alt(diff, Op("cast", diff))
    p1 = Op("power", cdiff, W("two", values=(2.0,)))
    mp1 = Alt(Op("mean", p1), Op("variance", data, mu))
    added_eps = Op("add", mp1, W("eps", values=(1e-5, 1e-6)))
    div_out = Op("divide", diff, Op("sqrt", added_eps))
    div_out2 = Op("multiply", diff, Op("rsqrt", added_eps))
    return Op("add", Op("multiply", Alt(div_out, div_out2), gamma), beta)


def rewrite_layer_norm(func):
    """tachikoma.py:796-857.  (As in the reference the epsilon of the new operator is nn.layer_norm's default 1e-5 whichever of
    the two matched constants the graph held, and the reduction axis is not inspected: callback :846-850.)"""
    def build(_, env, node):
        return R.layer_norm(env["data"], env["gamma"], env["beta"])
    return _rewrite(func, [_layer_norm_pattern()], build)


# ---- dense -> reshape -> bias [-> gelu] ----------------------------------------------------------------------------------------
def _dense_reshape_patterns():
    den = Op("nn.dense", W("data"), W("weight"))
    re_den = Op("reshape", den)
    added = Op("add", W("bias"), re_den)
    divisor = Op("divide", added, W("const1"))
    mul1 = Op("multiply", added, Op("add", Op("erf", divisor), W("const2")))
    return [Op("multiply", mul1, W("const3")), added]


def _find(e, op):
    """The last such call in post order, i.e. the one nearest the matched root (get_attr, tachikoma.py:905-915)."""
    found = None
    for n in R.post_order(e):
        if isinstance(n, R.Call) and n.op == op:
            found = n
    return found


def rewrite_dense_bias_gelu_reshape_last(func):
    """tachikoma.py:860-951: both forms (with and without the GELU), the reshape becomes the last operator so that
    dense + bias [+ gelu] fuse into one composite."""
    def build(i, env, node, has_gelu):
        newshape = _find(node, "reshape").attrs["newshape"]
        den = R.dense(env["data"], env["weight"])
        bias = env["bias"]
        if int(np.prod(bias.shape)) != den.shape[-1] and np.broadcast_shapes(bias.shape, den.shape) != tuple(den.shape):
            return None
        added = R.Call("add", [bias, den], {}, den.shape, den.dtype)
        if not has_gelu:
            return R.reshape(added, newshape)
        out = R.divide(added, env["const1"])
        out = R.add(R.erf(out), env["const2"])
        out = R.multiply(R.multiply(added, out), env["const3"])
        return R.reshape(out, newshape)
    # rewrite([with GELU, without], main): one callback after the other over the whole function (tachikoma.py:946-949)
    with_gelu, without = _dense_reshape_patterns()
    func = _rewrite(func, [with_gelu], lambda i, env, node: build(i, env, node, True))
    return _rewrite(func, [without], lambda i, env, node: build(i, env, node, False))


# ---- ResNet v1 downsample block --------------------------------------------------------------------------------------------------
def _resnetv1_pattern():
    data = W("data")

    def cbr(x, i, relu=True, **attrs):
        c = Op("add", Op("nn.conv2d", x, W(f"w{i}"), attrs=attrs), W(f"b{i}"))
        return Op("nn.relu", c) if relu else c

    conv1 = cbr(data, 1, kernel_size=[1, 1])
    conv2 = cbr(conv1, 2, kernel_size=[3, 3])
    conv3 = Op("nn.relu", Op("add", cbr(conv2, 3, relu=False, kernel_size=[1, 1]), data))
    left4 = cbr(conv3, 4, strides=[2, 2])
    left5 = cbr(left4, 5, kernel_size=[3, 3])
    left6 = cbr(left5, 6, relu=False, kernel_size=[1, 1])
    right7 = cbr(conv3, 7, relu=False, strides=[2, 2])
    return Op("nn.relu", Op("add", left6, right7))


def _conv_like(conv, data, weight, **override):
    a = dict(conv.attrs)
    a.update(override)
    return R.conv2d(data, weight, strides=a["strides"], padding=a["padding"], dilation=a["dilation"], groups=a["groups"],
                    channels=a["channels"], kernel_size=a["kernel_size"], data_layout=a["data_layout"],
                    kernel_layout=a["kernel_layout"], out_layout=a["out_layout"], out_dtype=a.get("out_dtype", ""))


def rewrite_resnetv1(func):
    """tachikoma.py:954-1119."""
    def build(_, env, node):
        convs = [n for n in R.post_order(node) if isinstance(n, R.Call) and n.op == "nn.conv2d"][-7:]  # attr_lst[-7:]
        by_w = {id(c.args[1]): c for c in convs}
        c = [by_w.get(id(env[f"w{i}"])) for i in range(1, 8)]
        if any(x is None for x in c) or c[0].attrs["data_layout"] != "NCHW":
            return None
        data = env["data"]
        try:
            conv1 = R.relu(R.add(_conv_like(c[0], data, env["w1"]), env["b1"]))
            conv2 = R.relu(R.add(_conv_like(c[1], conv1, env["w2"], strides=[2, 2]), env["b2"]))
            conv3 = R.add(_conv_like(c[2], conv2, env["w3"]), env["b3"])
            pool = R.max_pool2d(data, pool_size=(1, 1), strides=(2, 2), layout=c[2].attrs["data_layout"])
            conv3 = R.relu(R.add(conv3, pool))
            left4 = R.relu(R.add(_conv_like(c[3], conv3, env["w4"], strides=[1, 1]), env["b4"]))
            left5 = R.relu(R.add(_conv_like(c[4], left4, env["w5"]), env["b5"]))
            left6 = R.add(_conv_like(c[5], left5, env["w6"]), env["b6"])
            right7 = R.add(_conv_like(c[6], conv3, env["w7"], strides=[1, 1]), env["b7"])
            out = R.relu(R.add(left6, right7))
        except (AssertionError, ValueError):
            return None
        return out if tuple(out.shape) == tuple(node.shape) else None
    return _rewrite(func, [_resnetv1_pattern()], build)
