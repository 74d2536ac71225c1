"""CPU-only tests: host logic, C-ABI symbol table, JSON schema, binary module blob, trace container."""
import ctypes
import json
import os
import re
import struct
import subprocess
import sys

import numpy as np
import pytest

from oracle import interp
from tachikoma_b200 import models, relay as R, runtime as RT, tachikoma as T

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _small_graph():
    np.random.seed(0)
    x = R.var("in_0", (1, 8, 5, 5), "uint8")
    w = R.const(np.random.randint(-20, 20, (16, 8, 3, 3)).astype("int8"))
    c = R.qnn_conv2d(x, w, 0, 0, 0.2, 0.1, (3, 3), 16, padding=(1, 1))
    c = R.add(c, R.const(np.random.randint(-50, 50, (16, 1, 1)).astype("int32")))
    c = R.qnn_requantize(c, np.float32(0.02), 0, 0.2, 30, out_dtype="int32")
    c = R.cast(R.clip(c, 0.0, 255.0), "uint8")
    s = R.var("in_1", (1, 16, 5, 5), "uint8")
    c = R.clip(R.qnn_add(c, s, 0.2, 30, 0.3, 15, 0.2, 5), 0.0, 255.0)
    f = R.Function([x, s], c)
    ins = {"in_0": np.random.randint(0, 20, (1, 8, 5, 5)).astype("uint8"),
           "in_1": np.random.randint(0, 10, (1, 16, 5, 5)).astype("uint8")}
    return f, ins


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "tkb200.h")).read()
    declared = set(re.findall(r"\b(tkb_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = ctypes.CDLL(RT.lib_path())
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/tkb200.h but not exported"


def test_pattern_table_names():
    names = T.pattern_table()
    assert names[:2] == ["tachikoma.qnn.conv2d", "tachikoma.qnn.dense"]
    assert "tachikoma.conv2d_bias_relu" in names and "tachikoma.dense_bias_gelu" in names
    assert len(names) == 4 + 6 * 2 * 8  # tachikoma.py:418-458


def test_partition_fuses_chain_and_keeps_argpacker_order():
    f, _ = _small_graph()
    part = T.partition_for_tachikoma(f)
    sym, js, names, consts = T.serialize(part)
    g = json.loads(js)
    assert set(g) == {"nodes", "arg_nodes", "heads", "node_row_ptr"}
    kern = [n for n in g["nodes"] if n["op"] == "kernel"]
    assert [k["name"] for k in kern] == ["tachikoma.qnn.conv2d"]
    a = kern[0]["attrs"]
    assert a["bias_idx"] == [["2"]] and a["sum_idx"] == [["3"]]
    assert a["PartitionedFromPattern"] == [["qnn.conv2d_add_qnn.requantize_clip_cast_qnn.add_clip_"]]
    assert a["padding"] == [["1", "1", "1", "1"]] and a["kernel_layout"] == [["OIHW"]]
    assert names == [f"tachikoma_{sym}_const_{i}" for i in range(len(names))]
    # float mode follows comp_op_matcher.h:140-156
    partf = T.partition_for_tachikoma(f, mode="dnnl_float")
    _, jsf, _, _ = T.serialize(partf)
    af = [n for n in json.loads(jsf)["nodes"] if n["op"] == "kernel"][0]["attrs"]
    assert af["o_scl_idx"] == [["4"]] and af["act_scl_idx"] == [["5"]] and af["sum_scl_idx"] == [["6"]]
    assert af["dst_zp_idx"] == [["7"]] and af["activation"] == [["clip", "8", "9", "10"]]


def test_fused_json_equals_unfused_expression_on_oracle():
    f, ins = _small_graph()
    ref, _ = interp.run_expr(f, ins)
    part = T.partition_for_tachikoma(f)
    _, js, _, consts = T.serialize(part)
    outs, _ = interp.run_json(js, consts, [ins["in_0"], ins["in_1"]])
    np.testing.assert_equal(outs[0], ref)
    partf = T.partition_for_tachikoma(f, mode="dnnl_float")
    _, jsf, _, constsf = T.serialize(partf)
    outf, _ = interp.run_json(jsf, constsf, [ins["in_0"], ins["in_1"]])
    assert np.abs(outf[0].astype(int) - ref.astype(int)).max() <= 1


def test_create_symbols_binary_roundtrip_and_errors():
    f, ins = _small_graph()
    sym, js, names, consts = T.serialize(T.partition_for_tachikoma(f))
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    assert mod.get_symbol() == sym and mod.get_const_vars() == names
    assert mod.num_inputs == 2 and mod.num_outputs == 1 and mod.type_key == "tachikoma_json"
    assert mod.get_function("nope") is None and mod.get_function("__init_" + sym) is not None
    blob = mod.save_to_binary()
    # dmlc stream: u64 len + symbol, u64 len + json, u64 count + strings
    (n0,) = struct.unpack_from("<Q", blob, 0)
    assert blob[8:8 + n0].decode() == sym
    mod2 = RT.TachikomaJSONRuntime.load_from_binary(blob)
    assert mod2.get_symbol() == sym and mod2.get_const_vars() == names
    with pytest.raises(RT.TachikomaError, match="has not been initialized"):
        mod.run(ins["in_0"], ins["in_1"], np.zeros((1, 16, 5, 5), "uint8"))
    with pytest.raises(RT.TachikomaError, match="non-existent constant|mismatch"):
        RT.TachikomaJSONRuntime(sym, js, names[:-1])
    with pytest.raises(RT.TachikomaError):
        RT.TachikomaJSONRuntime(sym, "{not json", [])


def test_fixed_point_multiplier_matches_oracle_and_reference_pairs():
    # tests/python/contrib/test_hexagon/test_fixed_point_multiply.py:96-100
    for (m, s), v in {(1288490240, -2): 0.15, (1395864320, 1): 1.3, (1288490188, 0): 0.6}.items():
        got = RT.fixed_point_multiplier_shift(v)
        assert got[1] == s and abs(got[0] - m) <= 64
    rng = np.random.default_rng(0)
    for v in list(rng.uniform(1e-6, 8.0, 200)) + [0.0, 0.5, 1.0, 2.0, 1 / 16, 1 / 3, 3.0]:
        assert RT.fixed_point_multiplier_shift(v) == interp.fixed_point_multiplier_shift(v)


def test_trace_container_roundtrip_and_layout():
    recs = [("a____topo-index:3____output-num:0", np.arange(6, dtype=np.uint8).reshape(2, 3)),
            ("b", np.array(-5, dtype=np.int32))]
    blob = interp.save_ndarray_list(recs)
    assert struct.unpack_from("<Q", blob, 0)[0] == 0xF7E58D4F05049CB7  # file_utils.h:107
    back = interp.load_ndarray_list(blob)
    assert [n for n, _ in back] == [n for n, _ in recs]
    for (_, a), (_, b) in zip(recs, back):
        assert a.dtype == b.dtype and a.shape == b.shape and (a == b).all()


def test_models_partition_fully_and_match_unfused_oracle():
    for name, kw in [("mlp", dict(batch=4)), ("lenet5", dict(batch=2)), ("resnet18", dict(batch=1, image=32, num_classes=10))]:
        f = models.MODELS[name](**kw)
        x = models.synthetic_input(name, kw["batch"])
        if "image" in kw:
            x = np.ascontiguousarray(x[:, :, :kw["image"], :kw["image"]])
        ref, _ = interp.run_expr(f, {"data": x})
        part = T.partition_for_tachikoma(f)
        _, js, _, consts = T.serialize(part)
        ops = [n["name"] for n in json.loads(js)["nodes"] if n["op"] == "kernel"]
        assert not any(o.startswith("qnn.") for o in ops), ops  # every QNN chain landed in a composite
        outs, _ = interp.run_json(js, consts, [x])
        np.testing.assert_equal(outs[0], ref)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "tachikoma_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                src = open(os.path.join(dirpath, fn)).read()
                assert "oracle" not in src.replace("test oracle", ""), f"{fn} mentions the oracle"


def test_torch_is_buffer_plumbing_only():
    """PyTorch appears in the package only where device buffers / streams are handled (runtime.py); model building,
    partitioning, serialization and the trace reader are numpy only."""
    pkg = os.path.join(ROOT, "tachikoma_b200")
    for fn in os.listdir(pkg):
        # frontend.py reads TorchScript models (graph walk, packed-parameter unpacking): torch is its input format
        if fn.endswith(".py") and fn not in ("runtime.py", "frontend.py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(import torch|from torch)", src, re.M), f"{fn} imports torch"


# ---- trace reader / verifier: record-by-record replay on the oracle ---------------
def _oracle_trace(func, ins):
    part = T.partition_for_tachikoma(func)
    sym, js, names, consts = T.serialize(part)
    _, recs = interp.run_json(js, consts, [ins[p.name] for p in part.params], want_trace=True)
    return js, consts, recs


def test_trace_replay_accepts_a_valid_trace_and_pinpoints_tampering():
    from oracle import replay
    func, ins = _small_graph()
    js, consts, recs = _oracle_trace(func, ins)
    blob = interp.save_ndarray_list(recs)
    assert replay.verify_trace(js, blob, consts) == []
    # flip one output element of one record: exactly that node (and its consumers' chain check) is reported
    out_idx = [i for i, (n, _) in enumerate(recs) if n.endswith("____output-num:0")]
    i = out_idx[0]
    name, arr = recs[i]
    bad = arr.copy()
    bad.reshape(-1)[3] ^= 1
    tampered = list(recs)
    tampered[i] = (name, bad)
    problems = replay.verify_trace(js, interp.save_ndarray_list(tampered), consts)
    nid = int(re.search(r"topo-index:(\d+)", name).group(1))
    assert any(p.startswith("node %d " % nid) and "differ from the oracle" in p for p in problems), problems
    # a tampered weight is caught against the module's constants
    w_idx = next(i for i, (n, a) in enumerate(recs) if "input-num:1" in n)
    name, arr = recs[w_idx]
    bad = arr.copy()
    bad.reshape(-1)[0] += 1
    tampered = list(recs)
    tampered[w_idx] = (name, bad)
    problems = replay.verify_trace(js, interp.save_ndarray_list(tampered), consts)
    assert any("differs from constant node" in p for p in problems), problems
    # a dropped record is reported, not skipped
    problems = replay.verify_trace(js, interp.save_ndarray_list(recs[:-1]), consts)
    assert problems and "expected" in problems[-1] or "no records" in problems[-1]


@pytest.mark.parametrize("name,kw", [("mlp", dict(batch=4)), ("lenet5", dict(batch=2)),
                                     ("resnet18", dict(batch=1, image=32, num_classes=10))])
def test_trace_replay_models(name, kw):
    from oracle import replay
    func = models.MODELS[name](**kw)
    x = models.synthetic_input(name, kw["batch"])
    if "image" in kw:
        x = np.ascontiguousarray(x[:, :, :kw["image"], :kw["image"]])
    js, consts, recs = _oracle_trace(func, {"data": x})
    assert replay.verify_trace(js, interp.save_ndarray_list(recs), consts) == []


def test_serialize_matches_json_fixtures():
    """The JSON contract of both epilogue modes is pinned byte for byte: key order of the graph
    (json_node.h / graph_executor.cc Load), attribute nesting, constant names, ArgPacker tag order."""
    f, _ = _small_graph()
    gold = os.path.join(ROOT, "tests", "golden")
    for mode, name in (("int", "json_fixture_int.json"), ("dnnl_float", "json_fixture_dnnl_float.json")):
        _, js, names, _ = T.serialize(T.partition_for_tachikoma(f, mode=mode))
        want = open(os.path.join(gold, name)).read()
        assert js == want, f"serialize() drifted from tests/golden/{name}"
        g = json.loads(js)
        assert list(g) == ["nodes", "arg_nodes", "heads", "node_row_ptr"]
        kern = [n for n in g["nodes"] if n["op"] == "kernel"][0]
        a = kern["attrs"]
        assert all(isinstance(v, list) and isinstance(v[0], list) and all(isinstance(s, str) for s in v[0])
                   for k, v in a.items() if k not in ("shape", "dtype", "num_inputs", "num_outputs")), "attributes are [[str, ...]]"
        assert a["num_outputs"] == "1" and a["num_inputs"] == str(len(kern["inputs"]))
        consts = [n["name"] for n in g["nodes"] if n["op"] == "const"]
        assert consts == names
        if mode == "dnnl_float":
            order = ["bias_idx", "sum_idx", "o_scl_idx", "act_scl_idx", "sum_scl_idx", "dst_zp_idx"]
            assert [int(a[t][0][0]) for t in order] == list(range(2, 8))
            assert a["activation"] == [["clip", "8", "9", "10"]]
        else:
            order = ["bias_idx", "sum_idx", "src_zp_idx", "wgh_zp_idx", "src_scl_idx", "wgh_scl_idx", "rq_in_scl_idx",
                     "rq_in_zp_idx", "rq_out_scl_idx", "rq_out_zp_idx", "sum_lhs_scl_idx", "sum_lhs_zp_idx",
                     "sum_rhs_scl_idx", "sum_rhs_zp_idx", "sum_out_scl_idx", "sum_out_zp_idx"]
            assert [int(a[t][0][0]) for t in order] == list(range(2, 18))
            assert a["clip_a_max"] == [["%.17g" % 255.0]] and a["cast_dtype"] == [["uint8"]]


def test_wrapper_surface_of_the_reference_module():
    """python/tvm/relay/op/contrib/tachikoma.py names: present, identity on integer graphs
    (a DFPattern rewrite that matches nothing returns its input)."""
    f, _ = _small_graph()
    for fn in (T.rewrite_layer_norm, T.rewrite_dense_bias_gelu_reshape_last, T.rewrite_resnetv1, T.prune_tachikoma_subgraphs):
        assert fn(f) is f
    assert T.is_valid_subgraph(f.body)
    x = R.var("x", (1, 4), "uint8")
    assert not T.is_valid_subgraph(R.clip(x, 0.0, 9.0))
    fl = R.Function([x], R.Call("nn.relu", [x], {}, x.shape, "float32"))
    assert T.rewrite_resnetv1(fl) is fl  # float graph without the pattern: unchanged (tests/test_rewrites_cpu.py has the rest)
    assert T.get_optimal_layout_for_conv("NCHW", "OIHW") == "NHWC,OHWI,NHWC"


@pytest.mark.parametrize("dtype", ["uint8", "int8"])
def test_table_lookup_legalization_equals_elementwise_oracle(dtype):
    """legalize_qnn_unary (product, numpy) builds take(table, reinterpret(x)); the oracle evaluates the un-legalized
    operator element by element: same bytes for every bit pattern, also where f is undefined (sqrt / log of negatives)."""
    x = np.arange(256, dtype=np.uint8).view(dtype).reshape(4, 64)
    for op in R.QNN_UNARY_OPS:
        for (si, zi, so, zo) in [(0.125, 0, 0.125, 0), (0.05, 3, 0.02, -4 if dtype == "int8" else 7), (0.3, 100 if dtype == "uint8" else -100, 0.01, 0)]:
            v = R.var("x", x.shape, dtype)
            f = R.Function([v], R.qnn_unary(op, v, si, zi, so, zo))
            want, _ = interp.run_expr(f, {"x": x})
            leg = T.legalize_qnn_unary(f)
            ops = [e.op for e in R.post_order(leg.body) if isinstance(e, R.Call)]
            assert ops == ["reinterpret", "take"], ops
            got, _ = interp.run_expr(leg, {"x": x})
            np.testing.assert_array_equal(got, want, err_msg=f"{op} {dtype} {(si, zi, so, zo)}")
            part = T.partition_for_tachikoma(f)
            _, js, _, consts = T.serialize(part)
            outs, _ = interp.run_json(js, consts, [x])
            np.testing.assert_array_equal(outs[0], want)


def test_leaky_relu_oracle_matches_float_definition_within_one():
    x = np.arange(256, dtype=np.uint8).reshape(1, 256)
    v = R.var("x", x.shape, "uint8")
    for alpha, si, zi, so, zo in [(0.9, 0.125, 60, 0.6, 17), (0.1, 0.05, 128, 0.05, 128), (0.25, 0.2, 10, 0.1, 3)]:
        f = R.Function([v], R.qnn_leaky_relu(v, alpha, si, zi, so, zo))
        got, _ = interp.run_expr(f, {"x": x})
        d = (x.astype(np.float64) - zi) * si
        want = np.clip(np.round(np.where(d < 0, alpha * d, d) / so + zo), 0, 255)
        assert np.abs(got.astype(int) - want).max() <= 1
        _, js, _, consts = T.serialize(T.partition_for_tachikoma(f))
        outs, _ = interp.run_json(js, consts, [x])
        np.testing.assert_array_equal(outs[0], got)
