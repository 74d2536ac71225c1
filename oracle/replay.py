"""Trace reader / verifier: replays a tachikoma trace record by record on the
CPU oracle -- the "witness check" use of the trace the reference's README.md:6 describes.

TEST INFRASTRUCTURE, like everything under oracle/: the product path never imports this.

Input: the tachikoma JSON graph (the side-car of a trace, row T4: include/.../json_node.h:51-93 schema) and a
trace blob in the NDArray-list container (src/runtime/file_utils.cc:210-229).  Record names follow the debug
dump convention ``{op}____topo-index:{nid}____output-num:0`` (python/tvm/contrib/debugger/debug_result.py:114-127)
extended with ``____input-num:{i}``.  For every kernel node the verifier
  1. collects the node's recorded inputs and recorded output,
  2. re-computes the output from the RECORDED inputs alone with the oracle's restatement of the operator
     (so each record is checked independently of every other one), and compares bit for bit,
  3. checks the chain: an input that is another kernel node's output must equal that node's output record,
     a constant input must equal the module's constant when the constants are supplied.
"""
from __future__ import annotations

import json
import re

import numpy as np

from . import interp

_NAME = re.compile(r"^(?P<op>.*)____topo-index:(?P<nid>\d+)____(?P<kind>input|output)-num:(?P<i>\d+)$")


class TraceMismatch(AssertionError):
    pass


def index_records(records):
    """[(name, array)] -> {nid: {"op": str, "in": {i: array}, "out": {j: array}}}"""
    by_node = {}
    for name, arr in records:
        m = _NAME.match(name)
        if not m:
            raise TraceMismatch("unparsable record name %r" % name)
        node = by_node.setdefault(int(m["nid"]), {"op": m["op"], "in": {}, "out": {}})
        if node["op"] != m["op"]:
            raise TraceMismatch("records of node %s disagree on the operator name" % m["nid"])
        slot = node["in" if m["kind"] == "input" else "out"]
        if int(m["i"]) in slot:
            raise TraceMismatch("duplicate record %r" % name)
        slot[int(m["i"])] = arr
    return by_node


def verify_trace(graph_json, blob, consts=None, stop_at_first=False):
    """Replay ``blob`` against ``graph_json``.  Returns a list of problem strings (empty = the trace is a
    valid witness of the graph's integer semantics).  ``consts``: optional arrays in const order."""
    g = json.loads(graph_json)
    nodes = g["nodes"]
    recs = index_records(interp.load_ndarray_list(blob))
    const_of = {}
    ci = 0
    for nid in g["arg_nodes"]:
        if nodes[nid]["op"] == "const":
            if consts is not None:
                const_of[nid] = np.asarray(consts[ci])
            ci += 1
    problems = []

    def bad(msg):
        problems.append(msg)
        if stop_at_first:
            raise TraceMismatch(msg)

    for nid, n in enumerate(nodes):
        if n["op"] != "kernel":
            continue
        if nid not in recs:
            bad("node %d (%s): no records" % (nid, n["name"]))
            continue
        r = recs[nid]
        if r["op"] != n["name"]:
            bad("node %d: records name %r, graph says %r" % (nid, r["op"], n["name"]))
        if sorted(r["in"]) != list(range(len(n["inputs"]))) or sorted(r["out"]) != [0]:
            bad("node %d (%s): expected %d input records and 1 output record, found inputs %s outputs %s"
                % (nid, n["name"], len(n["inputs"]), sorted(r["in"]), sorted(r["out"])))
            continue
        ins = [r["in"][i] for i in range(len(n["inputs"]))]
        # (3) chain / constants
        for i, e in enumerate(n["inputs"]):
            src = nodes[e[0]]
            shape, dtype = src["attrs"]["shape"][0][e[1]], src["attrs"]["dtype"][0][e[1]]
            if list(ins[i].shape) != list(shape) or str(ins[i].dtype) != dtype:
                bad("node %d input %d: record is %s%s, graph entry is %s%s" % (nid, i, ins[i].dtype, list(ins[i].shape), dtype, list(shape)))
            if src["op"] == "kernel" and e[0] in recs and 0 in recs[e[0]]["out"]:
                if not np.array_equal(ins[i], recs[e[0]]["out"][0]):
                    bad("node %d input %d differs from the output record of its producer node %d" % (nid, i, e[0]))
            elif src["op"] == "const" and e[0] in const_of:
                if not np.array_equal(ins[i], const_of[e[0]].astype(dtype).reshape(shape)):
                    bad("node %d input %d differs from constant node %d" % (nid, i, e[0]))
        # (2) the operator itself, from the recorded inputs only
        try:
            want = interp._eval_json_node(n, [np.ascontiguousarray(a) for a in ins])
        except Exception as exc:  # malformed record shapes etc.
            bad("node %d (%s): oracle could not evaluate the recorded inputs: %s" % (nid, n["name"], exc))
            continue
        got = r["out"][0]
        if want.shape != got.shape or want.dtype != got.dtype:
            bad("node %d (%s): output record is %s%s, oracle gives %s%s" % (nid, n["name"], got.dtype, list(got.shape), want.dtype, list(want.shape)))
        elif not np.array_equal(want, got):
            diff = np.flatnonzero(want.reshape(-1) != got.reshape(-1))
            bad("node %d (%s): %d of %d output elements differ from the oracle (first at flat index %d: trace %d, oracle %d)"
                % (nid, n["name"], diff.size, want.size, diff[0], int(got.reshape(-1)[diff[0]]), int(want.reshape(-1)[diff[0]])))
    extra = sorted(set(recs) - {i for i, n in enumerate(nodes) if n["op"] == "kernel"})
    if extra:
        bad("records for node ids that are not kernel nodes of the graph: %s" % extra)
    return problems


# ---- debug-executor-compatible dump (python/tvm/contrib/debugger/debug_result.py:25, :114-149, :191-202) -------------
def load_debug_dump(directory):
    """Read the pair the debug executor writes: ``_tvmdbg_graph_dump.json`` (nodes carry ``op``, ``name``, ``inputs``
    as producer NAMES, ``attrs`` incl. ``T``, ``shape``) and ``output_tensors.params`` (NDArray-list blob keyed
    ``{name}____topo-index:{i}____output-num:{j}`` over ALL nodes).  Returns (graph dict, {key: array})."""
    import os
    with open(os.path.join(str(directory), "_tvmdbg_graph_dump.json")) as f:
        graph = json.load(f)
    with open(os.path.join(str(directory), "output_tensors.params"), "rb") as f:
        tensors = dict(interp.load_ndarray_list(f.read()))
    return graph, tensors


def verify_debug_dump(graph_json, directory, consts=None):
    """Replay a debug dump against the module's tachikoma JSON graph: the dump holds one tensor per node output, so the
    per-node input records of the trace format are rebuilt from the producers' outputs and handed to verify_trace."""
    graph, tensors = load_debug_dump(directory)
    g = json.loads(graph_json)
    problems = []
    if [n["name"] for n in graph["nodes"]] != [n["name"] for n in g["nodes"]]:
        return ["debug dump and graph disagree on the node list"]
    recs = []
    for nid, n in enumerate(g["nodes"]):
        if n["op"] != "kernel":
            continue
        for i, e in enumerate(n["inputs"]):
            key = "%s____topo-index:%d____output-num:%d" % (g["nodes"][e[0]]["name"], e[0], e[1])
            if key not in tensors:
                problems.append("debug dump lacks %s" % key)
                continue
            recs.append(("%s____topo-index:%d____input-num:%d" % (n["name"], nid, i), tensors[key]))
        key = "%s____topo-index:%d____output-num:0" % (n["name"], nid)
        if key not in tensors:
            problems.append("debug dump lacks %s" % key)
            continue
        recs.append((key, tensors[key]))
    if problems:
        return problems
    return verify_trace(graph_json, interp.save_ndarray_list(recs), consts)
