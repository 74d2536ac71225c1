"""Host-side multi-rank logic on CPU (gloo, world_size 2): batch sharding by rank, per-rank traces, final
host-side trace concatenation (SURVEY.md section 8 row M1 -- no data-path collective).  The per-rank
"device" here is the CPU oracle: what is under test is the sharding / concatenation code that bench.py and
a multi-GPU user run around the CUDA library."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import interp
from tachikoma_b200 import models, tachikoma as T, trace as TR

GLOBAL_BATCH = 10


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_trace(offset, count):
    part = T.partition_for_tachikoma(models.mlp(batch=count))
    sym, js, names, consts = T.serialize(part)
    x = models.synthetic_input("mlp", GLOBAL_BATCH)[offset:offset + count]
    out, recs = interp.run_json(js, consts, [x], want_trace=True)
    return out[0] if isinstance(out, (list, tuple)) else out, TR.save_trace(recs), js


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    off, cnt = TR.shard_batch(GLOBAL_BATCH, world, rank)
    out, blob, js = _rank_trace(off, cnt)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object((off, np.asarray(out), blob), gathered, dst=0)
    if rank == 0:
        gathered.sort(key=lambda t: t[0])
        merged = TR.concat_traces([g[2] for g in gathered], [g[0] for g in gathered], graph_json=js)
        q.put((np.concatenate([g[1] for g in gathered], axis=0), merged))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_batch_partitions_contiguously():
    for total, world in [(64, 8), (10, 3), (5, 8), (0, 2)]:
        spans = [TR.shard_batch(total, world, r) for r in range(world)]
        assert sum(c for _, c in spans) == total
        pos = 0
        for off, cnt in spans:
            assert off == pos
            pos += cnt
    with pytest.raises(ValueError):
        TR.shard_batch(4, 2, 2)


def test_trace_container_round_trip_matches_oracle_writer():
    _, blob, _ = _rank_trace(0, 3)
    recs = TR.load_trace(blob)
    assert TR.save_trace(recs) == blob
    assert blob == interp.save_ndarray_list(recs)  # the product writer and the oracle writer agree byte for byte


def test_two_ranks_gloo_shard_and_concat():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out, merged = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ref, _, _ = _rank_trace(0, GLOBAL_BATCH)
    np.testing.assert_equal(out, ref)  # sharded run == single-rank run of the whole batch
    recs = TR.load_trace(merged)
    names = [n for n, _ in recs]
    assert any(n.endswith("____rank:1____sample-offset:5") for n in names)
    # rank 1's copies of the weights were dropped, its activations kept
    w0 = [n for n in names if "input-num:1____rank:0" in n]
    w1 = [n for n in names if "input-num:1____rank:1" in n]
    assert w0 and not w1
    outs = [a for n, a in recs if "output-num:0" in n and "topo-index:%s" % names[-1].split("topo-index:")[1].split("_")[0] in n]
    np.testing.assert_equal(np.concatenate(outs, axis=0), ref)


def test_concat_keeps_identical_activations_of_other_ranks():
    """Two ranks fed the SAME samples (duplicate shards, debug runs): only records of const nodes are de-duplicated,
    every activation record of rank 1 survives -- including second operands (residual / qnn.add inputs), which an
    input-position heuristic would take for parameters."""
    from tachikoma_b200 import relay as R
    a = R.var("a", (2, 8), "uint8")
    b = R.var("b", (2, 8), "uint8")
    e = R.clip(R.qnn_add(a, b, np.float32(0.5), 3, np.float32(0.25), 1, np.float32(0.5), 2), 0.0, 255.0)
    part = T.partition_for_tachikoma(R.Function([a, b], e))
    sym, js, names, consts = T.serialize(part)
    xa = np.arange(16, dtype=np.uint8).reshape(2, 8)
    xb = (xa * 3).astype(np.uint8)
    _, recs = interp.run_json(js, consts, [xa, xb], want_trace=True)
    blob = TR.save_trace(recs)
    merged = TR.load_trace(TR.concat_traces([blob, blob], [0, 2], graph_json=js))
    per_rank = {0: [], 1: []}
    for n, _ in merged:
        per_rank[int(n.split("____rank:")[1].split("_")[0])].append(n.split("____rank:")[0])
    const_names = TR.const_record_names(js)
    all_names = [n for n, _ in recs]
    assert per_rank[0] == all_names
    assert per_rank[1] == [n for n in all_names if n not in const_names]
    assert any("input-num:1" in n for n in per_rank[1])  # the second qnn.add operand is an activation: kept
    assert const_names and all(n in all_names for n in const_names)
    # without the graph nothing is dropped
    assert len(TR.load_trace(TR.concat_traces([blob, blob]))) == 2 * len(recs)
