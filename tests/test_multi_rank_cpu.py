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
    return out[0] if isinstance(out, (list, tuple)) else out, TR.save_trace(recs)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    off, cnt = TR.shard_batch(GLOBAL_BATCH, world, rank)
    out, blob = _rank_trace(off, cnt)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object((off, np.asarray(out), blob), gathered, dst=0)
    if rank == 0:
        gathered.sort(key=lambda t: t[0])
        merged = TR.concat_traces([g[2] for g in gathered], [g[0] for g in gathered])
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
    _, blob = _rank_trace(0, 3)
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
    ref, _ = _rank_trace(0, GLOBAL_BATCH)
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
