"""Host-side trace utilities: the NDArray-list container (src/runtime/file_utils.cc:184-229 LoadParams /
SaveParams; per-array layout include/tvm/runtime/ndarray.h:449-494) and the multi-GPU pieces of the
path: contiguous batch sharding by rank and the final host-side
concatenation of the per-rank trace files.  No collective is involved anywhere on the data path.
"""
from __future__ import annotations

import struct

import numpy as np

LIST_MAGIC = 0xF7E58D4F05049CB7
ARRAY_MAGIC = 0xDD5E40F096B4A13F
_CODES = {"i": 0, "u": 1, "f": 2}
_KINDS = {0: "int", 1: "uint", 2: "float"}


def shard_batch(global_batch, world_size, rank):
    """Contiguous split of ``global_batch`` samples over ``world_size`` ranks -> (offset, count).
    The first ``global_batch % world_size`` ranks take one extra sample."""
    if not 0 <= rank < world_size:
        raise ValueError("rank %d outside world of %d" % (rank, world_size))
    base, extra = divmod(int(global_batch), int(world_size))
    count = base + (1 if rank < extra else 0)
    offset = rank * base + min(rank, extra)
    return offset, count


def save_trace(records):
    """records: iterable of (name, numpy array) -> NDArray-list blob (bytes)."""
    records = list(records)
    out = bytearray(struct.pack("<QQQ", LIST_MAGIC, 0, len(records)))
    for name, _ in records:
        b = name.encode()
        out += struct.pack("<Q", len(b)) + b
    out += struct.pack("<Q", len(records))
    for _, arr in records:
        arr = np.asarray(arr)
        a = np.ascontiguousarray(arr)
        out += struct.pack("<QQiii", ARRAY_MAGIC, 0, 1, 0, arr.ndim)
        out += struct.pack("<BBH", _CODES[a.dtype.kind], a.dtype.itemsize * 8, 1)
        out += struct.pack("<%dq" % arr.ndim, *arr.shape)
        out += struct.pack("<q", a.nbytes) + a.tobytes()
    return bytes(out)


def load_trace(blob):
    """NDArray-list blob -> list of (name, numpy array), in file order."""
    blob = memoryview(blob)
    magic, _, n = struct.unpack_from("<QQQ", blob, 0)
    if magic != LIST_MAGIC:
        raise ValueError("not an NDArray-list blob")
    pos = 24
    names = []
    for _ in range(n):
        (ln,) = struct.unpack_from("<Q", blob, pos)
        pos += 8
        names.append(bytes(blob[pos:pos + ln]).decode())
        pos += ln
    (m,) = struct.unpack_from("<Q", blob, pos)
    pos += 8
    if m != n:
        raise ValueError("name / array count mismatch")
    out = []
    for i in range(m):
        magic, _, _dev, _id, ndim = struct.unpack_from("<QQiii", blob, pos)
        if magic != ARRAY_MAGIC:
            raise ValueError("bad NDArray magic in record %d" % i)
        pos += 28
        code, bits, lanes = struct.unpack_from("<BBH", blob, pos)
        pos += 4
        shape = struct.unpack_from("<%dq" % ndim, blob, pos)
        pos += 8 * ndim
        (nbytes,) = struct.unpack_from("<q", blob, pos)
        pos += 8
        dt = np.dtype("%s%d" % (_KINDS[code], bits))
        arr = np.frombuffer(blob[pos:pos + nbytes], dtype=dt).reshape(shape)
        pos += nbytes
        out.append((names[i], arr))
    return out


def const_record_names(graph_json):
    """Names of the trace records whose tensor is produced by a ``const`` node of the tachikoma JSON graph
    (weights, bias, scales, zero points): ``{op}____topo-index:{nid}____input-num:{i}`` for every kernel-node input
    that points at a const node (schema: src/runtime/contrib/json/json_node.h:116-129)."""
    import json
    nodes = json.loads(graph_json)["nodes"]
    out = set()
    for nid, n in enumerate(nodes):
        if n.get("op") != "kernel":
            continue
        for i, ref in enumerate(n.get("inputs", [])):
            if nodes[ref[0]].get("op") == "const":
                out.add("%s____topo-index:%d____input-num:%d" % (n["name"], nid, i))
    return out


def concat_traces(blobs, sample_offsets=None, graph_json=None):
    """Join per-rank trace blobs (rank order) into one blob.  Record keys get the suffix
    ``____rank:{r}____sample-offset:{o}`` so that every record stays addressable and carries the global
    index of its first sample.  With ``graph_json`` (the module's graph, identical on all ranks) the records of
    CONSTANT tensors are kept from rank 0 only; which records those are is read from the graph (their producing
    node is a ``const`` node), never inferred from input position or byte equality: two ranks that happen to see
    identical activations keep all their activation records.  Without ``graph_json`` nothing is dropped."""
    blobs = list(blobs)
    if sample_offsets is None:
        sample_offsets = [0] * len(blobs)
    consts = const_record_names(graph_json) if graph_json is not None else set()
    out = []
    for r, blob in enumerate(blobs):
        for name, arr in load_trace(blob):
            if r > 0 and name in consts:
                continue
            out.append(("%s____rank:%d____sample-offset:%d" % (name, r, sample_offsets[r]), arr))
    return save_trace(out)
