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


def concat_traces(blobs, sample_offsets=None):
    """Join per-rank trace blobs (rank order) into one blob.  Record keys get the suffix
    ``____rank:{r}____sample-offset:{o}`` so that every record stays addressable and carries the global
    index of its first sample; constants (identical on all ranks) are kept from rank 0 only."""
    blobs = list(blobs)
    if sample_offsets is None:
        sample_offsets = [0] * len(blobs)
    merged = []
    seen_const = {}
    for r, blob in enumerate(blobs):
        for name, arr in load_trace(blob):
            key = "%s____rank:%d____sample-offset:%d" % (name, r, sample_offsets[r])
            merged.append((key, arr, name, r))
    # drop records of ranks > 0 that are byte-identical to rank 0's record of the same name (weights, scales)
    rank0 = {name: arr for _, arr, name, r in merged if r == 0}
    out = []
    for key, arr, name, r in merged:
        if r > 0 and name in rank0 and rank0[name].shape == arr.shape and rank0[name].dtype == arr.dtype \
                and name.find("input-num:0") < 0 and np.array_equal(rank0[name], arr) and _is_param(name, seen_const):
            continue
        out.append((key, arr))
    return save_trace(out)


def _is_param(name, cache):
    # inputs 1.. of a node are its parameters (weights, bias, scales, zero points) unless they are activations;
    # byte-equality with rank 0 (checked by the caller) is the actual criterion, this only excludes outputs.
    return "____output-num:" not in name
