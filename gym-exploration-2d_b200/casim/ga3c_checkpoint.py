"""Reader for the TensorFlow-1 "V2" checkpoints the reference ships for GA3C-CADRL
(gym_collision_avoidance/envs/policies/GA3C_CADRL/checkpoints/*/network_*.{index,data-00000-of-00001}) -- the
counterpart of NetworkVPCore.simple_load (GA3C_CADRL/network.py:56-62) behind GA3CCADRLPolicy.initialize_network
(GA3CCADRLPolicy.py:21-32), without TensorFlow.

Format (tensorflow/core/util/tensor_bundle): `<prefix>.index` is a LevelDB-style sorted string table (blocks of
prefix-compressed key/value entries + restart array, 1-byte compression tag + 4-byte crc per block, 48-byte footer
with the metaindex / index block handles and the magic 0xdb4775248b80fb57); its values are BundleEntryProto messages
(dtype, shape, shard_id, offset, size, crc32c); `<prefix>.data-*` holds the raw little-endian tensor bytes.
Only what these checkpoints use is implemented: uncompressed blocks, DT_FLOAT / DT_INT32 / DT_INT64 tensors.
"""
import os
import struct

import numpy as np

_MAGIC = 0xdb4775248b80fb57
_DTYPES = {1: np.float32, 2: np.float64, 3: np.int32, 9: np.int64}


def _varint(buf, pos):
    out, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7f) << shift
        if not b & 0x80:
            return out, pos
        shift += 7


def _block(data, offset, size):
    """Decode one table block -> list of (key bytes, value bytes)."""
    raw = data[offset:offset + size]
    ctype = data[offset + size]
    if ctype != 0:
        raise ValueError("compressed index block (type %d) is not supported" % ctype)
    n_restarts = struct.unpack_from("<I", raw, len(raw) - 4)[0]
    end = len(raw) - 4 - 4 * n_restarts
    pos, key, out = 0, b"", []
    while pos < end:
        shared, pos = _varint(raw, pos)
        non_shared, pos = _varint(raw, pos)
        vlen, pos = _varint(raw, pos)
        key = key[:shared] + bytes(raw[pos:pos + non_shared])
        pos += non_shared
        out.append((key, bytes(raw[pos:pos + vlen])))
        pos += vlen
    return out


def _proto_fields(buf):
    """Minimal protobuf wire decoder -> list of (field number, wire type, value)."""
    pos, out = 0, []
    while pos < len(buf):
        tag, pos = _varint(buf, pos)
        f, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _varint(buf, pos)
        elif wt == 1:
            v = buf[pos:pos + 8]
            pos += 8
        elif wt == 2:
            n, pos = _varint(buf, pos)
            v = buf[pos:pos + n]
            pos += n
        elif wt == 5:
            v = buf[pos:pos + 4]
            pos += 4
        else:
            raise ValueError("unsupported protobuf wire type %d" % wt)
        out.append((f, wt, v))
    return out


def _bundle_entry(buf):
    e = {"dtype": 0, "shape": [], "shard_id": 0, "offset": 0, "size": 0}
    for f, wt, v in _proto_fields(buf):
        if f == 1:
            e["dtype"] = v
        elif f == 2:                                   # TensorShapeProto { repeated Dim dim = 2 { int64 size = 1 } }
            for f2, _, v2 in _proto_fields(v):
                if f2 == 2:
                    size = 0
                    for f3, _, v3 in _proto_fields(v2):
                        if f3 == 1:
                            size = v3
                    e["shape"].append(size)
        elif f == 3:
            e["shard_id"] = v
        elif f == 4:
            e["offset"] = v
        elif f == 5:
            e["size"] = v
    return e


def read_checkpoint(prefix):
    """-> {variable name: numpy array} of every tensor in the checkpoint `prefix`(.index / .data-00000-of-00001)."""
    with open(prefix + ".index", "rb") as f:
        idx = f.read()
    if struct.unpack_from("<Q", idx, len(idx) - 8)[0] != _MAGIC:
        raise ValueError("%s.index is not a tensor-bundle index (bad magic)" % prefix)
    footer = idx[len(idx) - 48:]
    pos = 0
    _, pos = _varint(footer, pos)          # metaindex handle (offset, size)
    _, pos = _varint(footer, pos)
    ioff, pos = _varint(footer, pos)       # index handle
    isize, pos = _varint(footer, pos)
    entries = {}
    for _, handle in _block(idx, ioff, isize):
        boff, p = _varint(handle, 0)
        bsize, p = _varint(handle, p)
        for key, val in _block(idx, boff, bsize):
            if key == b"":                 # BundleHeaderProto
                continue
            entries[key.decode()] = _bundle_entry(val)
    shards = {}
    out = {}
    for name, e in entries.items():
        if e["dtype"] not in _DTYPES:
            continue
        sid = e["shard_id"]
        if sid not in shards:
            n_shards = 1 + max(x["shard_id"] for x in entries.values())
            with open("%s.data-%05d-of-%05d" % (prefix, sid, n_shards), "rb") as f:
                shards[sid] = f.read()
        a = np.frombuffer(shards[sid], dtype=_DTYPES[e["dtype"]], count=int(np.prod(e["shape"], dtype=np.int64)) if e["shape"] else 1,
                          offset=e["offset"])
        out[name] = a.reshape(e["shape"]).copy()
    return out


# order of the parameter blob casim_ga3c_set_weights takes (include/casim.h): TF variable names of
# GA3C_CADRL/network.py:65-98 (NetworkVP_rnn) and :45-51 (heads)
GA3C_VARIABLES = (("rnn/lstm_cell/kernel:0", (71, 256)), ("rnn/lstm_cell/bias:0", (256,)),
                  ("layer1/kernel:0", (68, 256)), ("layer1/bias:0", (256,)),
                  ("layer2/kernel:0", (256, 256)), ("layer2/bias:0", (256,)),
                  ("fullyconnected1/kernel:0", (256, 256)), ("fullyconnected1/bias:0", (256,)),
                  ("logits_p/kernel:0", (256, 11)), ("logits_p/bias:0", (11,)),
                  ("logits_v/kernel:0", (256, 1)), ("logits_v/bias:0", (1,)))


def ga3c_weights_from_checkpoint(prefix):
    """float32 [170764] blob for BatchedSim.ga3c_set_weights / env.set_ga3c_weights from a reference checkpoint prefix,
    e.g. ``.../GA3C_CADRL/checkpoints/IROS18/network_01900000`` (the default of GA3CCADRLPolicy.initialize_network)."""
    tensors = read_checkpoint(prefix)
    parts = []
    for name, shape in GA3C_VARIABLES:
        if name not in tensors:
            raise KeyError("checkpoint %s has no variable %s" % (prefix, name))
        t = tensors[name]
        if tuple(t.shape) != shape:
            raise ValueError("%s: shape %r, expected %r" % (name, t.shape, shape))
        parts.append(np.ascontiguousarray(t, np.float32).ravel())
    return np.concatenate(parts)


def checkpoint_prefix(checkpoints_root, checkpt_dir="IROS18", checkpt_name="network_01900000"):
    """GA3CCADRLPolicy.initialize_network's path rule (GA3CCADRLPolicy.py:21-32) under an explicit root directory."""
    return os.path.join(checkpoints_root, checkpt_dir, checkpt_name)
