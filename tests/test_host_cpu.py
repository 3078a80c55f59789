"""CPU: host-side logic of the drop-in boundary -- plugin descriptors -> type words, scenario tensors from Agent
objects, obstacle formats, shard arithmetic and the statistics all-reduce with world_size-2 gloo."""
import os
import socket
import sys

import numpy as np
import pytest

from casim import layout as L
from casim import plugins as P
from casim.plugins import Agent


def test_agent_descriptor_to_tensors():
    from casim.env import BatchedCollisionAvoidanceEnv
    a0 = Agent(-3, -3, 3, 3, 0.5, 1.0, 0.5, P.NonCooperativePolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 0)
    a1 = Agent(3, 3, -3, -3, 0.5, 1.0, None, P.StaticPolicy, P.FirstOrderDynamics,
               [P.OtherAgentsStatesSensor, P.LaserScanSensor, P.OccupancyGridSensor], 1)
    assert L.decode_type(a0.type_word) == (L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, 1)
    assert L.decode_type(a1.type_word) == (L.POLICY_STATIC, L.DYN_FIRST_ORDER, 7)
    env = BatchedCollisionAvoidanceEnv(num_envs=1, max_agents=4)
    env.set_agents([a0, a1])
    init, types, n = env._scenario()
    assert init.shape == (1, 4, 7) and list(n) == [2]
    assert list(init[0, 0]) == [-3, -3, 3, 3, 0.5, 1.0, 0.5] and np.isnan(init[0, 1, 6])
    assert types[0, 0] == a0.type_word and types[0, 1] == a1.type_word
    # the reference's obstacle format: corner index 1 = top-left, 3 = bottom-right (Map.py:117-118)
    env.set_static_map([[(10, 10), (2, 10), (2, 2), (10, 2)]])
    assert env.obstacles.tolist() == [[2, 10, 10, 2]]
    env.set_static_map(np.array([[2, 10, 10, 2.0]]))
    assert env.obstacles.tolist() == [[2, 10, 10, 2]]
    with pytest.raises(RuntimeError):
        a0.policy.find_next_action(None, None, 0)        # no Python compute path behind the descriptors
    with pytest.raises(RuntimeError):
        BatchedCollisionAvoidanceEnv(1, 2).step({})      # reset() first


def test_shard_bounds_cover_and_are_contiguous():
    from casim.dist import shard_bounds, shard_scenario
    for E in (1, 7, 4096, 4097):
        for W in (1, 2, 3, 8):
            blocks = [shard_bounds(E, W, r) for r in range(W)]
            assert blocks[0][0] == 0 and blocks[-1][1] == E
            assert all(blocks[r][1] == blocks[r + 1][0] for r in range(W - 1))
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1
    init, types, n, rects = np.zeros((10, 3, 7)), np.zeros((10, 3), np.int32), np.arange(10), np.zeros((4, 4))
    a, b, c, d = shard_scenario((init, types, n, rects), 2, 1)
    assert a.shape[0] == 5 and list(c) == [5, 6, 7, 8, 9] and d.shape == (4, 4)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from casim.dist import StatsReducer, shard_bounds
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    E = 11
    lo, hi = shard_bounds(E, world, rank)
    # per-env "statistics" that only the owner of the env contributes
    per_env = torch.arange(E, dtype=torch.float32).unsqueeze(1) * torch.ones(1, 8)
    local = per_env[lo:hi].sum(0)
    red = StatsReducer(device=None)
    red.reduce(local)
    total = red.wait()
    # max-over-ranks timing reduction used by bench.py
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    q.put((rank, total.tolist(), float(t.item()), (lo, hi)))
    dist.destroy_process_group()


def test_stats_all_reduce_world2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect = [float(sum(range(11)))] * 8
    for rank, total, tmax, (lo, hi) in res:
        assert total == expect and tmax == 2.0
    blocks = sorted(r[3] for r in res)
    assert blocks == [(0, 5), (5, 11)]


# ---------------------------------------------------------------------------------------------
# GA3C-CADRL checkpoint reader
def _write_bundle(prefix, tensors):
    """A minimal tensor-bundle writer (one data block, no compression) to exercise the reader without TensorFlow."""
    import struct

    def varint(v):
        out = bytearray()
        while True:
            b = v & 0x7f
            v >>= 7
            out.append(b | (0x80 if v else 0))
            if not v:
                return bytes(out)

    def block(entries):
        body, prev = bytearray(), b""
        for k, v in entries:            # prefix compression against the previous key (restart only at entry 0)
            shared = 0
            while shared < min(len(prev), len(k)) and prev[shared] == k[shared]:
                shared += 1
            body += varint(shared) + varint(len(k) - shared) + varint(len(v)) + k[shared:] + v
            prev = k
        body += struct.pack("<I", 0) + struct.pack("<I", 1)
        return bytes(body)

    data, entries = bytearray(), [(b"", b"\x08\x01")]
    for name in sorted(tensors):
        t = np.asarray(tensors[name])          # (ascontiguousarray would turn a scalar into shape (1,))
        dt = {np.dtype(np.float32): 1, np.dtype(np.int32): 3}[t.dtype]
        shape = b"".join(b"\x12" + varint(len(d)) + d for d in [b"\x08" + varint(s) for s in t.shape])
        e = b"\x08" + varint(dt) + b"\x12" + varint(len(shape)) + shape + b"\x20" + varint(len(data)) + b"\x28" + varint(t.nbytes)
        entries.append((name.encode(), e))
        data += t.tobytes()
    blk = block(entries)
    out = bytearray(blk) + b"\x00" + b"\x00" * 4                      # block + compression tag + crc (unchecked)
    meta_off = len(out)
    meta = block([])
    out += meta + b"\x00" + b"\x00" * 4
    idx_off = len(out)
    idx = block([(b"~", varint(0) + varint(len(blk)))])
    out += idx + b"\x00" + b"\x00" * 4
    footer = varint(meta_off) + varint(len(meta)) + varint(idx_off) + varint(len(idx))
    out += footer + b"\x00" * (40 - len(footer)) + struct.pack("<Q", 0xdb4775248b80fb57)
    open(prefix + ".index", "wb").write(bytes(out))
    open(prefix + ".data-00000-of-00001", "wb").write(bytes(data))


def test_checkpoint_reader_roundtrip(tmp_path):
    from casim.ga3c_checkpoint import GA3C_VARIABLES, ga3c_weights_from_checkpoint, read_checkpoint
    rng = np.random.default_rng(0)
    tensors = {name: rng.standard_normal(shape).astype(np.float32) for name, shape in GA3C_VARIABLES}
    tensors["step:0"] = np.array(7, np.int32)
    tensors["layer1/kernel/Adam:0"] = rng.standard_normal((68, 256)).astype(np.float32)      # optimiser slots are ignored
    prefix = str(tmp_path / "network_00000007")
    _write_bundle(prefix, tensors)
    got = read_checkpoint(prefix)
    assert set(got) == set(tensors)
    for k in tensors:
        assert np.array_equal(got[k], tensors[k]), k
    w = ga3c_weights_from_checkpoint(prefix)
    assert w.shape == (170764,) and np.array_equal(w[:71 * 256], tensors["rnn/lstm_cell/kernel:0"].ravel())
    assert np.array_equal(w[-257:-1], tensors["logits_v/kernel:0"].ravel())


def test_checkpoint_reader_on_reference_checkpoint():
    """The shipped IROS18 checkpoint (only where /root/reference exists) against tests/golden/ga3c_iros18.npz."""
    import os
    import helpers as H
    from casim.ga3c_checkpoint import GA3C_VARIABLES, checkpoint_prefix, ga3c_weights_from_checkpoint, read_checkpoint
    root = "/root/reference/gym_collision_avoidance/envs/policies/GA3C_CADRL/checkpoints"
    z = H.load("ga3c_iros18.npz")
    assert z["weights"].shape == (170764,) and np.isfinite(z["weights"]).all()
    if not os.path.isdir(root):
        pytest.skip("reference checkpoints not present on this machine")
    prefix = checkpoint_prefix(root)                       # defaults of GA3CCADRLPolicy.initialize_network
    assert np.array_equal(ga3c_weights_from_checkpoint(prefix), z["weights"])
    t = read_checkpoint(prefix)
    np.testing.assert_allclose([t[n].astype(np.float64).sum() for n, _ in GA3C_VARIABLES], z["sums"], rtol=0, atol=0)
    for d, name in (("ICRA21", "network_01990000"), ("run-20190727_192048-qedrf08y", "network_01900000")):
        assert ga3c_weights_from_checkpoint(checkpoint_prefix(root, d, name)).shape == (170764,)


def test_cell_mask_pack_roundtrip():
    """pack_cell_mask / unpack_cell_mask: bit (j * 64 + i) of 128 words, the layout of casim_explore_rollouts."""
    from casim.sim import pack_cell_mask, unpack_cell_mask
    rng = np.random.default_rng(0)
    m = rng.random((3, 2, 60, 60)) < 0.2
    w = pack_cell_mask(m)
    assert w.shape == (3, 2, 128) and w.dtype == np.int32
    assert np.array_equal(unpack_cell_mask(w, 60), m)
    one = np.zeros((60, 60), bool)
    one[7, 37] = True                       # y cell 7, x cell 37 -> bit 7*64+37 = 485 -> word 15, bit 5
    assert pack_cell_mask(one).view(np.uint32)[15] == (1 << 5) and pack_cell_mask(one).view(np.uint32).sum() == 32
