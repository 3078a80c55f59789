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
