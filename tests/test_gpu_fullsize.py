"""Full-size GPU parity on the shapes bench.py launches (BASELINE.json configs[2] / the north_star target `full`, and
configs[4] per GPU): the kernels run at 40 960 / 20 480 agents -- grid sizes, 64-bit index arithmetic, the per-env-map
branch (n_maps == E) and the weight-replica rotation of the tensor-core forward that the 96-env parity batches never
reach.  The oracle checks a 256-env sample; size-independent properties cover the rest: shard invariance (any
contiguous block of environments gives bit-identical results on its own) and determinism."""
import os
import sys

import numpy as np
import pytest

import helpers as H
from casim import layout as L

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _cfg(**kw):
    from casim import lib
    c = lib.default_config()
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def _np(sim):
    return {k: (None if v is None or isinstance(v, list) else v.detach().cpu().numpy()) for k, v in sim.buf.items()}


SENSORS = ("other_agents_states", "laserscan", "local_grid")


@pytest.mark.parametrize("per_env_maps", [False, True])
def test_full_workload_full_size(oracle_lib, per_env_maps):
    """`full` = 4096 envs x 10 agents, OtherAgentsStates + LaserScan + OccupancyGrid on the bench's own scenario
    (bench.scenario('full')), shared map and per-env maps."""
    import torch
    import bench
    from casim.sim import BatchedSim
    oracle = oracle_lib
    E, A = 4096, 10
    init, types, n, rects = bench.scenario("full", E, A, seed=4321)
    maps, rects_g = H.unpack_maps()
    rng = np.random.default_rng(5)
    cfg = _cfg(game_over_mode=1)

    def setup(sim, lo, hi):
        if per_env_maps:
            sim.set_map(mstack[lo:hi])
        else:
            sim.set_obstacles(rects)
        sim.reset(init[lo:hi], types[lo:hi], n[lo:hi])

    if per_env_maps:
        which = rng.integers(1, 3, E)
        mstack = np.stack([maps[int(w)] for w in which])
    else:
        mstack = oracle.rasterize(cfg, rects.astype(np.float64))
    sim = BatchedSim(E, A, cfg=cfg, sensors=SENSORS)
    setup(sim, 0, E)
    K = 4
    for _ in range(K):
        sim.step()
    torch.cuda.synchronize()
    keep = {k: sim.buf[k].clone() for k in SENSORS + ("rewards", "done", "game_over")}
    keep_state = sim.state.clone()
    # ---- shard invariance: what 2 GPUs would each compute, on their own handles ----
    for lo, hi in ((0, 1000), (1000, 4096)):
        s2 = BatchedSim(hi - lo, A, cfg=cfg, sensors=SENSORS)
        setup(s2, lo, hi)
        for _ in range(K):
            s2.step()
        torch.cuda.synchronize()
        assert torch.equal(s2.state, keep_state[:, lo:hi])
        for k, v in keep.items():
            assert torch.equal(s2.buf[k], v[lo:hi]), k
        s2.close()
    # ---- oracle on a 256-env sample: one more step from the GPU state ----
    sel = np.sort(rng.choice(E, 256, replace=False))
    ostate = np.ascontiguousarray(sim.state_numpy()[sel])
    omaps = mstack[sel] if per_env_maps else mstack
    oobs = oracle.step(cfg, ostate, n_agents=n[sel], maps=omaps, n_threads=8)
    sim.step()
    torch.cuda.synchronize()
    g = {k: (None if v is None else v[sel]) for k, v in _np(sim).items() if k != "stats"}
    H.compare_step("full sample", n[sel], sim.state_numpy()[sel], g, ostate, oobs)
    # the sensors saw something: occupied AND free grid cells, laser hits
    assert 0.02 < g["local_grid"].mean() < 0.98 and (g["laserscan"] > 0).any()
    # ---- determinism of the sensor kernels at this size ----
    a = sim.buf["local_grid"].clone()
    b = sim.buf["laserscan"].clone()
    sim.run_sensor(L.SENSOR_OCCUPANCY_GRID)
    sim.run_sensor(L.SENSOR_LASERSCAN)
    torch.cuda.synchronize()
    assert torch.equal(a, sim.buf["local_grid"]) and torch.equal(b, sim.buf["laserscan"])
    sim.close()


def test_c5_full_size_per_gpu(oracle_lib):
    """C5 per-GPU shape: 1024 envs x 20 agents (20 480 agents, 143 CTAs of the tensor-core forward).  fp32 forward vs the
    oracle on a 128-env sample; tensor-core forward: shard invariance (bitwise) + agreement with the fp32 forward."""
    import torch
    import bench
    from casim.sim import BatchedSim
    oracle = oracle_lib
    E, A = 1024, 20
    init, types, n, _ = bench.scenario("c5", E, A, seed=99)
    cfg = _cfg(game_over_mode=1, near_goal_threshold=0.2)
    w = bench.ga3c_weights(0)
    sim = BatchedSim(E, A, cfg=cfg, sensors=("other_agents_states",))
    sim.reset(init, types, n)
    sim.ga3c_set_weights(w)
    for _ in range(3):
        sim.step(policy_actions=sim.ga3c_forward(use_tensor_cores=False))
    torch.cuda.synchronize()
    rng = np.random.default_rng(3)
    sel = np.sort(rng.choice(E, 128, replace=False))
    state = sim.state_numpy()
    o = oracle.ga3c_forward(cfg, np.ascontiguousarray(state[sel]), w, n_agents=n[sel], n_threads=8)
    sim.ga3c_forward(use_tensor_cores=False)
    torch.cuda.synchronize()
    p32 = sim.buf["probs"].cpu().numpy()
    i32 = sim.buf["action_idx"].cpu().numpy()
    H.assert_close(sim.buf["nn_input"].cpu().numpy()[sel], o["nn_input"], "nn_input", rtol=1e-6, atol=1e-6)
    act = o["action_idx"] >= 0                      # finished agents are not queried (collision_avoidance_env.py:298-300)
    assert np.array_equal(i32[sel] >= 0, act) and act.mean() > 0.5
    assert np.abs(p32[sel][act] - o["probs"][act]).max() < 2e-5
    srt = np.sort(o["probs"], axis=-1)
    dec = act & ((srt[..., -1] - srt[..., -2]) > 1e-4)
    assert np.array_equal(i32[sel][dec], o["action_idx"][dec]) and dec.sum() > 0.5 * act.sum()
    # ---- tensor-core forward on the same states ----
    sim.ga3c_forward(use_tensor_cores=True)
    torch.cuda.synchronize()
    ptc = sim.buf["probs"].clone()
    itc = sim.buf["action_idx"].clone()
    atc = sim.buf["policy_actions"].clone()
    sim.ga3c_forward(use_tensor_cores=True)
    torch.cuda.synchronize()
    assert torch.equal(ptc, sim.buf["probs"]) and torch.equal(itc, sim.buf["action_idx"])          # deterministic
    p32t = torch.from_numpy(p32).cuda()
    i32t = torch.from_numpy(i32).cuda()
    for lo, hi in ((0, 300), (300, 1024)):
        s2 = BatchedSim(hi - lo, A, cfg=cfg, sensors=("other_agents_states",))
        s2.reset(init[lo:hi], types[lo:hi], n[lo:hi])
        s2.ga3c_set_weights(w)
        s2.state.copy_(sim.state[:, lo:hi])
        lv = itc[lo:hi] >= 0                                   # probabilities are written for the queried agents only
        # the default (fp32) forward: shard invariance is bitwise
        s2.ga3c_forward(use_tensor_cores=False)
        torch.cuda.synchronize()
        assert torch.equal(s2.buf["action_idx"], i32t[lo:hi])
        assert torch.equal(s2.buf["probs"][lv], p32t[lo:hi][lv])
        # the opt-in bf16 tensor-core forward: an agent's tile-mates change with the shard boundaries and the number of
        # LSTM steps a tile runs is its longest sequence, which shows in the last bits of the bf16 recurrence
        s2.ga3c_forward(use_tensor_cores=True)
        torch.cuda.synchronize()
        assert torch.equal(s2.buf["action_idx"] >= 0, lv)
        dp = (s2.buf["probs"][lv] - ptc[lo:hi][lv]).abs().max().item()
        assert dp < 5e-4, ("tensor-core forward: shard-to-shard difference", lo, hi, dp)
        s2.close()
    ptc = ptc.cpu().numpy()
    itc = itc.cpu().numpy()
    import test_gpu_ga3c as TG
    live = i32 >= 0
    assert np.array_equal(itc >= 0, live)
    assert np.abs(ptc[live] - p32[live]).max() < TG.TC_PROB_TOL
    srt = np.sort(p32, axis=-1)
    dec = live & ((srt[..., -1] - srt[..., -2]) > TG.TC_MIN_MARGIN)
    assert np.array_equal(itc[dec], i32[dec])
    assert (itc[live] == i32[live]).mean() >= TG.TC_MIN_AGREE
    sim.close()
