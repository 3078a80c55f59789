"""GPU parity tests: libcasim.so (through the C ABI) against (a) the golden vectors recorded from the real
reference and (b) the CPU oracle on seeded random batches.  Flags, done, closest-N ordering, laser hit
indices and occupancy cells bit-exact; continuous values within 1e-5 (helpers.RTOL / ATOL)."""
import numpy as np
import pytest

import helpers as H
from casim import layout as L

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_mod():
    import torch
    assert torch.cuda.is_available()
    return torch


def make_sim(E, A, cfg, sensors=("other_agents_states", "laserscan", "local_grid")):
    from casim.sim import BatchedSim
    return BatchedSim(E, A, cfg=cfg, sensors=sensors)


def gcfg(mode=0, **kw):
    from casim import lib
    c = lib.default_config()
    c.game_over_mode = int(mode)
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def bufs_numpy(sim):
    return {k: (None if v is None else v.detach().cpu().numpy()) for k, v in sim.buf.items()}


def test_library_loaded_and_default_config_matches_oracle(oracle_lib, torch_mod):
    from casim import lib
    a, b = lib.default_config(), oracle_lib.default_config()
    assert bytes(a) == bytes(b)


@pytest.mark.parametrize("name", ["core_single_step.npz", "sensors_single_step.npz", "dynamics_extra_single_step.npz",
                                  "learning_single_step.npz"])
def test_golden_single_step(oracle_lib, torch_mod, name):
    """Every golden case (reference outputs) through casim_step, grouped by (game-over mode, map)."""
    torch = torch_mod
    z = H.load(name)
    maps, rects = H.unpack_maps()
    ncases, A = z["state_in"].shape[:2]
    for mode in range(3):
        for mid in range(3):
            sel = np.where((z["game_over_mode"] == mode) & (z["map_id"] == mid))[0]
            if len(sel) == 0:
                continue
            E = len(sel)
            sim = make_sim(E, A, gcfg(mode))
            if mid:
                sim.set_obstacles(rects[mid].astype(np.float32))
            sim.bind(types=z["types"][sel], n_agents=z["n_agents"][sel])
            sim.set_trajectories(z["traj"][sel])
            sim.load_state_numpy(z["state_in"][sel])
            ext = torch.as_tensor(z["ext_actions"][sel], dtype=torch.float32).cuda()
            sim.step(ext_actions=ext, dt=0.1)
            torch.cuda.synchronize()
            g = bufs_numpy(sim)
            gs = sim.state_numpy()
            n = z["n_agents"][sel]
            what = "%s mode %d map %d" % (name, mode, mid)

            class Ref:
                pass
            ref = Ref()
            for k in ("done", "game_over", "collision_flags", "nearest_dist", "rewards", "obs_misc",
                      "other_agents_states", "laserscan", "local_grid"):
                setattr(ref, k, z[k][sel])
            ref.obs_misc = ref.obs_misc.copy()
            if not name.startswith("learning"):               # older fixtures: use_ppo was not recorded by the generator
                ref.obs_misc[..., 3] = g["obs_misc"][..., 3]
            H.compare_step(what, n, gs, g, z["state_out"][sel], ref)
            sim.close()


def test_golden_reset(oracle_lib, torch_mod):
    z = H.load("reset.npz")
    maps, rects = H.unpack_maps()
    A = z["init"].shape[1]
    for mid in range(3):
        sel = np.where(z["map_id"] == mid)[0]
        if len(sel) == 0:
            continue
        sim = make_sim(len(sel), A, gcfg(2))
        if mid:
            sim.set_obstacles(rects[mid].astype(np.float32))
        sim.reset(z["init"][sel], z["types"][sel], z["n_agents"][sel])
        torch_mod.cuda.synchronize()
        g = bufs_numpy(sim)
        n = z["n_agents"][sel]
        H.assert_state_close(sim.state_numpy(), z["state_out"][sel], n, "reset map %d" % mid)

        class Ref:
            pass
        ref = Ref()
        for k in ("obs_misc", "other_agents_states", "laserscan", "local_grid"):
            setattr(ref, k, z[k][sel])
        ref.obs_misc = ref.obs_misc.copy()
        ref.obs_misc[..., 3] = g["obs_misc"][..., 3]
        H.compare_obs("reset map %d" % mid, n, g, ref)
        sim.close()


def test_map_raster_and_occupancy_grid_golden(oracle_lib, torch_mod):
    """Rectangles -> raster equals the reference Map; OccupancyGridSensor equals cv2.warpAffine on 400 poses."""
    torch = torch_mod
    maps, rects = H.unpack_maps()
    z = H.load("occgrid.npz")
    grids = np.unpackbits(z["grids"], axis=-1)[..., :60]
    t = L.encode_type(L.POLICY_STATIC, L.DYN_UNICYCLE, L.SENSOR_OCCUPANCY_GRID)
    for mid in (1, 2):
        sel = np.where(z["map_id"] == mid)[0]
        sim = make_sim(len(sel), 1, gcfg(0), sensors=("local_grid",))
        m = sim.set_obstacles(rects[mid].astype(np.float32), want_map=True)
        assert np.array_equal(m[0].cpu().numpy(), maps[mid])
        st = np.zeros((len(sel), 1, L.STATE_WORDS))
        st[:, 0, L.PX], st[:, 0, L.PY], st[:, 0, L.HEADING] = z["poses"][sel].T
        st[:, 0, L.TYPE] = t
        st[:, 0, L.RADIUS] = 0.5
        sim.bind(types=np.full((len(sel), 1), t, np.int32))
        sim.load_state_numpy(st)
        sim.observe()
        torch.cuda.synchronize()
        got = sim.buf["local_grid"].cpu().numpy()[:, 0]
        assert int((got != grids[sel]).sum()) == 0
        # set_map from the raster gives the same planes
        sim2 = make_sim(len(sel), 1, gcfg(0), sensors=("local_grid",))
        sim2.set_map(maps[mid])
        sim2.bind(types=np.full((len(sel), 1), t, np.int32))
        sim2.load_state_numpy(st)
        sim2.observe()
        torch.cuda.synchronize()
        assert np.array_equal(sim2.buf["local_grid"].cpu().numpy()[:, 0], got)
        sim.close()
        sim2.close()


def test_c1_lockstep_trajectory(oracle_lib, torch_mod):
    """Config C1 (2-agent example): 100 steps; at every step the oracle is re-seeded with the GPU's fp32
    state, so each step is a single-step parity check on identical inputs; the first done step must be 37
    as in the reference."""
    oracle = oracle_lib
    z = H.load("c1_two_agents.npz")
    cfg = gcfg(2)
    sim = make_sim(1, 2, cfg, sensors=("other_agents_states",))
    sim.reset(z["init"][None], z["types"][None])
    torch_mod.cuda.synchronize()
    ocfg = oracle.default_config()
    ocfg.game_over_mode = 2
    ostate, oobs = oracle.reset(ocfg, z["init"][None], z["types"][None])
    H.assert_state_close(sim.state_numpy(), ostate, [2], "c1 reset")
    first_done = None
    for t in range(100):
        ostate = sim.state_numpy()
        oobs = oracle.step(ocfg, ostate, dt=0.1)
        sim.step()
        torch_mod.cuda.synchronize()
        H.compare_step("c1 t=%d" % t, [2], sim.state_numpy(), bufs_numpy(sim), ostate, oobs)
        # and against the free-running reference trajectory, loosely (fp32 state storage drifts ~1e-6)
        H.assert_close(sim.state_numpy()[0][:, [L.PX, L.PY]], z["states"][t + 1][:, [L.PX, L.PY]], "c1 ref t=%d" % t,
                       rtol=1e-4, atol=1e-4)
        assert np.array_equal(sim.buf["done"].cpu().numpy()[0], z["done"][t]), t
        if first_done is None and z["done"][t].any():
            first_done = t
    assert first_done == 37
    sim.close()


@pytest.mark.parametrize("A,E,seed", [(10, 512, 1), (2, 257, 2), (4, 300, 3), (20, 128, 4), (32, 64, 5), (7, 200, 6),
                                      (1, 40, 7)])
def test_random_batch_core_vs_oracle(oracle_lib, torch_mod, A, E, seed):
    torch = torch_mod
    oracle = oracle_lib
    rng = np.random.default_rng(seed)
    b = H.random_batch(rng, E, A, sensor_choices=(0, 1, 1), n_min=1)
    for mode in (0, 1, 2):
        cfg = gcfg(mode, max_other_agents_rows=min(9, max(A - 1, 1)) if A != 20 else 19,
                   max_other_agents_observed=9 if A != 20 else 19, collision_av_w_static_agent=int(mode == 1))
        sim = make_sim(E, A, cfg, sensors=("other_agents_states",))
        sim.bind(types=b["types"], n_agents=b["n_agents"])
        sim.set_trajectories(b["traj"])
        sim.load_state_numpy(b["state"])
        ext = torch.as_tensor(b["ext"], dtype=torch.float32).cuda()
        sim.step(ext_actions=ext)
        torch.cuda.synchronize()
        ostate = b["state"].copy()
        oobs = oracle.step(cfg, ostate, ext_actions=b["ext"], n_agents=b["n_agents"], traj=b["traj"])
        H.compare_step("random A=%d mode=%d" % (A, mode), b["n_agents"], sim.state_numpy(), bufs_numpy(sim), ostate, oobs)
        # per-step statistics of the shard
        H.assert_close(sim.buf["stats"].cpu().numpy(), oobs.stats, "stats", rtol=1e-5, atol=1e-4)
        sim.close()


@pytest.mark.parametrize("per_env_maps", [False, True])
def test_random_batch_sensors_vs_oracle(oracle_lib, torch_mod, per_env_maps):
    torch = torch_mod
    oracle = oracle_lib
    rng = np.random.default_rng(11)
    E, A = 96, 10
    b = H.random_batch(rng, E, A, sensor_choices=(0, 1, 2, 3, 4, 5, 6, 7, 7), spread=9.0)
    maps, rects = H.unpack_maps()
    cfg = gcfg(0)
    sim = make_sim(E, A, cfg)
    if per_env_maps:
        which = rng.integers(1, 3, E)
        mstack = np.stack([maps[int(w)] for w in which])
        sim.set_map(mstack)
    else:
        mstack = maps[1]
        sim.set_obstacles(rects[1].astype(np.float32))
    sim.bind(types=b["types"], n_agents=b["n_agents"])
    sim.set_trajectories(b["traj"])
    sim.load_state_numpy(b["state"])
    ext = torch.as_tensor(b["ext"], dtype=torch.float32).cuda()
    # observe (sensors from the stored state), then a step (sensors from the fp64 post-move pose)
    sim.observe()
    torch.cuda.synchronize()
    oobs = oracle.observe(cfg, b["state"], n_agents=b["n_agents"], maps=mstack)
    H.compare_obs("observe", b["n_agents"], bufs_numpy(sim), oobs)
    sim.step(ext_actions=ext)
    torch.cuda.synchronize()
    ostate = b["state"].copy()
    oobs = oracle.step(cfg, ostate, ext_actions=b["ext"], n_agents=b["n_agents"], traj=b["traj"], maps=mstack)
    H.compare_step("sensors step", b["n_agents"], sim.state_numpy(), bufs_numpy(sim), ostate, oobs)
    sim.close()


@pytest.mark.parametrize("A", [3, 12, 16, 24, 32])
def test_agent_count_instantiations_vs_oracle(oracle_lib, torch_mod, A):
    """Every k_core<., AP> instantiation the other tests do not reach (AP = 4 via an odd count, 12, 16, 24 and the
    CASIM_MAX_AGENTS = 32 maximum): ragged n_agents, all sensors, observe + 6 lock-step steps on a shared map."""
    torch = torch_mod
    oracle = oracle_lib
    rng = np.random.default_rng(100 + A)
    E = 24
    b = H.random_batch(rng, E, A, sensor_choices=(0, 1, 3, 7, 7), spread=9.0)
    b["n_agents"][0] = A                       # the full width at least once, a ragged rest
    assert int(b["n_agents"].min()) < A
    maps, rects = H.unpack_maps()
    cfg = gcfg(0)
    sim = make_sim(E, A, cfg)
    sim.set_obstacles(rects[1].astype(np.float32))
    sim.bind(types=b["types"], n_agents=b["n_agents"])
    sim.set_trajectories(b["traj"])
    sim.load_state_numpy(b["state"])
    ext = torch.as_tensor(b["ext"], dtype=torch.float32).cuda()
    sim.observe()
    torch.cuda.synchronize()
    oobs = oracle.observe(cfg, b["state"], n_agents=b["n_agents"], maps=maps[1])
    H.compare_obs("observe A=%d" % A, b["n_agents"], bufs_numpy(sim), oobs)
    for t in range(6):
        ostate = sim.state_numpy()
        oobs = oracle.step(cfg, ostate, ext_actions=b["ext"], n_agents=b["n_agents"], traj=b["traj"], maps=maps[1])
        sim.step(ext_actions=ext)
        torch.cuda.synchronize()
        H.compare_step("A=%d t=%d" % (A, t), b["n_agents"], sim.state_numpy(), bufs_numpy(sim), ostate, oobs)
    sim.close()


def test_empty_and_single_agent_environments_vs_oracle(oracle_lib, torch_mod):
    """Ragged down to nothing: environments with 0 agents (game_over at once, all(...) of an empty list,
    collision_avoidance_env.py:700-717) and with 1 agent (no pairs, no OtherAgentsStates rows) next to full ones."""
    torch = torch_mod
    oracle = oracle_lib
    rng = np.random.default_rng(5)
    E, A = 40, 6
    b = H.random_batch(rng, E, A, sensor_choices=(0, 1, 3, 7, 7), spread=9.0, n_min=0)
    b["n_agents"][:6] = [0, 1, 0, 1, 2, A]
    maps, rects = H.unpack_maps()
    cfg = gcfg(0)
    sim = make_sim(E, A, cfg)
    sim.set_obstacles(rects[1].astype(np.float32))
    sim.bind(types=b["types"], n_agents=b["n_agents"])
    sim.set_trajectories(b["traj"])
    sim.load_state_numpy(b["state"])
    ext = torch.as_tensor(b["ext"], dtype=torch.float32).cuda()
    sim.observe()
    torch.cuda.synchronize()
    oobs = oracle.observe(cfg, b["state"], n_agents=b["n_agents"], maps=maps[1])
    H.compare_obs("observe", b["n_agents"], bufs_numpy(sim), oobs)
    for t in range(4):
        ostate = sim.state_numpy()
        oobs = oracle.step(cfg, ostate, ext_actions=b["ext"], n_agents=b["n_agents"], traj=b["traj"], maps=maps[1])
        sim.step(ext_actions=ext)
        torch.cuda.synchronize()
        H.compare_step("t=%d" % t, b["n_agents"], sim.state_numpy(), bufs_numpy(sim), ostate, oobs)
        g = bufs_numpy(sim)
        empty = b["n_agents"] == 0
        assert g["game_over"][empty].all() and np.array_equal(g["game_over"], oobs.game_over)
    sim.close()


def test_auto_reset_and_multi_step_vs_oracle(oracle_lib, torch_mod):
    """30 lock-step steps with auto-reset: a game-over environment restarts from its initial condition."""
    oracle = oracle_lib
    rng = np.random.default_rng(21)
    E, A = 64, 6
    b = H.random_batch(rng, E, A, sensor_choices=(1,), policies=[L.POLICY_NONCOOPERATIVE, L.POLICY_STATIC,
                                                                 L.POLICY_NONCOOPERATIVE], dynamics=[L.DYN_UNICYCLE], spread=3.0)
    cfg = gcfg(0, auto_reset=1)
    sim = make_sim(E, A, cfg, sensors=("other_agents_states",))
    sim.reset(b["init"], b["types"], b["n_agents"])
    torch_mod.cuda.synchronize()
    prev_go = np.zeros((E,), np.uint8)
    n_resets = 0
    for t in range(30):
        ostate = sim.state_numpy()
        oobs = oracle.step(cfg, ostate, n_agents=b["n_agents"], init=b["init"], types0=b["types"], prev_game_over=prev_go)
        sim.step()
        torch_mod.cuda.synchronize()
        H.compare_step("auto-reset t=%d" % t, b["n_agents"], sim.state_numpy(), bufs_numpy(sim), ostate, oobs)
        n_resets += int(prev_go.sum())
        prev_go = oobs.game_over.copy()
    assert n_resets > 0
    sim.close()


def test_c2_full_size_properties(oracle_lib, torch_mod):
    """Config C2 size (4096 envs x 10 agents): shard invariance (any contiguous block of environments gives
    bit-identical results on its own), determinism, and agreement with the oracle on a sample of envs."""
    torch = torch_mod
    oracle = oracle_lib
    rng = np.random.default_rng(31)
    E, A = 4096, 10
    b = H.random_batch(rng, E, A, sensor_choices=(1,), policies=[L.POLICY_NONCOOPERATIVE] * 4 + [L.POLICY_STATIC],
                       dynamics=[L.DYN_UNICYCLE], spread=8.0, n_min=10)
    cfg = gcfg(0)
    sim = make_sim(E, A, cfg, sensors=("other_agents_states",))
    sim.reset(b["init"], b["types"], b["n_agents"])
    for _ in range(5):
        sim.step()
    torch.cuda.synchronize()
    full_state = sim.state.clone()
    full_oas = sim.buf["other_agents_states"].clone()
    full_rew = sim.buf["rewards"].clone()
    # same thing in two shards of 2048 environments (what two GPUs would each do)
    for lo, hi in ((0, 2048), (2048, 4096)):
        s2 = make_sim(hi - lo, A, cfg, sensors=("other_agents_states",))
        s2.reset(b["init"][lo:hi], b["types"][lo:hi], b["n_agents"][lo:hi])
        for _ in range(5):
            s2.step()
        torch.cuda.synchronize()
        assert torch.equal(s2.state, full_state[:, lo:hi])
        assert torch.equal(s2.buf["other_agents_states"], full_oas[lo:hi])
        assert torch.equal(s2.buf["rewards"], full_rew[lo:hi])
        s2.close()
    # oracle on a sample, one more step from the GPU state
    ostate_all = sim.state_numpy()
    sel = rng.choice(E, 256, replace=False)
    ostate = np.ascontiguousarray(ostate_all[sel])
    oobs = oracle.step(cfg, ostate, n_agents=b["n_agents"][sel])
    sim.step()
    torch.cuda.synchronize()
    g = {k: (None if v is None else v[sel]) for k, v in bufs_numpy(sim).items() if k != "stats"}
    H.compare_step("C2 sample", b["n_agents"][sel], sim.state_numpy()[sel], g, ostate, oobs)
    sim.close()


def test_occupancy_grid_far_outside_and_edge_agents(oracle_lib, torch_mod):
    """OccupancyGridSensor for agents on the map border and far off the grid (the rotated window leaves the padded
    fast-path planes -> bounds-checked path), on both maps, against the oracle: cells bit-exact."""
    oracle = oracle_lib
    maps, rects = H.unpack_maps()
    rng = np.random.default_rng(77)
    pos = [(14.96, 14.96), (-14.99, 14.2), (14.3, -14.97), (-15.0, -15.0), (15.2, 3.0), (-16.0, -2.0), (0.0, 22.0),
           (31.0, 31.0), (-45.0, 7.0), (250.0, -90.0), (3.0e4, 2.0e4), (14.0, 0.05), (0.0, 0.0), (9.9, 9.9)]
    E, A = len(pos), 2
    cfg = gcfg(0)
    t = L.encode_type(L.POLICY_STATIC, L.DYN_UNICYCLE, L.SENSOR_OCCUPANCY_GRID)
    st = np.zeros((E, A, L.STATE_WORDS))
    for e, (x, y) in enumerate(pos):
        for a in range(A):
            st[e, a, L.PX], st[e, a, L.PY] = x + 0.37 * a, y - 0.21 * a
            st[e, a, L.HEADING] = rng.uniform(-np.pi, np.pi)
            st[e, a, L.RADIUS] = 0.4
            st[e, a, L.PREF_SPEED] = 1.0
            st[e, a, L.TYPE] = t
    st = H.f32(st)
    st[..., L.TYPE] = t
    for mid in (1, 2):
        sim = make_sim(E, A, cfg, sensors=("local_grid",))
        sim.set_obstacles(rects[mid].astype(np.float32))
        sim.bind(types=st[..., L.TYPE].astype(np.int32), n_agents=np.full((E,), A, np.int32))
        sim.load_state_numpy(st)
        sim.observe()
        torch_mod.cuda.synchronize()
        oobs = oracle.observe(cfg, st, maps=maps[mid])
        g = sim.buf["local_grid"].cpu().numpy()
        assert np.array_equal(g, oobs.local_grid), (mid, np.argwhere(g != oobs.local_grid)[:5])
        assert g.any()
        sim.close()


def _grid_pose_batch(rng, E, A, spread):
    t = L.encode_type(L.POLICY_STATIC, L.DYN_UNICYCLE, L.SENSOR_OCCUPANCY_GRID)
    st = np.zeros((E, A, L.STATE_WORDS))
    st[..., L.PX] = rng.uniform(-spread, spread, (E, A))
    st[..., L.PY] = rng.uniform(-spread, spread, (E, A))
    h = rng.uniform(-np.pi, np.pi, (E, A))
    # exact multiples of pi/2 (fx == fy == 0 everywhere: plane 0) and tiny angles (long runs of fy == 0)
    special = rng.random((E, A))
    h = np.where(special < 0.10, rng.integers(-2, 3, (E, A)) * (np.pi / 2), h)
    h = np.where((special >= 0.10) & (special < 0.15), 0.0, h)
    h = np.where((special >= 0.15) & (special < 0.20), rng.uniform(-1e-3, 1e-3, (E, A)), h)
    st[..., L.HEADING] = h
    st[..., L.RADIUS] = 0.4
    st[..., L.PREF_SPEED] = 1.0
    st = H.f32(st)
    st[..., L.TYPE] = t
    return st


@pytest.mark.parametrize("kind", ["crossing", "hallway", "noise", "lines", "checker"])
def test_occupancy_grid_shared_map_tma_kernel(oracle_lib, torch_mod, kind):
    """k_grid_tma (shared map staged through TMA; uniform / edge / other tile classes) against the oracle on maps that
    exercise every class: rectangle maps (edges + corners), per-cell noise (all 'other'), 1-cell lines and a 3-cell
    checkerboard; poses all over the map incl. the border (bounds-checked path) and exact right angles.  Cells
    bit-exact; the packed delivery equals numpy.packbits of the dense grid."""
    oracle = oracle_lib
    maps, rects = H.unpack_maps()
    rng = np.random.default_rng({"crossing": 1, "hallway": 2, "noise": 3, "lines": 4, "checker": 5}[kind])
    if kind == "crossing":
        m = maps[1]
    elif kind == "hallway":
        m = maps[2]
    elif kind == "noise":
        m = (rng.random((300, 300)) < 0.3).astype(np.uint8)
    elif kind == "lines":
        m = np.zeros((300, 300), np.uint8)
        m[::37, :] = 1
        m[:, 11::53] = 1
        m[np.arange(300), np.arange(300)] = 1
        m[0, :] = m[-1, :] = 1
        m[:, 0] = m[:, -1] = 1
    else:
        yy, xx = np.mgrid[0:300, 0:300]
        m = (((yy // 3) + (xx // 3)) % 2).astype(np.uint8)
    E, A = 192, 5
    st = _grid_pose_batch(rng, E, A, 15.4)
    cfg = gcfg(0)
    sim = make_sim(E, A, cfg, sensors=("local_grid",))
    sim.set_map(m)
    packed = sim.enable_packed_grid()
    sim.bind(types=st[..., L.TYPE].astype(np.int32), n_agents=np.full((E,), A, np.int32))
    sim.load_state_numpy(st)
    sim.observe()
    torch_mod.cuda.synchronize()
    oobs = oracle.observe(cfg, st, maps=m)
    g = sim.buf["local_grid"].cpu().numpy()
    bad = np.argwhere(g != oobs.local_grid)
    assert len(bad) == 0, (kind, len(bad), bad[:5])
    assert g.any() and not g.all()
    assert np.array_equal(packed.cpu().numpy(), np.packbits(g.reshape(E, A, -1), axis=-1))
    # a second launch re-arms its own work counters
    sim.buf["local_grid"].zero_()
    sim.run_sensor(L.SENSOR_OCCUPANCY_GRID)
    torch_mod.cuda.synchronize()
    assert np.array_equal(sim.buf["local_grid"].cpu().numpy(), g)
    sim.close()


def test_occupancy_grid_packed_output_per_env_maps(oracle_lib, torch_mod):
    """The packed delivery through the gather kernel + k_grid_pack (per-env maps take that path)."""
    maps, rects = H.unpack_maps()
    rng = np.random.default_rng(9)
    E, A = 40, 3
    st = _grid_pose_batch(rng, E, A, 12.0)
    cfg = gcfg(0)
    sim = make_sim(E, A, cfg, sensors=("local_grid",))
    sim.set_map(np.stack([maps[1 + (e % 2)] for e in range(E)]))
    packed = sim.enable_packed_grid()
    sim.bind(types=st[..., L.TYPE].astype(np.int32), n_agents=np.full((E,), A, np.int32))
    sim.load_state_numpy(st)
    sim.observe()
    torch_mod.cuda.synchronize()
    g = sim.buf["local_grid"].cpu().numpy()
    assert g.any()
    assert np.array_equal(packed.cpu().numpy(), np.packbits(g.reshape(E, A, -1), axis=-1))
    sim.close()


@pytest.mark.parametrize("A,sensors", [(10, ("other_agents_states",)), (6, ()), (20, ("other_agents_states", "laserscan", "local_grid"))])
def test_step_n_equals_n_single_steps(oracle_lib, torch_mod, A, sensors):
    """casim_step_n: n steps in one launch (state in registers, auto-reset between the steps) == n casim_step calls,
    bit for bit: the per-step rewards / done / game_over / statistics rings, the final state and the last observation."""
    torch = torch_mod
    rng = np.random.default_rng(41 + A)
    E = 333
    b = H.random_batch(rng, E, A, sensor_choices=(7,) if sensors else (0,),
                       policies=[L.POLICY_NONCOOPERATIVE] * 3 + [L.POLICY_STATIC, L.POLICY_PEDESTRIAN],
                       dynamics=[L.DYN_UNICYCLE, L.DYN_FIRST_ORDER, L.DYN_UNICYCLE_MAX_ACC], spread=3.0)
    maps, rects = H.unpack_maps()
    cfg = gcfg(1, auto_reset=1)
    sims = []
    for _ in range(2):
        sim = make_sim(E, A, cfg, sensors=sensors)
        if "laserscan" in sensors:
            sim.set_obstacles(rects[1].astype(np.float32))
        sim.set_trajectories(b["traj"])
        sim.reset(b["init"], b["types"], b["n_agents"])
        sims.append(sim)
    one, many = sims
    for K in (1, 7, 24):
        ring = many.step_n(K)
        rews, dones, gos, stats = [], [], [], []
        for _ in range(K):
            one.step()
            rews.append(one.buf["rewards"].clone())
            dones.append(one.buf["done"].clone())
            gos.append(one.buf["game_over"].clone())
            stats.append(one.buf["stats"].clone())
        torch.cuda.synchronize()
        assert torch.equal(ring["rewards"], torch.stack(rews)), K
        assert torch.equal(ring["done"], torch.stack(dones)), K
        assert torch.equal(ring["game_over"], torch.stack(gos)), K
        assert torch.equal(ring["stats"], torch.stack(stats)), K
        assert torch.equal(many.state, one.state), K
        for k in ("obs_misc", "collision_flags", "nearest_dist") + tuple(sensors):
            assert torch.equal(many.buf[k], one.buf[k]), (K, k)
        assert int(torch.stack(gos).sum()) > 0           # restarts did happen inside the launch
    for sim in sims:
        sim.close()
