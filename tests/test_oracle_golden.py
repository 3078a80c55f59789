"""CPU: the C oracle (oracle/casim_oracle.c) against golden vectors produced by the real reference
(oracle/gen_golden.py).  Flags / ordering / grid cells exact, continuous values 1e-5."""
import os

import numpy as np
import pytest

import helpers as H
from casim import layout as L


def cfg_for(oracle, mode):
    c = oracle.default_config()
    c.game_over_mode = int(mode)
    return c


def run_single_step(oracle, z, k, maps):
    n = int(z["n_agents"][k])
    cfg = cfg_for(oracle, z["game_over_mode"][k])
    state = np.ascontiguousarray(z["state_in"][k:k + 1]).copy()
    mid = int(z["map_id"][k])
    obs = oracle.step(cfg, state, ext_actions=z["ext_actions"][k:k + 1], dt=0.1,
                      n_agents=z["n_agents"][k:k + 1], maps=maps[mid], traj=z["traj"][k:k + 1])
    return n, state, obs


@pytest.mark.parametrize("name", ["core_single_step.npz", "sensors_single_step.npz", "dynamics_extra_single_step.npz",
                                  "learning_single_step.npz"])
def test_single_step_vs_reference(oracle_lib, name):
    oracle = oracle_lib
    z = H.load(name)
    maps, _ = H.unpack_maps()
    ncases = z["state_in"].shape[0]
    for k in range(ncases):
        n, state, obs = run_single_step(oracle, z, k, maps)
        what = "%s case %d" % (name, k)
        H.assert_state_close(state, z["state_out"][k:k + 1], [n], what)
        assert np.array_equal(obs.done[0, :n], z["done"][k, :n]), what
        assert obs.game_over[0] == z["game_over"][k], what
        assert np.array_equal(obs.collision_flags[0, :n], z["collision_flags"][k, :n]), what
        H.assert_close(obs.rewards[0, :n], z["rewards"][k, :n], what + " rewards")
        H.assert_close(obs.nearest_dist[0, :n], z["nearest_dist"][k, :n], what + " nearest")
        H.assert_close(obs.obs_misc[0, :n, :3], z["obs_misc"][k, :n, :3], what + " obs_misc")
        if name.startswith("learning"):      # use_ppo recorded from the reference's attr expression (config.py:209-214)
            assert np.array_equal(obs.obs_misc[0, :n, 3], z["obs_misc"][k, :n, 3]), what + " use_ppo"
        H.assert_close(obs.other_agents_states[0, :n], z["other_agents_states"][k, :n], what + " other_agents")
        # closest-N ordering: the type / radius columns pin the row order exactly
        assert np.array_equal(obs.other_agents_states[0, :n, :, 9], z["other_agents_states"][k, :n, :, 9]), what
        H.assert_close(obs.laserscan[0, :n], z["laserscan"][k, :n], what + " laserscan", rtol=0, atol=1e-12)
        assert np.array_equal(obs.local_grid[0, :n], z["local_grid"][k, :n]), what + " local_grid"


def test_reset_vs_reference(oracle_lib):
    oracle = oracle_lib
    z = H.load("reset.npz")
    maps, _ = H.unpack_maps()
    cfg = oracle.default_config()
    for k in range(z["init"].shape[0]):
        n = int(z["n_agents"][k])
        state, obs = oracle.reset(cfg, z["init"][k:k + 1], z["types"][k:k + 1], n_agents=z["n_agents"][k:k + 1],
                                  maps=maps[int(z["map_id"][k])])
        what = "reset case %d" % k
        H.assert_state_close(state, z["state_out"][k:k + 1], [n], what)
        H.assert_close(obs.other_agents_states[0, :n], z["other_agents_states"][k, :n], what)
        H.assert_close(obs.laserscan[0, :n], z["laserscan"][k, :n], what, rtol=0, atol=1e-12)
        assert np.array_equal(obs.local_grid[0, :n], z["local_grid"][k, :n]), what
        H.assert_close(obs.obs_misc[0, :n, :3], z["obs_misc"][k, :n, :3], what)


@pytest.mark.parametrize("name,tol", [("c1_two_agents.npz", 1e-9), ("c1_two_agents_nep50.npz", 1e-5)])
def test_c1_trajectory_vs_reference(oracle_lib, name, tol):
    """Config C1: 100 free-running steps of the 2-agent example.  The nep50 variant is the reference
    executed with Python-float headings under numpy 2 (float32 dynamics): it must agree
    within the 1e-5 tolerance with identical flags."""
    oracle = oracle_lib
    z = H.load(name)
    cfg = oracle.default_config()
    cfg.game_over_mode = 2
    state, obs = oracle.reset(cfg, z["init"][None], z["types"][None])
    H.assert_close(state[0][:, H.CONT_WORDS], z["states"][0][:, H.CONT_WORDS], "t0", rtol=tol, atol=tol)
    first_done = None
    for t in range(z["rewards"].shape[0]):
        obs = oracle.step(cfg, state, dt=0.1)
        ref = z["states"][t + 1]
        for k in L.INT_WORDS:
            assert np.array_equal(state[0, :, k], ref[:, k]), (t, k)
        H.assert_close(state[0][:, H.CONT_WORDS], ref[:, H.CONT_WORDS], "t=%d" % t, rtol=tol, atol=tol)
        H.assert_close(obs.rewards[0], z["rewards"][t], "rew t=%d" % t, rtol=tol, atol=tol)
        assert np.array_equal(obs.done[0], z["done"][t])
        assert obs.game_over[0] == z["game_over"][t]
        H.assert_close(obs.other_agents_states[0], z["other_agents_states"][t + 1], "oas t=%d" % t, rtol=tol, atol=tol)
        if first_done is None and obs.done[0].any():
            first_done = t
    assert first_done == 37            # first done at step index 37


def test_appendix_b_known_answers(oracle_lib):
    """SURVEY.md Appendix B values (reference run with float headings -> 1e-5 tolerance on positions)."""
    oracle = oracle_lib
    cfg = oracle.default_config()
    cfg.game_over_mode = 2
    t = L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS)
    init = np.array([[[-3, -3, 3, 3, 0.5, 1.0, 0.5], [3, 3, -3, -3, 0.5, 1.0, 0.5]]], dtype=np.float64)
    state, _ = oracle.reset(cfg, init, np.array([[t, t]], np.int32))
    assert abs(state[0, 0, L.TIME_REMAINING] / 3.0 - 7.7352813742385695) < 1e-12
    obs = oracle.step(cfg, state)
    assert abs(state[0, 0, L.PX] - (-2.9292893186211586)) < 1e-7
    assert abs(obs.rewards[0, 0] - (-0.00076923076923077)) < 1e-15
    assert abs(state[0, 0, L.DIST_TO_GOAL] - 8.38528136962797) < 1e-7
    row = obs.other_agents_states[0, 0, 0]
    np.testing.assert_allclose(row, [5.8585786372423172, 5.8585786372423172, 8.2852813650173704, 0,
                                     -0.999999982885729, 0, 0.5, 1.0, 7.2852813650173704, 2.0], atol=1e-7)
    # ordering check: host at the origin, others at (3,0), (0,2) Static, (0,-2), (-1,0), all r = 0.2
    ts = L.encode_type(L.POLICY_STATIC, L.DYN_UNICYCLE, 0)
    tn = L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS)
    init = np.array([[[0, 0, 5, 0, 0.2, 1, 0], [3, 0, 5, 5, 0.2, 1, 0], [0, 2, 5, 5, 0.2, 1, 0],
                      [0, -2, 5, 5, 0.2, 1, 0], [-1, 0, 5, 5, 0.2, 1, 0]]], dtype=np.float64)
    state, obs = oracle.reset(cfg, init, np.array([[tn, tn, ts, tn, tn]], np.int32))
    rows = obs.other_agents_states[0, 0]
    np.testing.assert_allclose(rows[:4, 8], [2.6, 1.6, 1.6, 0.6], atol=1e-12)
    assert list(rows[:4, 9]) == [2, 2, 1, 2]          # agents 1, 3, 2, 4: the tie is won by the higher index
    np.testing.assert_allclose(rows[:4, 0:2], [[3, 0], [0, -2], [0, 2], [-1, 0]], atol=1e-12)


def test_map_raster_vs_reference(oracle_lib):
    oracle = oracle_lib
    maps, rects = H.unpack_maps()
    cfg = oracle.default_config()
    for mid in (1, 2):
        m = oracle.rasterize(cfg, rects[mid])
        assert np.array_equal(m, maps[mid])


def test_occupancy_grid_vs_cv2(oracle_lib):
    """OccupancyGridSensor restatement vs the reference sensor (cv2.warpAffine) on 400 random poses."""
    oracle = oracle_lib
    z = H.load("occgrid.npz")
    maps, _ = H.unpack_maps()
    cfg = oracle.default_config()
    grids = np.unpackbits(z["grids"], axis=-1)[..., :60]
    t = L.encode_type(L.POLICY_STATIC, L.DYN_UNICYCLE, L.SENSOR_OCCUPANCY_GRID)
    mism = 0
    for k in range(z["poses"].shape[0]):
        st = np.zeros((1, 1, L.STATE_WORDS))
        st[0, 0, L.PX], st[0, 0, L.PY], st[0, 0, L.HEADING] = z["poses"][k]
        st[0, 0, L.TYPE] = t
        st[0, 0, L.RADIUS] = 0.5
        obs = oracle.observe(cfg, st, maps=maps[int(z["map_id"][k])])
        mism += int((obs.local_grid[0, 0] != grids[k]).sum())
    assert mism == 0


def test_edt_vs_scipy_reference(oracle_lib):
    """edfMap.update (scipy.ndimage.distance_transform_edt) == sqrt(exact squared EDT) * 0.1, three maps."""
    z = H.load("explore.npz")
    maps, d2 = H.explore_maps(z)
    assert len(maps) == 3
    cfg = oracle_lib.default_config()
    for k, m in enumerate(maps):
        assert np.array_equal(oracle_lib.rasterize(cfg, z["rects_%d" % k]), m)
        assert np.array_equal(oracle_lib.edt(m), d2[k])


def test_explore_step_vs_reference(oracle_lib):
    """ig_greedy.find_next_action (detector emulation, targetMap.update, greedy one-step MI argmax) against the
    reference classes: visible cells / occupied-empty choice / feasibility / action exact, MI to 1e-9."""
    oracle = oracle_lib
    z = H.load("explore.npz")
    maps, _ = H.explore_maps(z)
    cfg = oracle.default_config()
    ncases = len(z["map_id"])
    assert ncases >= 40 and (z["n_targets"] > 0).sum() >= 10
    for mid in range(len(maps)):
        cases = [k for k in range(ncases) if int(z["map_id"][k]) == mid]
        b = H.explore_case_batch(z, cases)
        belief = b["belief0"].copy()
        out = oracle.explore_step(cfg, b["state"], b["oas"], belief, maps[mid])
        H.check_explore_vs_golden("oracle map %d" % mid, z, cases, cfg, belief, out, rtol=1e-9)
        assert int(out["counters"][3]) == 0          # nothing clamped: all samples stayed inside the arrays


def test_explore_known_answers(oracle_lib):
    """crossing map, pose (-5, 0, 0): 52 visible cells, reward 1.0740499287868415 on the prior."""
    oracle = oracle_lib
    z = H.load("explore.npz")
    maps, d2 = H.explore_maps(z)
    cfg = oracle.default_config()
    assert abs(np.sqrt(d2[0].max()) * 0.1 - 7.0710678118654755) < 1e-12
    pose = np.array([-5.0, 0.0, 0.0])
    vis = oracle.visible_cells(cfg, maps[0], pose)
    assert int(vis.sum()) == 52 == int(z["kb_num_visible"])
    jj, ii = np.nonzero(vis)
    cells = sorted(zip(ii.tolist(), jj.tolist()))
    assert cells[:6] == [(21, 29), (21, 30), (22, 29), (22, 30), (23, 28), (23, 29)]
    r = oracle.reward_from_pose(cfg, maps[0], pose, np.ones((60, 60)))
    assert abs(r - 1.0740499287868415) < 1e-12 and abs(r - float(z["kb_reward_prior"])) < 1e-12


def test_explore_rollouts_vs_reference(oracle_lib):
    """Dec-MCTS simulation rollouts (ig_mcts.get_next_pose / mcts_sim_state_storer / mcts_reward driven by
    pydecmcts Tree._simulate): observed-cell sets and feasibility exact, end pose to 1e-12, reward to 1e-9."""
    oracle = oracle_lib
    z = H.load("rollouts.npz")
    maps, cases = H.rollout_cases(z)
    cfg = oracle.default_config()
    assert len(cases) >= 20 and int(z["n_infeasible"].sum()) >= 50
    for n, cs in enumerate(cases):
        out = oracle.explore_rollouts(cfg, maps[cs["map_id"]], cs["pose"], cs["radius"], cs["belief"].astype(np.float64),
                                      cs["actions"], exclude=cs["exclude"] if cs["exclude"].any() else None)
        assert np.array_equal(out["obsvd"].astype(bool), cs["obsvd"]), "case %d observed cells" % n
        assert np.abs(out["end_pose"] - cs["end_pose"]).max() < 1e-12, "case %d end pose" % n
        H.assert_close(out["rewards"], cs["rewards"], "case %d rewards" % n, rtol=1e-9, atol=1e-12)
        lens = (cs["actions"] != 255).sum(1)
        assert int((lens - out["n_feasible"]).sum()) == int(z["n_infeasible"][n]), "case %d infeasible steps" % n


def test_team_update_vs_reference(oracle_lib):
    """ig_mcts.update_belief over 3 steps for teams of 3 explorers (sequencing of obsvd_targets: this step's for
    lower-index agents, the previous step's for the others and for done agents): belief to 1e-12, team cells exact."""
    oracle = oracle_lib
    z = H.load("team.npz")
    cfg = oracle.default_config()
    smap = np.unpackbits(z["static"])[:300 * 300].reshape(300, 300).astype(np.uint8)
    n_ig = int(z["n_ig"])
    assert (z["n_targets"] > 0).sum() >= 10 and z["done"].any()
    for sc in range(z["pose"].shape[0]):
        steps = H.team_scene(z, sc)
        A = steps[0]["state"].shape[1]
        belief = np.repeat(z["belief0"][sc].astype(np.float64)[None, None], A, axis=1).copy()
        tp, npv = np.zeros((1, A, 32, 2)), np.zeros((1, A), np.int32)
        for st, s in enumerate(steps):
            o = oracle.team_update(cfg, s["state"], s["oas"], belief, smap, tp, npv)
            tp, npv = o["targets"], o["n_targets"]
            for i in range(n_ig):
                H.assert_close(belief[0, i], s["belief_after"][i], "scene %d step %d agent %d belief" % (sc, st, i),
                               rtol=1e-12, atol=0)
                if s["done"][i]:
                    continue
                assert np.array_equal(o["team_mask"][0, i].astype(bool), s["team_mask"][i]), (sc, st, i)
                H.assert_close(o["team_reward"][0, i], s["team_reward"][i], "team reward", rtol=1e-9, atol=1e-12)
                assert int(npv[0, i]) == int(s["n_targets"][i])


def test_scenario_generators_vs_reference(oracle_lib):
    """train_agents_swap_circle / train_agents_random_positions(seed=s): the reference's own scenarios (numpy's and
    CPython's MT19937 streams restated): positions to 1e-12, policies / dynamics / sensors / counts exact."""
    oracle = oracle_lib
    z = H.load("scenarios.npz")
    A = z["init"].shape[1]
    assert len(z["kind"]) >= 120 and set(z["kind"].tolist()) == {0, 1, 2}
    for k in range(len(z["kind"])):
        o = oracle.scenario(int(z["kind"][k]), 1, A, int(z["number_of_agents"][k]), z["seed"][k:k + 1],
                            L.POLICY_LEARNING, L.DYN_FIRST_ORDER, L.DYN_UNICYCLE)
        n = int(z["n_agents"][k])
        assert int(o["n_agents"][0]) == n, k
        assert np.array_equal(o["types"][0], z["types"][k]), k
        H.assert_close(o["init"][0, :n, :6], z["init"][k, :n, :6], "scenario %d init" % k, rtol=1e-12, atol=1e-12)
        assert np.isnan(o["init"][0, :n, 6]).all() and not o["init"][0, n:].any()
        assert np.array_equal(o["coop"][0], z["coop"][k]), k


def test_ga3c_state_vector_vs_reference(oracle_lib):
    """agents_to_ga3c_cadrl_state (GA3CCADRLPolicy.py:45-106): slot ORDER exact (Python round(x, 2), stable sort on
    (-round(dist, 2), p_orthog), clip to the closest N), values to 1e-12."""
    oracle = oracle_lib
    z = H.load("ga3c_state.npz")
    w = oracle.ga3c_random_weights(0)
    ncases = z["state"].shape[0]
    n_ties = 0
    for k in range(ncases):
        cfg = oracle.default_config()
        cfg.max_other_agents_observed = int(z["max_obs"][k])
        n = int(z["n_agents"][k])
        out = oracle.ga3c_forward(cfg, z["state"][k:k + 1], w, n_agents=z["n_agents"][k:k + 1])
        ref = z["nn_input"][k, :n]
        got = out["nn_input"][0, :n]
        assert np.array_equal(got[:, 0], ref[:, 0]), k          # number of other agents observed
        # slot order: the (radius, combined radius) columns identify which agent sits in which slot
        np.testing.assert_allclose(got, ref, rtol=0, atol=1e-12, err_msg="case %d" % k)
        d = np.round(ref[:, 5:].reshape(n, 10, 7)[:, :, 6], 2)
        n_ties += int(sum((np.diff(np.sort(row[:int(c)])) == 0).sum() for row, c in zip(d, ref[:, 0])))
    assert n_ties > 50            # the fixtures do exercise the tie-break


def test_ga3c_network_vs_torch(oracle_lib):
    """Network restatement (TF1 LSTMCell gate order i, j, f, o + forget_bias 1, masked sequence, 3 x Dense(256) ReLU,
    Dense(11), softmax floor) against an independent torch implementation.  PARITY UNPINNED w.r.t. TensorFlow."""
    import torch
    oracle = oracle_lib
    w = oracle.ga3c_random_weights(3)
    rng = np.random.default_rng(5)
    B = 64
    x = np.zeros((B, 75), np.float32)
    x[:, 0] = rng.integers(0, 11, B)
    x[:, 1:5] = rng.uniform(0, 5, (B, 4))
    for b in range(B):
        L_ = int(x[b, 0])
        x[b, 5:5 + 7 * L_] = rng.uniform(-3, 3, 7 * L_)
    probs = oracle.ga3c_network(w, x)
    # torch version
    parts, off = {}, 0
    for name, shape in oracle.GA3C_LAYOUT:
        size = int(np.prod(shape))
        parts[name] = torch.from_numpy(w[off:off + size].reshape(shape).copy())
        off += size
    K, bvec = parts["rnn/lstm_cell/kernel"], parts["rnn/lstm_cell/bias"]
    # torch LSTMCell gate order is i, f, g, o; TF1 LSTMCell is i, j(=g), f, o with forget_bias added to f
    def reorder(t):
        i, j, f, o = t.split(64, dim=-1)
        return torch.cat([i, f, j, o], dim=-1)
    cell = torch.nn.LSTMCell(7, 64)
    with torch.no_grad():
        cell.weight_ih.copy_(reorder(K[:7]).T)
        cell.weight_hh.copy_(reorder(K[7:]).T)
        bb = reorder(bvec).clone()
        bb[64:128] += 1.0
        cell.bias_ih.copy_(bb)
        cell.bias_hh.zero_()
        xt = torch.from_numpy(x)
        avg = torch.tensor([0.0] + [0, 0, 1, 0.5] + [0, 0, 0, 0, 0.5, 0, 1] * 10)
        std = torch.tensor([1.0] + [5, 3.14, 1, 1] + [5, 5, 1, 1, 1, 5, 1] * 10)
        xn = (xt - avg) / std
        h = torch.zeros(B, 64)
        c = torch.zeros(B, 64)
        seq = xn[:, 5:].reshape(B, 10, 7)
        for t in range(10):
            h2, c2 = cell(seq[:, t], (h, c))
            m = (xt[:, 0] > t).unsqueeze(1)
            h, c = torch.where(m, h2, h), torch.where(m, c2, c)
        a = torch.cat([xn[:, 1:5], h], 1)
        for name in ("layer1", "layer2", "fullyconnected1"):
            a = torch.relu(a @ parts[name + "/kernel"] + parts[name + "/bias"])
        logits = a @ parts["logits_p/kernel"] + parts["logits_p/bias"]
        ref = (torch.softmax(logits, 1) + 1e-4) / (1 + 11e-4)
    np.testing.assert_allclose(probs, ref.numpy(), rtol=2e-4, atol=2e-6)
    assert np.array_equal(probs.argmax(1), ref.numpy().argmax(1))
    np.testing.assert_allclose(oracle.ga3c_actions()[[0, 4, 6, 10]],
                               [[1.0, -np.pi / 6], [1.0, np.pi / 6], [0.5, 0.0], [0.0, np.pi / 6]], atol=1e-15)


def test_ga3c_network_vs_tf_graph_golden(oracle_lib):
    """a24 pin: the C restatement against tests/golden/ga3c_probs.npz -- the TF graph of GA3C_CADRL/network.py:45-98
    evaluated in numpy float64 (oracle/gen_golden_ga3c_probs.py, written independently of the C oracle) on the
    reference's shipped IROS18 weights and 1574 state vectors produced by the reference's own
    agents_to_ga3c_cadrl_state.  The C oracle computes in float32 like TF: 5e-6 absolute on probabilities (the float32
    evaluation of the same graph differs from float64 by 3e-6), argmax identical on every vector."""
    oracle = oracle_lib
    z = H.load("ga3c_probs.npz")
    w = H.load("ga3c_iros18.npz")["weights"]
    assert len(z["x"]) >= 200
    p = oracle.ga3c_network(w, z["x"])
    assert np.abs(p - z["probs64"]).max() < 5e-6
    assert np.array_equal(p.argmax(1), z["action"])
    assert np.abs(z["probs32"] - z["probs64"]).max() < 5e-6


# ---------------------------------------------------------------------------------------------
# RVO policy
def _rvo_sim_call(fn, g, i):
    import ctypes
    P = ctypes.POINTER(ctypes.c_double)
    n = int(g["n"][i])
    arrs = [np.ascontiguousarray(g[k][i, :n], dtype=np.float64) for k in ("pos", "vel", "pref", "radius", "ms")]
    out = np.zeros(2)
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_int, P, P, P, P, P, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_double,
                   ctypes.c_int, ctypes.c_double, P]
    assert fn(n, *[a.ctypes.data_as(P) for a in arrs], int(g["ego"][i]), float(g["collab"][i]), 0.1, float(g["nd"][i]),
              int(g["mn"][i]), 5.0, out.ctypes.data_as(P)) == 0
    return out


def test_rvo_simulator_restatement_is_bit_exact_vs_reference_library(oracle_lib):
    """oracle/casim_oracle_rvo.c vs the new ego positions recorded from the reference's own C++ RVO2 library
    (oracle/gen_golden_rvo.py, 600 random crowds incl. collisions, negative / zero collaboration, neighbour caps)."""
    g = np.load(os.path.join(H.GOLDEN, "rvo.npz"))
    fn = oracle_lib.lib().casim_oracle_rvo_ref_step
    for i in range(len(g["n"])):
        got = _rvo_sim_call(fn, g, i)
        assert np.array_equal(got, g["new_pos"][i]), "case %d: %r != %r" % (i, got, g["new_pos"][i])


def test_rvo_reference_library_if_built(oracle_lib):
    """When oracle/_ref/librvo2_ref.so (the reference's sources compiled by `make -C oracle ref`) travelled with the
    repo, compare call for call on fresh random inputs, not only on the fixtures."""
    import ctypes
    path = os.path.join(os.path.dirname(H.GOLDEN), os.pardir, "oracle", "_ref", "librvo2_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/librvo2_ref.so not built")
    ref = ctypes.CDLL(path).rvo2_ref_step
    fn = oracle_lib.lib().casim_oracle_rvo_ref_step
    rng = np.random.default_rng(2024)
    for c in range(300):
        n = int(rng.integers(2, 25))
        sp = float(rng.choice([1.0, 2.5, 6.0]))
        g = {"n": [n], "ego": [int(rng.integers(0, n))], "collab": [float(rng.choice([1.0, 0.5, -1.0]))], "nd": [np.inf],
             "mn": [int(rng.choice([10, 4, 30]))]}
        f = lambda a: np.asarray(a, np.float32).astype(np.float64)[None]
        g["pos"], g["vel"] = f(rng.uniform(-sp, sp, (n, 2))), f(rng.uniform(-1, 1, (n, 2)))
        g["pref"], g["radius"], g["ms"] = f(rng.uniform(-1, 1, (n, 2))), f(rng.uniform(0.2, 0.9, n)), f(rng.uniform(0.5, 1.5, n))
        assert np.array_equal(_rvo_sim_call(ref, g, 0), _rvo_sim_call(fn, g, 0)), "case %d" % c


def test_rvo_policy_restatement_vs_reference_policy(oracle_lib):
    """casim_oracle_rvo_actions vs the UNMODIFIED RVOPolicy.find_next_action run over the reference's Cython binding
    (tests/golden/rvo.npz keys state / coop / n_agents / actions)."""
    from casim import lib
    g = np.load(os.path.join(H.GOLDEN, "rvo.npz"))
    act = oracle_lib.rvo_actions(lib.default_config(), g["state"], g["coop"], g["n_agents"])
    np.testing.assert_allclose(act, g["actions"], rtol=0, atol=1e-12)
    assert ((act[..., 0] == 0) == (g["actions"][..., 0] == 0)).all()      # the pi/6 turn limit zeroes the same speeds


def test_rvo_policy_with_obstacles_vs_reference_policy(oracle_lib):
    """Obstacle ORCA lines (Agent.cpp:66-284) + obstacle tree (KdTree.cpp:113-329): actions recorded from the unmodified
    RVOPolicy.find_next_action called with the crossing-map obstacles."""
    from casim import lib
    g = np.load(os.path.join(H.GOLDEN, "rvo.npz"))
    act = oracle_lib.rvo_actions(lib.default_config(), g["obst_state"], g["obst_coop"], g["obst_n_agents"],
                                 polygons=list(g["obst_polygons"]))
    np.testing.assert_allclose(act, g["obst_actions"], rtol=0, atol=1e-12)
    free = oracle_lib.rvo_actions(lib.default_config(), g["obst_state"], g["obst_coop"], g["obst_n_agents"])
    assert (np.abs(free - act).max(-1) > 1e-6).sum() > 50            # the obstacles do constrain many of the agents


def test_rvo_obstacle_simulator_vs_reference_library_if_built(oracle_lib):
    """Random star-shaped (non-convex) polygons, 2-vertex walls and the crossing map, incl. edge splitting in the
    obstacle tree: bit-exact new positions against oracle/_ref/librvo2_ref.so when it travelled with the repo."""
    import ctypes
    path = os.path.join(os.path.dirname(H.GOLDEN), os.pardir, "oracle", "_ref", "librvo2_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/librvo2_ref.so not built")
    ref = ctypes.CDLL(path).rvo2_ref_step_obst
    orc = oracle_lib.lib().casim_oracle_rvo_ref_step_obst
    P, PI = ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int32)
    for fn in (ref, orc):
        fn.restype = ctypes.c_int
        fn.argtypes = [ctypes.c_int, P, P, P, P, P, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_int,
                       ctypes.c_double, ctypes.c_int, PI, P, P]

    def call(fn, arrs, ego, collab, polys):
        out = np.zeros(2)
        sizes = np.array([len(p) for p in polys], np.int32)
        verts = np.ascontiguousarray(np.concatenate(polys), np.float64)
        assert fn(len(arrs[3]), *[a.ctypes.data_as(P) for a in arrs], ego, collab, 0.1, float("inf"), 10, 5.0, len(polys),
                  sizes.ctypes.data_as(PI), verts.ctypes.data_as(P), out.ctypes.data_as(P)) == 0
        return out
    rng = np.random.default_rng(5)
    f = lambda a: np.ascontiguousarray(np.asarray(a, np.float32).astype(np.float64))   # noqa: E731
    for c in range(400):
        n = int(rng.integers(1, 8))
        if c % 3 == 0:
            polys = [np.array(p, float) for p in g_cross()]
        elif c % 3 == 1:
            polys = []
            for k in range(int(rng.integers(1, 5))):
                m = int(rng.integers(3, 7))
                ang, rad = np.sort(rng.uniform(0, 2 * np.pi, m)), rng.uniform(0.5, 2.5, m)
                polys.append(rng.uniform(-6, 6, 2) + np.stack([rad * np.cos(ang), rad * np.sin(ang)], 1))
        else:
            polys = [np.array([[x, y], [x + w, y + h]]) for x, y, w, h in rng.uniform(-5, 5, (3, 4))]
        pos, vel, sp = f(rng.uniform(-4, 4, (n, 2))), f(rng.uniform(-1, 1, (n, 2))), f(rng.uniform(0.5, 1.5, n))
        arrs = [pos, vel, f(rng.uniform(-1, 1, (n, 2))), f(rng.uniform(0.25, 0.9, n)), sp]
        ego, collab = int(rng.integers(0, n)), float(rng.choice([1.0, 0.5, 0.0, -0.5]))
        assert np.array_equal(call(ref, arrs, ego, collab, polys), call(orc, arrs, ego, collab, polys)), "case %d" % c


def g_cross():
    return [[(10, 10), (2, 10), (2, 2), (10, 2)], [(-2, 10), (-10, 10), (-10, 2), (-2, 2)],
            [(10, -2), (2, -2), (2, -10), (10, -10)], [(-2, -2), (-10, -2), (-10, -10), (-2, -10)]]
