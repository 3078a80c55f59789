"""GPU: the gym-shaped boundary (casim.env.BatchedCollisionAvoidanceEnv + casim.plugins) -- the reference's
example scenario through reset()/step() with Agent objects, checked against SURVEY.md Appendix B known answers
(recorded from the reference) and against the oracle in lock-step."""
import numpy as np
import pytest

import helpers as H
from casim import layout as L
from casim import plugins as P
from casim.config import Config
from casim.plugins import Agent

pytestmark = pytest.mark.gpu


class _Cfg(Config):
    TRAIN_SINGLE_AGENT = False
    STATES_IN_OBS = ['dist_to_goal', 'rel_goal', 'radius', 'heading_ego_frame', 'pref_speed', 'other_agents_states']


def test_c1_example_through_env_api(oracle_lib):
    """BASELINE.json configs[0]: example.py 2-agent scenario, NonCooperativePolicy + UnicycleDynamics +
    OtherAgentsStatesSensor, one env (envs/test_cases.py:79-86)."""
    import torch
    from casim.env import BatchedCollisionAvoidanceEnv
    agents = [Agent(-3, -3, 3, 3, 0.5, 1.0, 0.5, P.NonCooperativePolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 0),
              Agent(3, 3, -3, -3, 0.5, 1.0, 0.5, P.NonCooperativePolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 1)]
    env = BatchedCollisionAvoidanceEnv(num_envs=1, max_agents=2, config=_Cfg)
    env.set_agents(agents)
    obs = env.reset()
    assert set(obs.keys()) == set(_Cfg.STATES_IN_OBS)
    cfg = env.cfg
    first_done = None
    for t in range(60):
        ostate = env.sim.state_numpy()
        oobs = oracle_lib.step(cfg, ostate)
        obs, rewards, game_over, info = env.step({})
        torch.cuda.synchronize()
        assert rewards.shape == (2,) and info["which_agents_done"].shape == (2,)
        H.assert_close(rewards.cpu().numpy(), oobs.rewards[0], "rewards t=%d" % t)
        assert np.array_equal(info["which_agents_done"].cpu().numpy(), oobs.done[0].astype(bool))
        H.assert_close(obs[0]["other_agents_states"], oobs.other_agents_states[0, 0], "oas t=%d" % t)
        H.assert_close(obs["dist_to_goal"].cpu().numpy()[0], ostate[0, :, L.DIST_TO_GOAL], "dist_to_goal t=%d" % t)
        if t == 0:   # SURVEY.md Appendix B (reference run)
            assert abs(agents[0].pos_global_frame[0] - (-2.9292893186211586)) < 1e-5
            assert abs(float(rewards[0]) - (-0.00076923076923077)) < 1e-8
            assert abs(float(obs[0]["dist_to_goal"]) - 8.38528136962797) < 1e-5
            np.testing.assert_allclose(obs[0]["rel_goal"], [5.929289318621159, 5.929289318621159], rtol=1e-6)
            np.testing.assert_allclose(obs[0]["other_agents_states"][0],
                                       [5.8585786372423172, 5.8585786372423172, 8.2852813650173704, 0, -0.999999982885729, 0,
                                        0.5, 1.0, 7.2852813650173704, 2.0], atol=1e-5)
        if first_done is None and bool(info["which_agents_done"].any()):
            first_done = t
            assert agents[0].in_collision and agents[1].in_collision
            assert abs(float(rewards[0]) - (-0.7692307692307693)) < 1e-6
    assert first_done == 37
    env.close()


def test_env_runs_exploration_and_ga3c_policies(oracle_lib):
    """ig_greedy and GA3C-CADRL agents are driven by the env itself (casim_explore_step / casim_ga3c_forward inside
    step()), batched over several environments, with the reference's obstacle format."""
    import torch
    from casim.env import BatchedCollisionAvoidanceEnv
    E = 8
    envs_agents = []
    rng = np.random.default_rng(0)
    for e in range(E):
        x = float(rng.uniform(-9, -4))
        envs_agents.append([
            Agent(x, 0.3, 9, 0, 0.5, 1.0, 0.0, P.ig_greedy, P.FirstOrderDynamics, [P.OtherAgentsStatesSensor], 0),
            Agent(x + 2.5, -0.2, 0, 0, 0.2, 1.0, 0.0, P.StaticPolicy, P.UnicycleDynamics, [], 1),
            Agent(0.5, -8, 0.5, 8, 0.4, 1.0, None, P.GA3CCADRLPolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 2)])
    crossing = [[(10, 10), (2, 10), (2, 2), (10, 2)], [(-2, 10), (-10, 10), (-10, 2), (-2, 2)],
                [(10, -2), (2, -2), (2, -10), (10, -10)], [(-2, -2), (-10, -2), (-10, -10), (-2, -10)]]
    env = BatchedCollisionAvoidanceEnv(num_envs=E, max_agents=3, config=_Cfg)
    env.set_agents(envs_agents)
    env.set_static_map(crossing)
    env.set_ga3c_weights(oracle_lib.ga3c_random_weights(2))
    env.reset()
    p0 = env.sim.state[0, :, :, 0:2].clone()
    for t in range(10):
        obs, rewards, game_over, info = env.step()
    torch.cuda.synchronize()
    assert rewards.shape == (E, 3) and game_over.shape == (E,)
    moved = (env.sim.state[0, :, :, 0:2] - p0).norm(dim=-1).cpu().numpy()
    assert (moved[:, 0] > 0.1).all() and (moved[:, 1] == 0).all() and (moved[:, 2] > 0.05).all()
    belief = env.sim.belief.cpu().numpy()
    assert (belief[:, 0] > 1.0).any() and (belief[:, 0] < 1.0).any()      # occupied and empty updates happened
    assert (belief[:, 1:] == 1.0).all()                                   # only the ig_greedy agent owns a belief
    assert obs[3][0]["other_agents_states"].shape == (9, 10)
    env.close()


@pytest.mark.parametrize("pinned", [True, False])
def test_step_host_delivery_matches_device_step(pinned):
    """casim_step_host: results delivered into host memory (zero-copy stores for pinned scalar buffers, staged
    copies for pageable ones and for sensor planes) equal, bit for bit, what casim_step leaves in device memory."""
    import torch
    from casim import lib, scenarios as S
    from casim.sim import BatchedSim
    from casim.config import NUM_STATS
    E, A = 96, 10
    cfg = lib.default_config()
    cfg.auto_reset = 1
    sensors = ("other_agents_states", "laserscan")
    init, types, n, rects = S.crossing(E, A, seed=7, sensors=L.SENSOR_OTHER_AGENTS | L.SENSOR_LASERSCAN)
    types = np.array(types).copy()
    types[:, 1] = L.encode_type(L.POLICY_EXTERNAL, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS | L.SENSOR_LASERSCAN)
    types[:, 2] = L.encode_type(L.POLICY_LEARNING, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS | L.SENSOR_LASERSCAN)
    sims = []
    for _ in range(2):
        sim = BatchedSim(E, A, cfg=cfg, sensors=sensors, want_debug=True, want_misc=True, want_stats=True)
        sim.set_obstacles(rects)
        sim.reset(init, types, n)
        sims.append(sim)
    ref, tst = sims

    def hostbuf(shape, dtype):
        t = torch.full(shape, 77, dtype=dtype)
        return t.pin_memory() if pinned else t
    host = {"rewards": hostbuf((E, A), torch.float32), "done": hostbuf((E, A), torch.uint8),
            "game_over": hostbuf((E,), torch.uint8), "stats": hostbuf((NUM_STATS,), torch.float32),
            "collision_flags": hostbuf((E, A), torch.uint8), "nearest_dist": hostbuf((E, A), torch.float32),
            "obs_misc": hostbuf((E, A, 4), torch.float32),
            "other_agents_states": hostbuf(tuple(ref.buf["other_agents_states"].shape), torch.float32),
            "laserscan": hostbuf(tuple(ref.buf["laserscan"].shape), torch.float32)}
    gen = torch.Generator().manual_seed(3)
    act_host = torch.zeros((E, A, 2), dtype=torch.float32)
    act_host = act_host.pin_memory() if pinned else act_host
    act_dev = torch.zeros((E, A, 2), dtype=torch.float32, device=ref.device)
    bound = None
    for t in range(25):
        act_host.copy_(torch.rand((E, A, 2), generator=gen))            # the caller's actions change every step
        ref.step(ext_actions=act_host.to(ref.device))
        torch.cuda.synchronize()
        if pinned:                  # the pre-bound form (arguments resolved once) and the plain call alternate
            bound = bound if t else tst.bind_host_step(host, ext_actions_host=act_host, ext_actions_dev=act_dev)
            (bound if t % 2 == 0 else lambda: tst.step_host(host, ext_actions_host=act_host, ext_actions_dev=act_dev))()
        else:
            tst.step_host(host, ext_actions_host=act_host, ext_actions_dev=act_dev)
        assert torch.equal(act_dev.cpu(), act_host), "device copy of the actions not refreshed"
        for name, hb in host.items():
            want = ref.buf[name].cpu().numpy()
            got = hb.numpy()
            assert np.array_equal(want, got, equal_nan=True), "%s differs at step %d (pinned=%s)" % (name, t, pinned)
        assert torch.equal(ref.state, tst.state)
    assert host["done"].any() and host["stats"][6] == E * A
    for s in sims:
        s.close()


def test_pipelined_host_step_equals_blocking_steps():
    """casim_step_host_pipelined: the host buffers handed back by call t hold step t-1, bit for bit what a plain
    casim_step sequence on a twin simulator produces; the packed grid equals numpy.packbits of the dense one."""
    import torch
    from casim import lib, scenarios as S
    from casim.sim import BatchedSim
    E, A = 256, 10
    sens = ("other_agents_states", "laserscan", "local_grid")
    init, types, n, rects = S.crossing(E, A, seed=11, sensors=L.SENSOR_OTHER_AGENTS | L.SENSOR_LASERSCAN | L.SENSOR_OCCUPANCY_GRID)
    cfg = lib.default_config()
    cfg.auto_reset = 1
    sims = []
    for _ in range(2):
        sim = BatchedSim(E, A, cfg=cfg, sensors=sens, want_debug=False)
        sim.set_obstacles(rects)
        sim.reset(init, types, n)
        sims.append(sim)
    a, b = sims
    pipe = a.host_pipeline(["rewards", "done", "game_over", "obs_misc", "other_agents_states", "laserscan", "local_grid"])
    assert pipe.d2h_bytes < 1000 * E * A
    T = 7
    refs = []
    for t in range(T):
        got = pipe.step()
        b.step()
        torch.cuda.synchronize()
        refs.append({k: v.cpu().clone() for k, v in b.buf.items() if isinstance(v, torch.Tensor)})
        if t == 0:
            assert got is None
            continue
        want = refs[t - 1]
        for k in ("rewards", "done", "game_over", "obs_misc", "other_agents_states", "laserscan"):
            assert torch.equal(got[k], want[k]), (t, k)
        assert np.array_equal(got["local_grid_packed"].numpy(), np.packbits(want["local_grid"].numpy().reshape(E, A, -1), axis=-1)), t
    got = pipe.flush()
    assert torch.equal(got["other_agents_states"], refs[-1]["other_agents_states"])
    assert np.array_equal(got["local_grid_packed"].numpy(), np.packbits(refs[-1]["local_grid"].numpy().reshape(E, A, -1), axis=-1))
    for sim in sims:
        sim.close()


def test_reset_without_a_scenario_starts_the_forks_default():
    """Config.SCENARIOS_FOR_TRAINING = ["IG_agent_crossing"] (envs/config.py:91 of the fork): reset() without set_agents /
    set_scenario_tensors / set_scenario_generator builds that fixed scenario (test_cases.py:3209-3237) in every
    environment, map included -- what _init_agents does (collision_avoidance_env.py:403-445)."""
    import torch
    from casim import test_cases as tc
    from casim.env import BatchedCollisionAvoidanceEnv
    env = BatchedCollisionAvoidanceEnv(num_envs=3, max_agents=6)
    env.reset()
    torch.cuda.synchronize()
    st = env.sim.state_numpy()
    assert np.array_equal(env.sim._keep["n_agents"].cpu().numpy(), [5, 5, 5])
    assert np.allclose(st[:, :5, L.PX], [[-5, 0, 5, 6, -6]] * 3) and np.allclose(st[:, :5, L.PY], [[0, 0, 0, 12, -12]] * 3)
    pol = st[:, :5, L.TYPE].astype(np.int64) & 0xff
    assert (pol[:, :3] == L.POLICY_IG_GREEDY).all() and (pol[:, 3:] == L.POLICY_STATIC).all()
    assert env.sim.has_map and env._mcts is not None                 # crossing map set, Dec-MCTS planner armed
    agents, obstacles = tc.IG_agent_crossing()
    assert len(agents) == 5 and obstacles == tc.CROSSING_OBSTACLES
    env.close()
    small = BatchedCollisionAvoidanceEnv(num_envs=1, max_agents=4)   # the scenario does not fit: said so, loudly
    with pytest.raises(RuntimeError, match="max_agents"):
        small.reset()
    small.close()
    assert [a.policy.policy_id for a in tc.get_testcase_two_agents()] == [L.POLICY_LEARNING, L.POLICY_GA3C_CADRL]
    assert tc.get_testcase_two_agents_external_rvo()[1].policy.policy_id == L.POLICY_RVO
    assert all(a.sensors[0].bit == L.SENSOR_LASERSCAN for a in tc.get_testcase_two_agents_laserscanners())
