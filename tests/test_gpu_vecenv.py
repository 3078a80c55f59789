"""DummyVecEnv semantics of BatchedVecEnv and the masked restart behind it (casim_reset_envs):
an environment that reports game_over is reset inside the same call, the observation row returned for it is the new
episode's first observation, the finished episode's last observation travels in infos['terminal_observation'], and
every other environment is bit-for-bit untouched (envs/wrappers.py:101-106 MultiagentDummyVecEnv; stable_baselines
dummy_vec_env.step_wait).  Also the gym-facing attributes (collision_avoidance_env.py:103-134)."""
import numpy as np
import pytest

import helpers as H
from casim import layout as L

pytestmark = pytest.mark.gpu

KEYS = ['dist_to_goal', 'rel_goal', 'radius', 'heading_ego_frame', 'pref_speed', 'other_agents_states', 'laserscan', 'use_ppo']


def _scenario(E, A, seed=5):
    rng = np.random.default_rng(seed)
    ang = rng.uniform(-np.pi, np.pi, (E, A))
    rad = rng.uniform(1.0, 2.5, (E, A))
    start = np.stack([rad * np.cos(ang), rad * np.sin(ang)], -1)
    init = np.concatenate([start, -start, rng.uniform(0.2, 0.4, (E, A, 1)), rng.uniform(0.8, 1.4, (E, A, 1)),
                           np.full((E, A, 1), np.nan)], -1).astype(np.float32)
    types = np.full((E, A), L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS | L.SENSOR_LASERSCAN),
                    np.int32)
    n = rng.integers(2, A + 1, E).astype(np.int32)
    return init, types, n


def test_masked_reset_leaves_other_envs_untouched(oracle_lib):
    import torch
    from casim import lib
    from casim.sim import BatchedSim
    E, A = 200, 6
    init, types, n = _scenario(E, A)
    maps, rects = H.unpack_maps()
    cfg = lib.default_config()
    sens = ("other_agents_states", "laserscan", "local_grid")
    types = types | (L.SENSOR_OCCUPANCY_GRID << 16)
    sim = BatchedSim(E, A, cfg=cfg, sensors=sens)
    sim.set_obstacles(rects[1].astype(np.float32))
    sim.reset(init, types, n)
    first = {k: sim.buf[k].clone() for k in sens + ("obs_misc",)}
    state_first = sim.state.clone()
    for _ in range(25):
        sim.step()
    torch.cuda.synchronize()
    before = {k: sim.buf[k].clone() for k in sens + ("obs_misc", "rewards", "done", "game_over")}
    state_before = sim.state.clone()
    mask = sim.buf["game_over"].clone()
    mask[::3] = 1                                     # plus an arbitrary third of the batch
    mb = mask.bool()
    assert 0 < int(mb.sum()) < E
    sim.reset_envs(mask)
    torch.cuda.synchronize()
    assert torch.equal(sim.state[:, ~mb], state_before[:, ~mb])
    assert torch.equal(sim.state[:, mb], state_first[:, mb])
    for k in first:
        assert torch.equal(sim.buf[k][~mb], before[k][~mb]), k          # untouched, bit for bit
        assert torch.equal(sim.buf[k][mb], first[k][mb]), k             # == the observation reset() produced
    for k in ("rewards", "done", "game_over"):
        assert torch.equal(sim.buf[k], before[k]), k
    sim.close()


def test_vecenv_dummyvecenv_semantics():
    import torch
    from casim.env import BatchedCollisionAvoidanceEnv, make
    from casim.wrappers import BatchedVecEnv, MultiagentFlattenDictWrapper
    E, A = 64, 5
    init, types, n = _scenario(E, A, seed=8)
    maps, rects = H.unpack_maps()

    from casim.config import Config

    class Cfg(Config):
        TRAIN_SINGLE_AGENT = False

    def build():
        env = make("CollisionAvoidance-v0", num_envs=E, max_agents=A, config=Cfg)
        env.set_scenario_tensors(init, types, n)
        env.set_static_map(rects[1].astype(np.float32))
        return env, MultiagentFlattenDictWrapper(env, KEYS, max_num_agents=A)

    env, w = build()
    assert isinstance(env, BatchedCollisionAvoidanceEnv)
    venv = BatchedVecEnv(w)
    ref_env, ref_w = build()            # the same batch stepped WITHOUT restarts, as the per-episode ground truth
    obs0 = venv.reset()
    ref_obs0 = ref_w.reset()
    assert torch.equal(obs0, ref_obs0)
    # gym-facing attributes
    assert env.action_space.shape == (2,) and float(env.action_space.high[1]) == 4.0
    assert set(env.observation_space.spaces) == set(env.config.STATES_IN_OBS)
    assert w.observation_space.shape == (w.size,) and w.single_agent_observation_space.shape == (w.size // A,)
    seen_restart = 0
    alive = torch.ones(E, dtype=torch.bool, device=obs0.device)      # envs still in their FIRST episode
    for t in range(60):
        obs, rew, done, info = venv.step()
        robs, rrew, rdone, _ = ref_w.step()
        torch.cuda.synchronize()
        # while an env is in its first episode it is indistinguishable from the never-restarted batch
        assert torch.equal(rew[alive], rrew[alive]) and torch.equal(done[alive], rdone[alive])
        fin = done & alive
        if bool(done.any()):
            assert torch.equal(info["terminal_mask"], done)
            assert torch.equal(info["terminal_observation"][fin], robs[fin])        # last obs of the finished episode
            assert torch.equal(obs[done], obs0[done])                                # first obs of the new episode
            seen_restart += int(done.sum())
        keep = alive & ~done
        assert torch.equal(obs[keep], robs[keep])
        alive = keep
    assert seen_restart >= 8
    # multiEnvObservationArrayToDict (wrappers.py:58-73): only 'use_ppo' agents are filled -> none here
    d = w.multiEnvObservationArrayToDict(obs)
    assert d.shape == (E, A) and all(x is None for x in d.ravel())
    venv.close()
    ref_env.close()


def test_vecenv_generator_draws_next_scenario_for_finished_envs():
    import torch
    from casim import plugins as P
    from casim.env import BatchedCollisionAvoidanceEnv
    from casim.wrappers import BatchedVecEnv, MultiagentFlattenDictWrapper
    E, A = 32, 4
    from casim.config import Config

    class Cfg(Config):
        TRAIN_SINGLE_AGENT = False

    env = BatchedCollisionAvoidanceEnv(num_envs=E, max_agents=A, config=Cfg)
    env.set_scenario_generator(L.SCENARIO_RANDOM_POSITIONS, 4, P.NonCooperativePolicy, P.UnicycleDynamics)
    keys = ['dist_to_goal', 'rel_goal', 'radius', 'pref_speed', 'other_agents_states']
    venv = BatchedVecEnv(MultiagentFlattenDictWrapper(env, keys, max_num_agents=A))
    venv.reset()
    init0 = env._gen_out[0].clone()
    ep0 = env._env_episode.copy()
    restarted = np.zeros(E, bool)
    for t in range(300):
        _, _, done, _ = venv.step()
        restarted |= done.cpu().numpy()
        if restarted.sum() >= 4:
            break
    torch.cuda.synchronize()
    assert restarted.sum() >= 1
    init1 = env._gen_out[0]
    bits = lambda t: t.contiguous().view(torch.int32)       # noqa: E731  (the heading column is NaN = "towards the goal")
    same = (bits(init1) == bits(init0)).flatten(1).all(1).cpu().numpy()
    assert np.array_equal(~same, restarted)                          # exactly the finished envs got a new scenario
    assert np.array_equal(env._env_episode > ep0, restarted)
    # a restarted env's scenario is the generator's answer for ITS seed: regenerate from scratch and compare
    seeds = (env._env_episode * E + np.arange(E) + 1).astype(np.uint32)
    chk = env.sim.generate_scenario(L.SCENARIO_RANDOM_POSITIONS, 4, seeds, L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE)
    assert torch.equal(bits(chk[0]), bits(init1))
    venv.close()
