"""casim_generate_scenario (the reference's seeded scenario generators on the device) against the scenarios the
reference itself generated (tests/golden/scenarios.npz) and against the CPU oracle at full batch size; and a reset +
episode on generated scenarios without any host-side initial conditions."""
import numpy as np
import pytest

import helpers as H
from casim import layout as L

pytestmark = pytest.mark.gpu


def _sim(E, A, **kw):
    from casim import lib
    from casim.sim import BatchedSim
    cfg = lib.default_config()
    return BatchedSim(E, A, cfg=cfg, **kw)


def test_scenarios_golden_vs_reference():
    import torch
    z = H.load("scenarios.npz")
    A = z["init"].shape[1]
    for kind in (0, 1, 2):
        for nn in sorted(set(z["number_of_agents"].tolist())):
            sel = np.nonzero((z["kind"] == kind) & (z["number_of_agents"] == nn))[0]
            sim = _sim(len(sel), A, sensors=())
            init, types, n, coop = sim.generate_scenario(kind, nn, z["seed"][sel], L.POLICY_LEARNING, L.DYN_FIRST_ORDER,
                                                         L.DYN_UNICYCLE)
            torch.cuda.synchronize()
            init, types, n, coop = init.cpu().numpy(), types.cpu().numpy(), n.cpu().numpy(), coop.cpu().numpy()
            assert np.array_equal(n, z["n_agents"][sel]) and np.array_equal(types, z["types"][sel])
            assert np.array_equal(coop, z["coop"][sel].astype(np.float32))
            for k, g in enumerate(sel):
                m = int(n[k])
                H.assert_close(init[k, :m, :6], z["init"][g, :m, :6], "kind %d n %d seed %d" % (kind, nn, z["seed"][g]),
                               rtol=1e-6, atol=1e-6)
                assert np.isnan(init[k, :m, 6]).all() and not init[k, m:].any()
            sim.close()


def test_scenarios_full_size_vs_oracle_and_episode(oracle_lib):
    """4096 environments x 10 agents: bit-level agreement with the oracle on policies / counts, fp32-rounding
    agreement on positions, minimum separations of is_pose_valid, and an episode started from the device buffers."""
    import torch
    oracle = oracle_lib
    E, A = 4096, 10
    seeds = (np.arange(E, dtype=np.uint32) * 7919 + 11).astype(np.uint32)
    for kind in (L.SCENARIO_SWAP_CIRCLE, L.SCENARIO_RANDOM_POSITIONS, L.SCENARIO_PAIRWISE_SWAP):
        sim = _sim(E, A, sensors=("other_agents_states",))
        init, types, n, coop = sim.generate_scenario(kind, A, seeds, L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE)
        torch.cuda.synchronize()
        o = oracle.scenario(kind, E, A, A, seeds, L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, L.DYN_UNICYCLE)
        gi = init.cpu().numpy()
        assert np.array_equal(types.cpu().numpy(), o["types"]) and np.array_equal(n.cpu().numpy(), o["n_agents"])
        H.assert_close(gi[..., :6], o["init"][..., :6], "kind %d init" % kind, rtol=1e-6, atol=1e-6)
        start = gi[..., 0:2].astype(np.float64)
        d = np.linalg.norm(start[:, :, None] - start[:, None, :], axis=-1) + 10 * np.eye(A)[None]
        assert d.min() > (2.0 if kind == 2 else 1.5) - 1e-5                       # is_pose_valid(distance=1.5) on every pair of start positions
        pol = o["types"] & 0xff
        frac_nc = (pol[:, 1:] == L.POLICY_NONCOOPERATIVE).mean()
        assert (0.45 < frac_nc < 0.55) if kind == 1 else (0.15 < frac_nc < 0.25)       # random.choice of two / P(u > 0.8)
        # RVO agents need the RVO step; run the episode with them as NonCooperative (same slots, same geometry)
        t2 = types.clone()
        t2 = (t2 & ~0xff) | L.POLICY_NONCOOPERATIVE
        t2 = (t2 & 0xffff) | (L.SENSOR_OTHER_AGENTS << 16)
        sim.reset(init, t2, n)
        p0 = sim.state[0, :, :, 0:2].clone()
        for _ in range(20):
            sim.step()
        torch.cuda.synchronize()
        moved = (sim.state[0, :, :, 0:2] - p0).norm(dim=-1)
        # (agents that collide early stop moving: the floor is one step, the typical agent travels 20 steps x 0.1 s x 1 m/s)
        assert float(moved.min()) > 0.05 and float(moved.median()) > 1.9 and bool(torch.isfinite(sim.state).all())
        sim.close()


def test_env_generates_episodes_like_init_agents():
    """BatchedCollisionAvoidanceEnv.set_scenario_generator: reset() builds the episode on the device with
    seed = episode_number, as _init_agents does in evaluate mode (collision_avoidance_env.py:403-415); at E = 1 the
    episodes are the reference's (scenarios.npz seeds 1, 2, 3), and the RVO / NonCooperative mix steps."""
    import torch
    from casim import plugins as P
    from casim.env import BatchedCollisionAvoidanceEnv
    from casim.config import Config

    class Cfg(Config):
        TRAIN_SINGLE_AGENT = False

    z = H.load("scenarios.npz")
    for kind in (L.SCENARIO_SWAP_CIRCLE, L.SCENARIO_PAIRWISE_SWAP):
        env = BatchedCollisionAvoidanceEnv(num_envs=1, max_agents=z["init"].shape[1], config=Cfg)
        env.set_scenario_generator(kind, 6, P.LearningPolicy, P.FirstOrderDynamics, P.UnicycleDynamics)
        for episode in (1, 2, 3):
            env.reset()
            g = int(np.nonzero((z["kind"] == kind) & (z["number_of_agents"] == 6) & (z["seed"] == episode))[0][0])
            n = int(z["n_agents"][g])
            assert int(env.sim._keep["n_agents"][0]) == n
            assert np.array_equal(env.sim._keep["types"].cpu().numpy()[0], z["types"][g])
            H.assert_close(env.sim._keep["init"].cpu().numpy()[0, :n, :6], z["init"][g, :n, :6], "episode %d" % episode,
                           rtol=1e-6, atol=1e-6)
            for _ in range(5):
                obs, rewards, game_over, info = env.step(np.full((1, z["init"].shape[1], 2), 0.5, np.float32))
            torch.cuda.synchronize()
            st = env.sim.state_numpy()[0]
            moved = np.linalg.norm(st[:n, [L.PX, L.PY]] - z["init"][g, :n, 0:2], axis=1)
            assert np.isfinite(st).all() and moved[1:].sum() > 0.5                # (RVO agents may yield: slow, even still)
        env.close()
    E = 256
    env = BatchedCollisionAvoidanceEnv(num_envs=E, max_agents=8, config=Cfg)
    env.set_scenario_generator(L.SCENARIO_RANDOM_POSITIONS, 8, P.NonCooperativePolicy, P.UnicycleDynamics, seed_base=1000)
    env.reset()
    first = env.sim._keep["init"].clone()
    for _ in range(10):
        obs, rewards, game_over, info = env.step()
    env.reset()
    assert not torch.equal(first, env.sim._keep["init"])          # a new episode is a new scenario
    assert rewards.shape == (E, 8) and bool(torch.isfinite(env.sim.state).all())
    env.close()
