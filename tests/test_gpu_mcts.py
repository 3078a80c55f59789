"""The batched Dec-MCTS planner with the GPU rollout evaluator (casim_explore_rollouts): the reference-grown trees
of tests/golden/mcts.npz node for node, and a C4-size planning step (2048 trees) for determinism and sanity."""
import numpy as np
import pytest

import helpers as H
from casim import layout as L

pytestmark = pytest.mark.gpu


def _cfg():
    from casim import lib
    c = lib.default_config()
    c.game_over_mode = 0
    return c


def _sim(E, A, cfg):
    from casim.sim import BatchedSim
    sim = BatchedSim(E, A, cfg=cfg, sensors=("other_agents_states",))
    sim.enable_exploration()
    return sim


def test_mcts_golden_trees_on_gpu():
    import torch
    from casim.mcts import BatchedDecMCTS, sim_evaluator
    z = H.load("mcts.npz")
    cfg = _cfg()
    smap = np.unpackbits(z["static"])[:300 * 300].reshape(300, 300).astype(np.uint8)
    Hn, nsims, ntree, gamma, seed = int(z["H"]), int(z["nsims"]), int(z["ntree"]), float(z["gamma"]), int(z["seed"])
    dist_a = None
    for k in (0, 1):
        sim = _sim(1, 1, cfg)
        sim.set_map(smap)
        state = np.zeros((1, 1, L.STATE_WORDS))
        state[0, 0, L.PX], state[0, 0, L.PY], state[0, 0, L.HEADING] = z["poses"][k]
        state[0, 0, L.RADIUS] = float(z["radius"])
        state[0, 0, L.PREF_SPEED] = 1.0
        state[0, 0, L.TYPE] = L.encode_type(L.POLICY_IG_GREEDY, L.DYN_FIRST_ORDER, L.SENSOR_OTHER_AGENTS)
        sim.bind(types=state[:, :, L.TYPE].astype(np.int32), n_agents=np.ones((1,), np.int32))
        sim.load_state_numpy(state)
        sim.belief.copy_(torch.from_numpy(z["belief"])[None, None])
        T, ev, masks_of, _ = sim_evaluator(sim)
        assert T == 1
        np.random.seed(seed + k)
        tree = BatchedDecMCTS(T, ev, horizon=Hn, c_p=1.0, comm_n=5, reference_order=True)
        if k == 1:
            tree.receive_comms(dist_a)
        avgs = [float(tree.grow(nsims=nsims, gamma=gamma)[0]) for _ in range(ntree)]
        n = len(z["parent_%d" % k])
        assert int(tree.n_nodes[0]) == n
        assert np.array_equal(tree.parent[0, :n], z["parent_%d" % k])
        assert np.array_equal(tree.path[0, :n], z["path_%d" % k])
        H.assert_close(tree.mu[0, :n], z["mu_%d" % k], "mu of tree %d" % k, rtol=1e-5, atol=1e-7)
        H.assert_close(np.array(avgs), z["avg_%d" % k], "avg reward of tree %d" % k, rtol=1e-5, atol=1e-7)
        assert int(tree.best_action_index()[0]) == int(z["action_%d" % k])
        if k == 0:
            dist_a = tree.distribution(with_masks=masks_of)
        sim.close()


def test_mcts_c4_planning_step():
    """C4 size: 1024 envs x (2 explorers + 2 targets): 20 iterations x 10 simulations x horizon 10 for all 2048 trees;
    the two explorers of an environment exchange their distributions half way (Dec-MCTS comms)."""
    import torch
    from casim import scenarios as S
    from casim.mcts import BatchedDecMCTS, sim_evaluator, action_value
    E = 1024
    init, types, n, rects = S.exploration(E, seed=5)
    cfg = _cfg()
    sim = _sim(E, 4, cfg)
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    for _ in range(2):
        sim.explore_step()
        sim.step(policy_actions=sim.buf["policy_actions"])

    def plan(seed):
        T, ev, masks_of, (te, ta) = sim_evaluator(sim)
        assert T == 2 * E
        tree = BatchedDecMCTS(T, ev, horizon=10, c_p=1.0, seed=seed)
        for _ in range(10):
            tree.grow(nsims=10, gamma=1.0)
        d = tree.distribution(with_masks=masks_of)
        # trees are in (env, agent) order and every env has exactly two explorers: the partner is the neighbour
        assert np.array_equal(te[0::2], te[1::2])
        partner = np.arange(T) ^ 1
        from casim.mcts import ActionDistribution
        other = ActionDistribution(d.seqs[partner], d.q[partner], d.masks[partner])
        tree.receive_comms(other)
        avg = None
        for _ in range(10):
            avg = tree.grow(nsims=10, gamma=1.0)
        return tree, avg

    t1, avg1 = plan(3)
    t2, avg2 = plan(3)
    assert np.array_equal(t1.n_nodes, t2.n_nodes) and np.array_equal(t1.mu, t2.mu) and np.array_equal(avg1, avg2)
    a = t1.best_action_index()
    assert a.min() >= 0 and a.max() <= 8
    assert t1.launches == 21 and (t1.n_nodes > 10).all() and (t1.n_nodes <= 1 + 9 * 21).all()
    act = action_value(a)
    assert act.shape == (2 * E, 2) and np.isin(act[:, 0], (0.0, 2.0, 4.0)).all()
    assert float(t1.mu[:, 1:].max()) > 0.3 and np.isfinite(t1.mu).all()
    sim.close()


def test_device_trees_equal_host_trees_node_for_node():
    """DeviceDecMCTS (casim_mcts_select / casim_mcts_update: selection, expansion, statistics, back-propagation and the
    sampling of the other robots' claimed cells on the GPU) against BatchedDecMCTS (the host planner the golden test
    pins to the reference's trees) from the same seed: every node array identical, with and without communication,
    discount < 1, a tree that reaches the horizon, and a re-allocation of the node arrays on the way."""
    from casim import scenarios as S
    from casim.mcts import ActionDistribution, BatchedDecMCTS, DeviceDecMCTS, sim_evaluator
    E = 96
    init, types, n, rects = S.exploration(E, seed=11)
    cfg = _cfg()
    sim = _sim(E, 4, cfg)
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    for _ in range(3):
        sim.explore_step()
        sim.step(policy_actions=sim.buf["policy_actions"])
    T, ev, masks_of, (te, ta) = sim_evaluator(sim)
    sel = np.zeros((E, 4), bool)
    sel[te, ta] = True
    partner = np.arange(T) ^ 1

    def same(dev, host, what):
        dev.sync_host()
        assert np.array_equal(dev.n_nodes, host.n_nodes), what
        nmax = int(host.n_nodes.max())
        live = np.arange(nmax)[None, :] < host.n_nodes[:, None]
        for k in ("parent", "stage", "nchild", "has_rollout", "mu", "N", "best_reward"):
            assert np.array_equal(getattr(dev, k)[:, :nmax][live], getattr(host, k)[:, :nmax][live]), (what, k)
        for k in ("child", "path", "best_seq"):
            assert np.array_equal(getattr(dev, k)[:, :nmax][live], getattr(host, k)[:, :nmax][live]), (what, k)
        assert np.array_equal(dev.last_node, host.last_node), what

    # horizon 2: leaves at stage == H are re-simulated; hashed: the simulation actions drawn inside casim_mcts_select
    for Hn, gamma, nsims, hashed in ((10, 0.9, 10, False), (2, 1.0, 4, False), (10, 0.95, 10, True)):
        host = BatchedDecMCTS(T, ev, horizon=Hn, c_p=1.0, seed=7, hashed_draws=hashed)
        dev = DeviceDecMCTS(sim, sel, horizon=Hn, c_p=1.0, seed=7, nsims=nsims, n_grows=4,    # too small on purpose
                            device_draws=hashed)
        for _ in range(12):
            host.grow(nsims=nsims, gamma=gamma)
        dev.reserve(12)
        dev.grow_n(12, nsims=nsims, gamma=gamma)
        same(dev, host, "H=%d no comms" % Hn)
        dh, dd = host.distribution(with_masks=masks_of), dev.distribution(with_masks=masks_of)
        assert np.array_equal(dh.seqs, dd.seqs) and np.array_equal(dh.q, dd.q) and np.array_equal(dh.masks, dd.masks)
        assert np.array_equal(host.best_action_index(), dev.best_action_index())
        # communication: every tree receives its partner's distribution; a third of the trees have no partner entry
        valid = (np.arange(T) % 3) != 0
        for tr, dist in ((host, dh), (dev, dd)):
            tr.receive_comms(ActionDistribution(dist.seqs[partner], dist.q[partner], dist.masks[partner]), valid=valid)
            tr.receive_comms(ActionDistribution(dist.seqs, dist.q, dist.masks), valid=np.zeros(T, bool), draw_only=valid)
        avg_h = None
        for _ in range(9):
            avg_h = host.grow(nsims=nsims, gamma=gamma)
        dev.reserve(9)
        dev.grow_n(8, nsims=nsims, gamma=gamma)
        avg_d = dev.grow(nsims=nsims, gamma=gamma)
        same(dev, host, "H=%d comms" % Hn)
        assert np.array_equal(avg_h, avg_d)
        assert dev.launches == host.launches == 22
        assert (host.n_nodes > 12).all()
        if Hn == 2:
            assert (host.stage[np.arange(T), host.last_node] == 2).any()
    # an under-sized node array is refused, not overrun
    small = DeviceDecMCTS(sim, sel, horizon=10, seed=1, nsims=4, n_grows=1)
    small.grow_n(6, gamma=1.0)
    with pytest.raises(Exception, match="node arrays full"):
        small.sync_host()
    sim.close()


def test_env_runs_ig_mcts_policy():
    """ig_mcts agents driven by the env itself: team belief update + Dec-MCTS planning inside step(); two explorers
    and two static targets per environment, explorers first (the reference's layout)."""
    import torch
    from casim import plugins as P
    from casim.env import BatchedCollisionAvoidanceEnv
    from casim.config import Config

    class Cfg(Config):
        pass

    E = 6
    rng = np.random.default_rng(1)
    envs_agents = []
    for e in range(E):
        x = float(rng.uniform(-9, -5))
        ags = [P.Agent(x, 0.4, 9, 0, 0.5, 1.0, 0.0, P.ig_mcts, P.FirstOrderDynamics, [P.OtherAgentsStatesSensor], 0),
               P.Agent(0.3, -8 + e * 0.3, 0, 9, 0.5, 1.0, 1.5, P.ig_mcts, P.FirstOrderDynamics, [P.OtherAgentsStatesSensor], 1),
               P.Agent(x + 3.0, -0.3, 0, 0, 0.2, 1.0, 0.0, P.StaticPolicy, P.UnicycleDynamics, [], 2),
               P.Agent(0.0, -3.0, 0, 0, 0.2, 1.0, 0.0, P.StaticPolicy, P.UnicycleDynamics, [], 3)]
        for a in ags[:2]:
            a.policy.set_param(Ntree=6, Nsims=4, mcts_horizon=6, Ncycles=2, mcts_gamma=1)
        envs_agents.append(ags)
    crossing = [[(10, 10), (2, 10), (2, 2), (10, 2)], [(-2, 10), (-10, 10), (-10, 2), (-2, 2)],
                [(10, -2), (2, -2), (2, -10), (10, -10)], [(-2, -2), (-10, -2), (-10, -10), (-2, -10)]]

    def run():
        env = BatchedCollisionAvoidanceEnv(num_envs=E, max_agents=4, config=Cfg)
        env.set_agents(envs_agents)
        env.set_static_map(crossing)
        env.reset()
        p0 = env.sim.state[0, :, :, 0:2].clone()
        acts = []
        for t in range(4):
            obs, rewards, game_over, info = env.step()
            acts.append(env.sim.buf["policy_actions"].cpu().numpy().copy())
        torch.cuda.synchronize()
        out = (env.sim.state.cpu().numpy().copy(), env.sim.belief.cpu().numpy().copy(), np.array(acts),
               (env.sim.state[0, :, :, 0:2] - p0).norm(dim=-1).cpu().numpy(), env._mcts)
        env.close()
        return out

    state, belief, acts, moved, planner = run()
    assert np.isin(acts[:, :, :2, 0], (0.0, 2.0, 4.0)).all()
    assert np.isin(np.round(acts[:, :, :2, 1] / (0.5 * np.pi)), (-1, 0, 1)).all() and (acts[:, :, 2:] == 0).all()
    assert (moved[:, :2] > 0.2).any() and (moved[:, 2:] == 0).all()
    assert (belief[:, :2] != 1.0).any() and (belief[:, 2:] == 1.0).all()
    # explorer 1 plans after explorer 0 and receives its distribution (the only comms the reference's indexing yields)
    assert (planner.best[0]["q"].sum(1) > 0).all() and \
        any(v.any() for _, v, _ in planner.trees[1][0].comms) and not any(v.any() for _, v, _ in planner.trees[0][0].comms)
    state2, belief2, acts2, _, _ = run()
    assert np.array_equal(acts, acts2) and np.array_equal(state, state2) and np.array_equal(belief, belief2)


def test_env_step_with_ig_mcts_vs_reference_env():
    """The reference's own CollisionAvoidanceEnv.step on the fork's exploration scenario (tc.IG_agent_crossing: three
    ig_mcts explorers + two static targets, crossing map; tests/golden/mcts_env.npz) against this env's step with the
    same planner parameters and the same np.random stream: planned actions exact, poses / target maps to 1e-5."""
    import torch
    from casim import plugins as P
    from casim.env import BatchedCollisionAvoidanceEnv
    from casim.config import Config

    class Cfg(Config):
        TRAIN_SINGLE_AGENT = False        # as in the harness that recorded the golden (rewards for every agent)

    z = H.load("mcts_env.npz")
    n_ig = int(z["n_ig"])
    agents = []
    for k, row in enumerate(z["init"]):
        if k < n_ig:
            a = P.Agent(row[0], row[1], row[2], row[3], row[4], row[5], row[6], P.ig_mcts, P.FirstOrderDynamics,
                        [P.OtherAgentsStatesSensor, P.LaserScanSensor], k)
            a.policy.set_param(Ntree=int(z["Ntree"]), Nsims=int(z["Nsims"]), mcts_horizon=int(z["H"]),
                               Ncycles=int(z["Ncycles"]), mcts_gamma=1)
            a.policy.reference_order = True
        else:
            a = P.Agent(row[0], row[1], row[2], row[3], row[4], row[5], row[6], P.StaticPolicy, P.FirstOrderDynamics, [], k)
        agents.append(a)
    crossing = [[(10, 10), (2, 10), (2, 2), (10, 2)], [(-2, 10), (-10, 10), (-10, 2), (-2, 2)],
                [(10, -2), (2, -2), (2, -10), (10, -10)], [(-2, -2), (-10, -2), (-10, -10), (-2, -10)]]
    env = BatchedCollisionAvoidanceEnv(num_envs=1, max_agents=len(agents), config=Cfg)
    env.set_agents(agents)
    env.set_static_map(crossing)
    env.reset()
    for step in range(z["pose"].shape[0]):
        np.random.seed(int(z["seed"]) + step)
        obs, rewards, game_over, info = env.step()
        torch.cuda.synchronize()
        acts = env.sim.buf["policy_actions"].cpu().numpy()[0]
        assert np.array_equal(acts[:n_ig].astype(np.float64), z["action"][step].astype(np.float32).astype(np.float64)), step
        st = env.sim.state_numpy()[0]
        pose = st[:, [L.PX, L.PY, L.HEADING]]
        H.assert_close(pose, z["pose"][step], "pose after step %d" % step, rtol=1e-5, atol=1e-5)
        b = env.sim.belief.cpu().numpy()[0, :n_ig].astype(np.float64)
        H.assert_close(b, z["belief"][step], "target maps after step %d" % step, rtol=1e-5, atol=0)
        assert np.array_equal(env.sim.buf["team_n"][0].cpu().numpy()[0, :n_ig], z["n_targets"][step])
        H.assert_close(rewards.cpu().numpy().reshape(-1), z["rewards"][step], "rewards step %d" % step, rtol=1e-5, atol=1e-6)
    env.close()
