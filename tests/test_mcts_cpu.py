"""Host logic of the batched Dec-MCTS planner (casim/mcts.py) against trees grown by the reference's own pydecmcts
Tree + ig_mcts callbacks (tests/golden/mcts.npz, oracle/gen_golden_explore.py:main_mcts).  The rollout evaluator is
the CPU oracle here (the GPU evaluator is covered by tests/test_gpu_rollouts.py and test_gpu_mcts.py): same random
choices -> the same tree, node for node."""
import numpy as np

import helpers as H


def _oracle_evaluator(oracle, cfg, smap, poses, radius, belief):
    from casim.sim import pack_cell_mask, unpack_cell_mask

    def evaluate(actions, exclude):
        T, K, _ = actions.shape
        rew, nf = np.zeros((T, K)), np.zeros((T, K), np.int32)
        for t in range(T):
            ex = None if exclude is None else unpack_cell_mask(exclude[t], cfg.ig_map_dim)
            o = oracle.explore_rollouts(cfg, smap, poses[t], radius, belief, actions[t], exclude=ex)
            rew[t], nf[t] = o["rewards"], o["n_feasible"]
        return rew, nf

    def masks_of(seqs):
        T = seqs.shape[0]
        return np.stack([pack_cell_mask(oracle.explore_rollouts(cfg, smap, poses[t], radius, belief, seqs[t])["obsvd"].astype(bool))
                         for t in range(T)])

    return evaluate, masks_of


def _check_tree(z, k, tree, t, avgs, rtol):
    n = len(z["parent_%d" % k])
    assert int(tree.n_nodes[t]) == n
    assert np.array_equal(tree.parent[t, :n], z["parent_%d" % k])
    assert np.array_equal(tree.stage[t, :n], z["stage_%d" % k])
    assert np.array_equal(tree.path[t, :n], z["path_%d" % k])
    H.assert_close(tree.N[t, :n], z["N_%d" % k], "N of tree %d" % k, rtol=1e-12, atol=0)
    H.assert_close(tree.mu[t, :n], z["mu_%d" % k], "mu of tree %d" % k, rtol=rtol, atol=1e-12)
    H.assert_close(np.array(avgs), z["avg_%d" % k], "avg reward of tree %d" % k, rtol=rtol, atol=1e-12)
    d = tree.distribution()
    nq = len(z["q_%d" % k])
    assert int(d.n[t]) == nq
    H.assert_close(d.q[t, :nq], z["q_%d" % k], "q of tree %d" % k, rtol=10 * rtol, atol=1e-12)
    assert int(tree.best_action_index()[t]) == int(z["action_%d" % k])


def test_mcts_trees_vs_reference(oracle_lib):
    from casim.mcts import BatchedDecMCTS
    oracle = oracle_lib
    z = H.load("mcts.npz")
    cfg = oracle.default_config()
    smap = np.unpackbits(z["static"])[:300 * 300].reshape(300, 300).astype(np.uint8)
    Hn, nsims, ntree, gamma, seed = int(z["H"]), int(z["nsims"]), int(z["ntree"]), float(z["gamma"]), int(z["seed"])
    belief = z["belief"].astype(np.float64)
    radius = float(z["radius"])
    assert (z["stage_1"].max() >= 3) and len(z["parent_1"]) < 1 + 9 * (ntree + 1)      # tree B has infeasible expansions
    dist_a = None
    for k in (0, 1):
        ev, masks_of = _oracle_evaluator(oracle, cfg, smap, z["poses"][k:k + 1], radius, belief)
        np.random.seed(seed + k)
        tree = BatchedDecMCTS(1, ev, horizon=Hn, c_p=1.0, comm_n=5, reference_order=True)
        if k == 1:
            tree.receive_comms(dist_a)
        avgs = [float(tree.grow(nsims=nsims, gamma=gamma)[0]) for _ in range(ntree)]
        _check_tree(z, k, tree, 0, avgs, rtol=1e-9)
        if k == 0:
            dist_a = tree.distribution(with_masks=masks_of)


def test_mcts_batch_is_independent_trees(oracle_lib):
    """Two trees grown in one batch (vectorised bookkeeping, shared node capacity) equal the same trees grown alone."""
    from casim.mcts import BatchedDecMCTS
    oracle = oracle_lib
    z = H.load("mcts.npz")
    cfg = oracle.default_config()
    smap = np.unpackbits(z["static"])[:300 * 300].reshape(300, 300).astype(np.uint8)
    belief = z["belief"].astype(np.float64)
    radius = float(z["radius"])
    Hn = 4

    class Scripted:          # the same draws whatever the batch composition: a per-tree stream
        def __init__(self, seeds):
            self.g = [np.random.default_rng(s) for s in seeds]

        def integers(self, lo, hi, shape, dtype=np.int64):
            return np.stack([g.integers(lo, hi, shape[1:], dtype=dtype) for g in self.g])

        def random(self, n):
            return np.array([g.random() for g in self.g])

    def run(sel):
        ev, _ = _oracle_evaluator(oracle, cfg, smap, z["poses"][sel], radius, belief)
        tree = BatchedDecMCTS(len(sel), ev, horizon=Hn, seed=0)
        tree.rng = Scripted([100 + s for s in sel])
        for _ in range(8):
            tree.grow(nsims=3, gamma=0.9)
        return tree

    both, a, b = run([0, 1]), run([0]), run([1])
    for t, one in ((0, a), (1, b)):
        n = int(one.n_nodes[0])
        assert int(both.n_nodes[t]) == n
        for name in ("parent", "stage", "path", "mu", "N", "best_reward", "best_seq", "has_rollout"):
            assert np.array_equal(getattr(both, name)[t, :n], getattr(one, name)[0, :n]), name
    assert both.launches == 9          # one evaluator call per iteration (+ the root expansion)


def test_hash_draws_known_answers_and_uniformity():
    """casim.mcts.hash_draws (the numpy twin of csrc/casim_mcts.cu:mcts_draw, the in-kernel stand-in for
    np.random.choice(9)) against a plain-integer splitmix64, plus a uniformity check."""
    from casim.mcts import draw_seed, hash_draws
    M = (1 << 64) - 1

    def ref(seed, index):
        x = (seed + 0x9E3779B97F4A7C15 * (index + 1)) & M
        x ^= x >> 30
        x = (x * 0xBF58476D1CE4E5B9) & M
        x ^= x >> 27
        x = (x * 0x94D049BB133111EB) & M
        x ^= x >> 31
        return ((x >> 32) * 9) >> 32

    seed = draw_seed((3, 1, 4))
    assert seed == draw_seed((3, 1, 4)) and seed != draw_seed((3, 1, 5)) and 0 <= seed < (1 << 64)
    T, nsims, H = 5, 3, 7
    for it in (0, 1, 17):
        d = hash_draws(seed, it, T, nsims, H)
        assert d.shape == (T, nsims, H) and d.dtype == np.uint8
        flat = d.ravel()
        for i in (0, 1, 2, 50, T * nsims * H - 1):
            assert int(flat[i]) == ref(seed, it * T * nsims * H + i), (it, i)
    big = hash_draws(seed, 2, 2048, 10, 10)
    freq = np.bincount(big.ravel(), minlength=9) / big.size
    assert big.max() == 8 and np.abs(freq - 1 / 9).max() < 0.005
