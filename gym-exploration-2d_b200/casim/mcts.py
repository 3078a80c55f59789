"""Dec-MCTS planner for the explorer agents of a batch: the reference's ``Tree`` (policies/pydecmcts/DecMCTS.py:62-420)
with ig_mcts's callbacks (policies/ig_mcts.py:79-109, 173-265), grown for every tree in lock-step -- one node per
tree per ``grow`` -- with the expensive part (expansion feasibility and the ``nsims`` simulations, 86 % of the
reference's step time) evaluated by ONE ``casim_explore_rollouts`` launch per iteration.

What stays on the host is the tree bookkeeping, as arrays over [tree, node]: UCT selection (``_select``, :140-153),
expansion (``_expansion``, :201-231), statistics + discounted back-propagation (``grow``, :334-357), the top-n action
distribution (``_update_distribution``, :162-180) and the communication of best rollouts between the robots of one
environment (``send_comms`` / ``receive_comms`` / ``_get_system_state``, :182-194, :364-380).

Node numbering follows the reference (creation order; node 0 here = node 1 there), so that a tree grown here and a
tree grown by the reference with the same random choices are the same tree.
"""
import numpy as np

N_ACTIONS = 9
END = 255
VEL_LIST = (0.0, 2.0, 4.0)                       # ig_mcts.py:259-265 mcts_avail_actions
DPHI_LIST = (-0.5 * np.pi, 0.0, 0.5 * np.pi)


def action_value(idx):
    """action index -> the [vel, dphi] array of mcts_avail_actions (vel-major order)."""
    idx = np.asarray(idx)
    return np.stack([np.take(VEL_LIST, idx // 3), np.take(DPHI_LIST, idx % 3)], -1)


def draw_seed(seed):
    """any numpy seed (int / tuple) -> the 64-bit seed of the counter-based draws"""
    return int(np.random.SeedSequence(seed).generate_state(1, np.uint64)[0])


def hash_draws(seed64, iteration, T, nsims, H):
    """The draws casim_mcts_select makes itself when it is given none (csrc/casim_mcts.cu:mcts_draw): splitmix64 of
    (seed, global draw index), top 32 bits scaled to 0..8.  -> uint8 [T,nsims,H]"""
    with np.errstate(over="ignore"):
        idx = np.uint64(iteration) * np.uint64(T * nsims * H) + np.arange(T * nsims * H, dtype=np.uint64)
        x = np.uint64(seed64) + np.uint64(0x9E3779B97F4A7C15) * (idx + np.uint64(1))
        x ^= x >> np.uint64(30)
        x *= np.uint64(0xBF58476D1CE4E5B9)
        x ^= x >> np.uint64(27)
        x *= np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
        return (((x >> np.uint64(32)) * np.uint64(9)) >> np.uint64(32)).astype(np.uint8).reshape(T, nsims, H)


class ActionDistribution:
    """DecMCTS.py:21-59 for every tree: ``seqs`` [T,n,H] uint8 best rollouts of the top-n nodes (END padded), ``q`` [T,n]
    normalised weights (0 = slot unused), ``masks`` [T,n,128] int32 observed-cell masks of those rollouts."""

    def __init__(self, seqs, q, masks=None):
        self.seqs, self.masks = seqs, masks
        q = np.asarray(q, np.float64)
        tot = q.sum(1, keepdims=True)
        used = (seqs[:, :, 0] != END)
        n_used = used.sum(1, keepdims=True)
        uniform = np.where(used, 1.0 / np.maximum(n_used, 1), 0.0)        # sum(q) == 0 -> uniform (:39-40)
        self.q = np.where(tot > 0, q / np.where(tot > 0, tot, 1.0), uniform)
        self.n = n_used[:, 0]


class BatchedDecMCTS:
    """``evaluate(actions [T,K,H] uint8, exclude [T,128] int32 | None) -> (rewards [T,K], n_feasible [T,K])`` is the
    rollout evaluator (``sim_evaluator`` below wraps casim_explore_rollouts).  T trees, all rooted at their agent's
    current pose."""

    def __init__(self, n_trees, evaluate, horizon=10, c_p=1.0, comm_n=5, seed=0, reference_order=False, hashed_draws=False):
        self.T, self.H, self.c_p, self.comm_n = int(n_trees), int(horizon), float(c_p), int(comm_n)
        self.evaluate = evaluate
        self.hashed_draws = draw_seed(seed) if hashed_draws else None   # draw like DeviceDecMCTS(device_draws=True)
        self.iteration = 0
        self.reference_order = bool(reference_order)       # draw from np.random exactly like the reference (tests)
        self.rng = np.random.default_rng(seed)
        self.comms = []                                    # list of (ActionDistribution, valid [T] bool)
        self.launches = 0
        self._alloc(1 + N_ACTIONS * 4)
        T = self.T
        self.n_nodes = np.ones(T, np.int64)
        # root (Tree.__init__, :120-133): mu = N = best_reward = 0, stage 0, then _expansion(root)
        self._expand(np.zeros(T, np.int64), self._rollouts(np.zeros(T, np.int64), 0, None)[1])

    # ---------------------------------------------------------------- storage
    def _alloc(self, M):
        T, H = self.T, self.H
        new = dict(parent=np.full((T, M), -1, np.int64), stage=np.zeros((T, M), np.int64),
                   mu=np.zeros((T, M)), N=np.zeros((T, M)), best_reward=np.zeros((T, M)),
                   child=np.full((T, M, N_ACTIONS), -1, np.int64), nchild=np.zeros((T, M), np.int64),
                   path=np.full((T, M, H), END, np.uint8), has_rollout=np.zeros((T, M), bool),
                   best_seq=np.full((T, M, H), END, np.uint8))
        if hasattr(self, "M"):
            for k, v in new.items():
                v[:, :self.M] = getattr(self, k)
        for k, v in new.items():
            setattr(self, k, v)
        self.M = M

    def _rows(self):
        return np.arange(self.T)

    # ---------------------------------------------------------------- the one GPU call per iteration
    def _rollouts(self, node, nsims, exclude):
        """Sequences for `node` of every tree: its 9 one-step extensions (expansion feasibility) followed by `nsims`
        random completions up to the horizon (_simulate, :233-269).  Returns (sim_seqs [T,nsims,H], feasible [T,9],
        rewards [T,nsims]).  The sequences keep the REQUESTED actions: replaying one reproduces its infeasible
        steps, whereas the reference's stored [0, 0] substitute (:243) would be a feasible stand-still step."""
        T, H = self.T, self.H
        r = self._rows()
        path, stage = self.path[r, node], self.stage[r, node]
        seqs = np.repeat(path[:, None, :], N_ACTIONS + nsims, axis=1)
        can = stage < H
        ext = seqs[:, :N_ACTIONS]
        ext[r[can], :, stage[can]] = np.arange(N_ACTIONS, dtype=np.uint8)[None, :]
        if nsims:
            if self.reference_order:        # per tree, per simulation, per step: np.random.choice(len(options)), ig_mcts.py:257
                draws = np.full((T, nsims, H), END, np.uint8)
                for t in range(T):
                    for i in range(nsims):
                        for d in range(stage[t], H):
                            draws[t, i, d] = np.random.choice(N_ACTIONS)
            elif self.hashed_draws is not None:
                draws = hash_draws(self.hashed_draws, self.iteration, T, nsims, H)
                self.iteration += 1
            else:
                draws = self.rng.integers(0, N_ACTIONS, (T, nsims, H), dtype=np.uint8)
            tail = np.arange(H)[None, None, :] >= stage[:, None, None]
            seqs[:, N_ACTIONS:] = np.where(tail, draws, seqs[:, N_ACTIONS:])
        rewards, n_feasible = self.evaluate(seqs, exclude)
        self.launches += 1
        feasible = can[:, None] & (n_feasible[:, :N_ACTIONS] == (stage + 1)[:, None])
        return seqs[:, N_ACTIONS:], feasible, np.asarray(rewards[:, N_ACTIONS:], np.float64)

    def reserve(self, n_grows):
        """Room for `n_grows` more iterations (each adds at most 9 nodes per tree) without re-allocating."""
        need = int(self.n_nodes.max()) + N_ACTIONS * int(n_grows)
        if need > self.M:
            self._alloc(need)

    def _expand(self, node, feasible):
        """_expansion (:201-231): a child per feasible action, in action order; nothing at stage == horizon."""
        need = int((self.n_nodes + N_ACTIONS).max())
        if need > self.M:
            self._alloc(max(need, 2 * self.M))
        tt, aa = np.nonzero(feasible)                       # row-major: per tree in action order
        if tt.size == 0:
            return
        k = (np.cumsum(feasible, 1) - 1)[tt, aa]            # position among the tree's new children
        par = node[tt]
        new = self.n_nodes[tt] + k
        st = self.stage[tt, par]
        self.parent[tt, new] = par
        self.stage[tt, new] = st + 1
        self.path[tt, new] = self.path[tt, par]
        self.path[tt, new, st] = aa
        self.child[tt, par, k] = new
        cnt = feasible.sum(1)
        self.nchild[self._rows(), node] += cnt
        self.n_nodes += cnt

    # ---------------------------------------------------------------- one iteration
    def _system_exclude(self):
        """_get_system_state (:182-194): one weighted-random best rollout per communicating robot; the union of their
        observed cells is what mcts_reward subtracts (ig_mcts.py:246-253)."""
        if not self.comms:
            return None
        out = np.zeros((self.T, 128), np.int32)
        for dist, valid, draw in self.comms:
            n = dist.q.shape[1]
            if self.reference_order:
                pick = np.array([np.random.choice(n, p=dist.q[t]) if (valid[t] or draw[t]) else 0 for t in range(self.T)])
            else:
                u = self.rng.random(self.T)[:, None]
                pick = np.minimum((u > np.cumsum(dist.q, 1)).sum(1), n - 1)
            if valid.any():
                out |= np.where(valid[:, None], dist.masks[self._rows(), pick], 0)
        return out

    def grow(self, nsims=10, gamma=0.9):
        """Tree.grow (:273-360) for every tree.  Returns the average simulation reward [T]."""
        T = self.T
        r = self._rows()
        exclude = self._system_exclude()                 # sampled first, like the reference
        # SELECTION: descend by UCT while the node has children (:287-288, _select :140-153, _UCT :14-18)
        node = np.zeros(T, np.int64)
        while True:
            nc = self.nchild[r, node]
            go = nc > 0
            if not go.any():
                break
            ch = self.child[r, node]                                          # [T,9], -1 = none
            chs = np.where(ch >= 0, ch, 0)
            mu_j, n_j = self.mu[r[:, None], chs], self.N[r[:, None], chs]
            n_p = self.N[r, node][:, None]
            with np.errstate(divide="ignore", invalid="ignore"):
                uct = mu_j + 2.0 * self.c_p * (2.0 * np.log(n_p) / n_j) ** 0.5
            uct = np.where(n_j == 0, np.inf, uct)
            uct = np.where(ch >= 0, uct, -np.inf)
            best = ch[r, np.argmax(uct, 1)]                                   # first maximum, like np.argmax on the list
            node = np.where(go, best, node)
        # EXPANSION + SIMULATION: one launch
        seqs, feasible, rewards = self._rollouts(node, nsims, exclude)
        self._expand(node, feasible)
        avg = self._seq_mean(rewards)
        k_best = np.argmax(rewards, 1)                                        # first strict maximum (:330-332)
        best_reward = rewards[r, k_best]
        best_seq = seqs[r, k_best]
        self.mu[r, node] = avg
        self.best_reward[r, node] = best_reward
        self.N[r, node] = 1.0
        self.best_seq[r, node] = best_seq
        self.has_rollout[r, node] = True
        # BACK-PROPAGATION (:341-357)
        cur = node.copy()
        while True:
            up = self.parent[r, cur]
            go = up >= 0
            if not go.any():
                break
            t, u = r[go], up[go]
            n_old = self.N[t, u]
            self.mu[t, u] = (gamma * self.mu[t, u] * n_old + avg[go]) / (n_old + 1.0)
            self.N[t, u] = gamma * n_old + 1.0
            better = best_reward[go] > self.best_reward[t, u]
            tb, ub = t[better], u[better]
            self.best_reward[tb, ub] = best_reward[go][better]
            self.best_seq[tb, ub] = best_seq[go][better]
            self.has_rollout[tb, ub] = True
            cur = np.where(go, up, cur)
        self.last_node = node
        return avg

    @staticmethod
    def _seq_mean(rewards):
        """avg_reward accumulated left to right then divided (:323, :334), not a pairwise sum"""
        acc = np.zeros(rewards.shape[0])
        for i in range(rewards.shape[1]):
            acc = acc + rewards[:, i]
        return acc / rewards.shape[1]

    # ---------------------------------------------------------------- results and communication
    def distribution(self, with_masks=None):
        """_update_distribution (:162-180): the comm_n non-root nodes of highest mu (ties: lowest node id, as the
        stable reversed sort gives), those that hold a best rollout, weighted by mu^2.  ``with_masks`` = an evaluator
        returning observed-cell masks for [T,n,H] sequences (needed only when the distribution is communicated)."""
        T, n = self.T, self.comm_n
        r = self._rows()
        top, mu_top, has_top, seq_top = self._top_nodes(n)
        ok = np.isfinite(mu_top) & has_top
        seqs = np.where(ok[:, :, None], seq_top, END).astype(np.uint8)
        q = np.where(ok, mu_top ** 2, 0.0)
        # compact the kept entries to the front (the reference builds X, q from the kept nodes only)
        order = np.argsort(~ok, axis=1, kind="stable")
        seqs, q, top = seqs[r[:, None], order], q[r[:, None], order], top[r[:, None], order]
        d = ActionDistribution(seqs, q, with_masks(seqs) if with_masks is not None else None)
        d.nodes = np.where(np.take_along_axis(ok, order, 1), top, -1)
        return d

    def _top_nodes(self, n):
        """-> (node ids [T,n], their mu (-inf = no such node), has_rollout, best_seq [T,n,H]): highest mu first, ties by
        lowest node id, the root excluded."""
        r = self._rows()
        mu = np.where(np.arange(self.M)[None, :] < self.n_nodes[:, None], self.mu, -np.inf)
        mu[:, 0] = -np.inf
        top = np.argsort(-mu, axis=1, kind="stable")[:, :n]                   # [T,n]
        return top, mu[r[:, None], top], self.has_rollout[r[:, None], top], self.best_seq[r[:, None], top]

    def best_action_index(self):
        """ig_mcts.find_next_action's result (ig_mcts.py:106-109): the first action of the best rollout of the
        highest-mu node = the first step of that node's path.  -1 where the distribution is empty (the reference
        raises IndexError there)."""
        d = self.distribution()
        a = d.seqs[:, 0, 0].astype(np.int64)
        return np.where(d.n > 0, a, -1)

    def receive_comms(self, dist, valid=None, draw_only=None):
        """receive_comms (:367-380) for every tree at once: ``dist`` holds, per tree, ONE other robot's distribution
        (with masks); ``valid`` [T] marks the trees that do have such a robot.  Call once per communicating robot.
        ``draw_only`` [T]: entries whose sampled path is discarded but whose random draw still happens (a tree that
        was handed its own distribution: _get_system_state samples it, then overwrites it, :192-193)."""
        valid = np.ones(self.T, bool) if valid is None else np.asarray(valid, bool)
        draw = np.zeros(self.T, bool) if draw_only is None else np.asarray(draw_only, bool)
        self.comms.append((dist, valid & (dist.n > 0), draw & (dist.n > 0)))


class DeviceDecMCTS(BatchedDecMCTS):
    """The same planner with the node arrays resident on the GPU: one grow iteration = casim_mcts_select ->
    casim_explore_rollouts -> casim_mcts_update on the current stream, no host round trip (csrc/casim_mcts.cu).  The random
    draws (simulation actions, one uniform per communicating robot) are generated on the host in the host planner's
    order and uploaded once per ``grow_n`` call, so a tree grown here equals the tree ``BatchedDecMCTS`` grows from the
    same seed, node for node.  ``select`` [E,A] bool marks the agents that plan (one tree each, (env, agent) order)."""

    ARRAYS = ("parent", "stage", "mu", "N", "best_reward", "child", "nchild", "path", "has_rollout", "best_seq")

    def __init__(self, sim, select, horizon=10, c_p=1.0, comm_n=5, seed=0, xdt=1, nsims=10, n_grows=100, device_draws=True):
        import torch
        from . import lib as _lib
        self._torch, self._lib = torch, _lib
        self.sim, self.xdt, self.nsims = sim, int(xdt), int(nsims)
        te, ta = np.nonzero(np.asarray(select, bool))
        self.te, self.ta = te, ta
        self.T, self.H, self.c_p, self.comm_n = len(te), int(horizon), float(c_p), int(comm_n)
        if self.T == 0:
            raise _lib.CasimError("DeviceDecMCTS: no agent selected")
        if getattr(sim, "belief", None) is None:
            raise _lib.CasimError("DeviceDecMCTS needs sim.enable_exploration() first")
        self.reference_order = False
        self.rng = np.random.default_rng(seed)
        # simulation actions: drawn inside casim_mcts_select (counter-based hash, = hash_draws) or, for a tree that must
        # equal the host planner's default stream, on the host with self.rng and uploaded
        self.device_draws = bool(device_draws)
        self.seed64, self.iteration = draw_seed(seed), 0
        self.comms = []
        self.launches = 0
        self.K = N_ACTIONS + self.nsims
        dev, T, H, K = sim.device, self.T, self.H, self.K
        self.M = 0
        self.d = dict(slot=torch.as_tensor(te * sim.A + ta, dtype=torch.int32, device=dev),
                      n_nodes=torch.ones(T, dtype=torch.int32, device=dev),
                      node=torch.zeros(T, dtype=torch.int32, device=dev),
                      overflow=torch.zeros(1, dtype=torch.int32, device=dev))
        self._alloc(1 + N_ACTIONS * (int(n_grows) + 1))
        self.actions = torch.full((sim.E, sim.A, K, H), END, dtype=torch.uint8, device=dev)
        self.rewards = torch.zeros((sim.E, sim.A, K), dtype=torch.float32, device=dev)
        self.n_feasible = torch.zeros((sim.E, sim.A, K), dtype=torch.int32, device=dev)
        self.exclude = None
        self.avg = torch.zeros(T, dtype=torch.float64, device=dev)
        # mcts_reward's per-cell term on the (fixed) belief, tabulated once per grow_n: casim_explore_mi_table
        self.mi_table = torch.zeros(tuple(sim.belief.shape), dtype=torch.int64, device=dev)
        self._table_ready = False
        self._comm_dev = None
        self._host_valid = False
        # root (Tree.__init__, :120-133): _expansion(root) from the feasibility of its 9 one-step extensions
        self._iteration(None, None, 0.0, root_only=True)

    # ---------------------------------------------------------------- storage
    def _alloc(self, M):
        torch, dev, T, H = self._torch, self.sim.device, self.T, self.H
        spec = dict(parent=((T, M), torch.int32, -1), child=((T, M, N_ACTIONS), torch.int32, -1),
                    nchild=((T, M), torch.uint8, 0), stage=((T, M), torch.uint8, 0), has_rollout=((T, M), torch.uint8, 0),
                    mu=((T, M), torch.float64, 0), N=((T, M), torch.float64, 0), best_reward=((T, M), torch.float64, 0),
                    path=((T, M, H), torch.uint8, END), best_seq=((T, M, H), torch.uint8, END))
        for k, (shape, dt, fill) in spec.items():
            new = torch.full(shape, fill, dtype=dt, device=dev)
            if self.M:
                new[:, :self.M] = self.d[k]
            self.d[k] = new
        self.M = M

    def reserve(self, n_grows):
        self._torch.cuda.current_stream().synchronize()
        need = int(self.d["n_nodes"].max().item()) + N_ACTIONS * int(n_grows)
        if need > self.M:
            self._alloc(need)

    def _struct(self, gamma):
        L, d = self._lib, self.d
        s = L.MctsTrees()
        s.T, s.M, s.H, s.nsims, s.comm_n = self.T, self.M, self.H, self.nsims, self.comm_n
        s.c_p, s.gamma = self.c_p, float(gamma)
        for k in ("slot", "n_nodes", "node", "overflow") + self.ARRAYS:
            setattr(s, k, d[k].data_ptr())
        if self._comm_dev is not None:
            q, masks, valid = self._comm_dev
            s.n_comms, s.comm_q, s.comm_masks, s.comm_valid = q.shape[0], q.data_ptr(), masks.data_ptr(), valid.data_ptr()
        return s

    def _upload_comms(self):
        """receive_comms entries -> [C,T,n] weights, [C,T,n,128] masks, [C,T] valid on the device"""
        torch, dev = self._torch, self.sim.device
        if not self.comms:
            self._comm_dev, self.exclude = None, None
            return
        for dist, _, _ in self.comms:
            if dist.q.shape != (self.T, self.comm_n) or dist.masks is None:
                raise self._lib.CasimError("DeviceDecMCTS: a received distribution needs [T, comm_n] weights and masks")
        q = np.stack([np.ascontiguousarray(dist.q, np.float64) for dist, _, _ in self.comms])
        masks = np.stack([np.ascontiguousarray(dist.masks, np.int32) for dist, _, _ in self.comms])
        valid = np.stack([np.asarray(v, np.uint8) for _, v, _ in self.comms])
        self._comm_dev = (torch.from_numpy(q).to(dev), torch.from_numpy(masks).to(dev), torch.from_numpy(valid).to(dev))
        if self.exclude is None:
            self.exclude = torch.zeros((self.sim.E, self.sim.A, 128), dtype=torch.int32, device=dev)

    # ---------------------------------------------------------------- one iteration, all on the device
    def _iteration(self, draws, comm_u, gamma, root_only=False):
        L, sim = self._lib, self.sim
        stream = ctypes_stream(self._torch)
        s = self._struct(gamma)
        ex = None if (root_only or self.exclude is None) else self.exclude
        L.check(sim.lib.casim_mcts_select(sim.handle, ctypes_byref(s), L.ptr(draws), self.seed64, self.iteration,
                                          L.ptr(comm_u), L.ptr(self.actions), L.ptr(ex), int(root_only), stream),
                "casim_mcts_select")
        if not root_only and draws is None:
            self.iteration += 1
        use_table = self._table_ready and not root_only
        if use_table:
            L.check(sim.lib.casim_explore_set_mi_table(sim.handle, L.ptr(self.mi_table)), "casim_explore_set_mi_table")
        # only the trees' agent slots get rollout CTAs (the other explorers of an environment plan in their own batch)
        L.check(sim.lib.casim_explore_set_rollout_slots(sim.handle, L.ptr(self.d["slot"]), self.T), "casim_explore_set_rollout_slots")
        try:
            L.check(sim.lib.casim_explore_rollouts(sim.handle, L.ptr(sim.state), L.ptr(sim.belief), L.ptr(self.actions), self.K,
                                                   self.H, self.xdt, L.ptr(ex), 0, L.ptr(self.rewards), None, None,
                                                   L.ptr(self.n_feasible), stream), "casim_explore_rollouts")
        finally:                                 # the launch has taken the pointers: no other caller sees a stale table / list
            L.check(sim.lib.casim_explore_set_rollout_slots(sim.handle, None, 0), "casim_explore_set_rollout_slots")
            if use_table:
                L.check(sim.lib.casim_explore_set_mi_table(sim.handle, None), "casim_explore_set_mi_table")
        L.check(sim.lib.casim_mcts_update(sim.handle, ctypes_byref(s), L.ptr(self.actions), L.ptr(self.rewards),
                                          L.ptr(self.n_feasible), L.ptr(self.avg), int(root_only), stream), "casim_mcts_update")
        self.launches += 1
        self._host_valid = False

    def grow_n(self, n, nsims=None, gamma=0.9):
        """``n`` iterations of Tree.grow (:273-360) for every tree, without touching the host in between."""
        torch, dev, T = self._torch, self.sim.device, self.T
        if nsims is not None and int(nsims) != self.nsims:
            raise self._lib.CasimError("DeviceDecMCTS was built for nsims = %d" % self.nsims)
        self._upload_comms()
        self._lib.check(self.sim.lib.casim_explore_mi_table(self.sim.handle, self._lib.ptr(self.sim.state),
                                                            self._lib.ptr(self.sim.belief), self._lib.ptr(self.mi_table),
                                                            ctypes_stream(torch)), "casim_explore_mi_table")
        self._table_ready = True
        C = len(self.comms)
        us = np.zeros((n, max(C, 1), T))
        draws = None if self.device_draws else np.empty((n, T, self.nsims, self.H), np.uint8)
        for it in range(n):                              # the host planner's order: _system_exclude, then _rollouts
            for c in range(C):
                us[it, c] = self.rng.random(T)
            if draws is not None:
                draws[it] = self.rng.integers(0, N_ACTIONS, (T, self.nsims, self.H), dtype=np.uint8)
        d_us = torch.from_numpy(us).to(dev) if C else None
        d_draws = None if draws is None else torch.from_numpy(draws).to(dev)
        for it in range(n):
            self._iteration(None if d_draws is None else d_draws[it], d_us[it] if C else None, gamma)
        self._keep = (d_us, d_draws)                     # alive until the stream has consumed them

    def grow(self, nsims=10, gamma=0.9):
        self.grow_n(1, nsims, gamma)
        self._torch.cuda.current_stream().synchronize()
        return self.avg.cpu().numpy()

    # ---------------------------------------------------------------- results
    def sync_host(self):
        """Node arrays -> the numpy attributes of the host planner (distribution / best_action_index read them)."""
        if self._host_valid:
            return
        self._torch.cuda.current_stream().synchronize()
        if int(self.d["overflow"].item()):
            raise self._lib.CasimError("DeviceDecMCTS: node arrays full (M = %d); reserve() more iterations" % self.M)
        self.n_nodes = self.d["n_nodes"].cpu().numpy().astype(np.int64)
        used = int(self.n_nodes.max())
        for k in self.ARRAYS:
            a = self.d[k][:, :used].cpu().numpy()
            if k in ("parent", "child", "stage", "nchild"):
                a = a.astype(np.int64)
            elif k == "has_rollout":
                a = a.astype(bool)
            setattr(self, k, a)
        self.last_node = self.d["node"].cpu().numpy().astype(np.int64)
        self._host_valid = True

    def _top_nodes(self, n):
        """The same ranking with a stable sort on the device; only the n winners per tree come back."""
        torch, d = self._torch, self.d
        used = int(d["n_nodes"].max().item())                                 # (synchronises the stream)
        if int(d["overflow"].item()):
            raise self._lib.CasimError("DeviceDecMCTS: node arrays full (M = %d); reserve() more iterations" % self.M)
        ids = torch.arange(used, device=self.sim.device)[None, :]
        mu = torch.where(ids < d["n_nodes"][:, None], d["mu"][:, :used], float("-inf"))
        mu[:, 0] = float("-inf")
        top = torch.argsort(-mu, dim=1, stable=True)[:, :n]
        if top.shape[1] < n:                                                  # fewer nodes than entries: pad with the root
            top = torch.cat([top, torch.zeros((self.T, n - top.shape[1]), dtype=top.dtype, device=top.device)], 1)
        mu_top = torch.gather(mu, 1, top)
        has = torch.gather(d["has_rollout"][:, :used], 1, top) != 0
        seq = torch.gather(d["best_seq"][:, :used], 1, top[:, :, None].expand(-1, -1, self.H))
        return top.cpu().numpy(), mu_top.cpu().numpy(), has.cpu().numpy(), seq.cpu().numpy()


def ctypes_stream(torch):
    import ctypes
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ctypes_byref(s):
    import ctypes
    return ctypes.byref(s)


def sim_evaluator(sim, xdt=1, select=None):
    """The GPU evaluator: trees = the explorer agents of ``sim`` (policy id IG_GREEDY) in (env, agent) order, or the
    agents marked in ``select`` [E,A] bool.  Returns (n_trees, evaluate, masks_of, (env index, agent index))."""
    import torch
    from . import layout as L
    if select is None:
        types = sim._keep["types"].cpu().numpy()
        select = (types & 0xff) == L.POLICY_IG_GREEDY
    te, ta = np.nonzero(select)
    E, A = sim.E, sim.A

    def scatter(x, fill, dtype):
        full = np.full((E, A) + x.shape[1:], fill, dtype)
        full[te, ta] = x
        return full

    def evaluate(actions, exclude):
        ex = None if exclude is None else scatter(exclude, 0, np.int32)
        out = sim.explore_rollouts(scatter(actions, END, np.uint8), exclude=ex, xdt=xdt, want_end_pose=False,
                                   want_obsvd=False)
        torch.cuda.current_stream().synchronize()
        return out["rewards"].cpu().numpy()[te, ta].astype(np.float64), out["n_feasible"].cpu().numpy()[te, ta]

    def masks_of(seqs):
        out = sim.explore_rollouts(scatter(seqs, END, np.uint8), xdt=xdt, want_end_pose=False, want_obsvd=True)
        return out["obsvd"].cpu().numpy()[te, ta]

    return len(te), evaluate, masks_of, (te, ta)


class TeamPlanner:
    """ig_mcts.find_next_action for every explorer of a batch, sequenced like collision_avoidance_env._take_action_dmcts
    (:341-379): ``Ncycles`` cycles; in each cycle the explorers of an environment plan one after the other (so the r-th
    explorer of every environment forms one batch of trees), each receiving the others' latest action distributions --
    this cycle's from the explorers that planned before it, the previous cycle's / step's from the others -- and growing
    its tree ``Ntree`` times.  Cycle 0 of a step starts a new tree (new_step, ig_mcts.py:89-93).
    Which distributions an agent receives follows ig_mcts.py:96-98 literally: ``for j in range(len(dmcts_agents)):
    agents[j].policy.best_paths`` -- agents 0 .. n_explorers-2 (sic: an index into the list is used as an agent index),
    of which its own entry is overwritten by its own state in _get_system_state; explorers are expected first in the
    agent list, as in the reference's scenarios."""

    def __init__(self, sim, params):
        from . import layout as L
        self.sim, self.p = sim, params
        types = sim._keep["types"].cpu().numpy()
        self.ig = (types & 0xff) == L.POLICY_IG_GREEDY
        self.rank = np.where(self.ig, np.cumsum(self.ig, 1) - 1, -1)
        self.n_ig = self.ig.sum(1)
        self.R = int(self.n_ig.max()) if self.ig.any() else 0
        E, H, n = sim.E, params.mcts_horizon, 5
        self.best = [dict(seqs=np.full((E, n, H), END, np.uint8), q=np.zeros((E, n)), masks=np.zeros((E, n, 128), np.int32))
                     for _ in range(self.R)]
        self.trees = [None] * self.R
        self.steps = 0

    def reset(self):
        for b in self.best:
            b["seqs"][:] = END
            b["q"][:] = 0
            b["masks"][:] = 0
        self.steps = 0

    def reset_envs(self, mask):
        """Forget the communicated action distributions of the environments that start a new episode (mask [E] bool)."""
        for b in self.best:
            b["seqs"][mask] = END
            b["q"][mask] = 0
            b["masks"][mask] = 0

    def plan(self, done):
        """done [E,A] bool (agents that are not asked for an action, collision_avoidance_env.py:298-299).
        Returns float32 [E,A,2] actions ([vel, dphi]; zeros for everyone who did not plan)."""
        sim, p = self.sim, self.p
        actions = np.zeros((sim.E, sim.A, 2), np.float32)
        for cycle in range(p.Ncycles):
            for r in range(self.R):
                sel = self.ig & (self.rank == r) & ~done
                if not sel.any():
                    continue
                T, ev, masks_of, (te, ta) = sim_evaluator(sim, p.xDT, select=sel)
                if cycle == 0:
                    ref_order = getattr(p, "reference_order", False)
                    if getattr(p, "device_trees", True) and not ref_order:
                        tree = DeviceDecMCTS(sim, sel, horizon=p.mcts_horizon, c_p=p.mcts_cp, comm_n=5,
                                             seed=(p.seed, self.steps, r), xdt=p.xDT, nsims=p.Nsims,
                                             n_grows=p.Ntree * p.Ncycles)
                    else:
                        tree = BatchedDecMCTS(T, ev, horizon=p.mcts_horizon, c_p=p.mcts_cp, comm_n=5,
                                              seed=(p.seed, self.steps, r), reference_order=ref_order)
                        tree.reserve(p.Ntree * p.Ncycles)
                    self.trees[r] = (tree, te)
                else:
                    tree, te_old = self.trees[r]
                    assert np.array_equal(te_old, te)
                    tree.comms = []                      # receive_comms replaces the entry of the same robot (:379)
                for j in range(self.R - 1):             # sic, see the class docstring
                    b = self.best[j]
                    dist = ActionDistribution(b["seqs"][te], b["q"][te], b["masks"][te])
                    listed = j < self.n_ig[te] - 1
                    if j == r:                          # its own entry: sampled, then overwritten by its own state
                        tree.receive_comms(dist, valid=np.zeros(T, bool), draw_only=listed)
                    else:
                        tree.receive_comms(dist, valid=listed)
                if isinstance(tree, DeviceDecMCTS):
                    tree.grow_n(p.Ntree, nsims=p.Nsims, gamma=p.mcts_gamma)
                else:
                    for _ in range(p.Ntree):
                        tree.grow(nsims=p.Nsims, gamma=p.mcts_gamma)
                d = tree.distribution(with_masks=masks_of)
                self.best[r]["seqs"][te], self.best[r]["q"][te], self.best[r]["masks"][te] = d.seqs, d.q, d.masks
                a = tree.best_action_index()
                actions[te, ta] = np.where((a >= 0)[:, None], action_value(np.maximum(a, 0)), 0.0)
        self.steps += 1
        return actions
