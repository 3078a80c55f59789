"""TEST INFRASTRUCTURE -- generates tests/golden/explore.npz by running the UNMODIFIED reference classes
edfMap / targetMap / ig_greedy (gym_collision_avoidance/envs/information_models/, policies/ig_greedy.py)
in the build container.  Run through ``python oracle/gen_golden.py explore``.

The one shim: ig_greedy.find_next_action calls targetMap.update(pose, targets) with a bare pose
(ig_greedy.py:39) which raises once a target is detected; the harness wraps update() so that it receives
([pose], [targets]) -- the form ig_mcts.update_belief uses (ig_mcts.py:117-133).  No reference source is edited.
"""
import os
import types

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(_HERE), "tests", "golden")

CROSSING = [[2, 10, 10, 2], [-10, 10, -2, 2], [2, -2, 10, -10], [-10, -2, -2, -10]]
HALLWAY = [[3, 7, 10, -7], [-10, 7, -3, -7], [-10.5, 10, -10, -10], [10.5, 10, 10, -10]]
# not symmetric in y: exposes the vertical mirroring between Map.static_map and edfMap indexing
ASYM = [[-12, 12, -4, 9], [1, 6, 3, -1], [5, -3, 12, -5], [-9, -4, -6, -12], [-2, 13, 9, 11.5]]
MAPS = [CROSSING, HALLWAY, ASYM]


def obstacles_of(rects):
    """rect (x1, y1, x3, y3) -> the reference's obstacle format: 4 corners of which Map.get_occupancy_grid reads
    index 1 (top-left) and index 3 (bottom-right) (Map.py:117-118, envs/test_cases.py:3081-3084)"""
    return [[(r[2], r[1]), (r[0], r[1]), (r[0], r[3]), (r[2], r[3])] for r in rects]


def make_policy(ns, ref_map, radius):
    pol = ns.ig_greedy()
    ego = types.SimpleNamespace(radius=float(radius))
    pol.init_maps(ego, ref_map, (30, 30), 0.1, 60.0, 5.0, dt=0.1)
    orig = pol.targetMap.update
    pol.targetMap.update = lambda pose, targets, frame='global': orig([pose], [targets], frame=frame)
    return pol


def free_pose(rng, edf, radius):
    for _ in range(10000):
        p = np.float32(rng.uniform(-13.5, 13.5, 2)).astype(np.float64)
        if edf.get_edf_value_from_pose(p) > radius + 0.05:
            return p
    raise RuntimeError("no free pose")


def main(ns, ncases=48, seed=2024):
    rng = np.random.default_rng(seed)
    ref_maps = [ns.Map(30, 30, 0.1, obstacles_of(r)) for r in MAPS]
    out = {"rects_%d" % k: np.array(r, np.float64) for k, r in enumerate(MAPS)}
    for k, m in enumerate(ref_maps):
        out["static_%d" % k] = np.packbits(m.static_map.astype(np.uint8))
        e = ns.edfMap(m, 0.1, (30, 30))
        d2 = np.rint((e.map / 0.1) ** 2).astype(np.int32)
        assert np.array_equal(np.sqrt(d2.astype(np.float64)) * 0.1, e.map)      # EDT == sqrt(int) * cell, exactly
        out["edf_d2_%d" % k] = d2
    rec = {k: [] for k in ("map_id", "pose", "radius", "oas", "belief0", "vis", "occ", "action", "cand_mi", "best_mi",
                           "n_targets")}
    for case in range(ncases):
        mid = case % len(MAPS)
        radius = float(np.float32(rng.choice([0.2, 0.35, 0.5])))
        pol = make_policy(ns, ref_maps[mid], radius)
        p = free_pose(rng, pol.edfMap, radius)
        heading = float(np.float32(rng.uniform(-np.pi, np.pi)))
        if case % 7 == 3:       # hug a wall so that some candidate next poses are infeasible
            for _ in range(200):
                q = free_pose(rng, pol.edfMap, radius)
                if pol.edfMap.get_edf_value_from_pose(q) < radius + 0.3:
                    p = q
                    break
        # other_agents_states rows: zero-padded [9,10]; a few Static (type 1) and dynamic (type 2) agents
        oas = np.zeros((9, 10))
        n_rows = int(rng.integers(0, 5))
        for r in range(n_rows):
            if rng.random() < 0.7:   # in front of the agent, often inside FOV / range
                d, a = rng.uniform(0.5, 6.0), heading + rng.uniform(-0.7, 0.7)
            else:
                d, a = rng.uniform(0.5, 9.0), rng.uniform(-np.pi, np.pi)
            oas[r, 0], oas[r, 1] = np.float32(d * np.cos(a)), np.float32(d * np.sin(a))
            oas[r, 2:9] = np.float32(rng.uniform(-1, 1, 7))
            oas[r, 9] = 1.0 if rng.random() < 0.7 else 2.0
        # non-trivial belief: products of rOcc / rEmp powers, rounded to fp32 (the GPU stores fp32 odds)
        a = rng.integers(0, 4, (60, 60)) * (rng.random((60, 60)) < 0.3)
        b = rng.integers(0, 6, (60, 60)) * (rng.random((60, 60)) < 0.5)
        belief0 = np.float32(1.5 ** a * 0.66 ** b)
        if case % 5 == 0:
            belief0[:] = 1.0
        pol.targetMap.map = belief0.astype(np.float64).copy()
        obs = {"pos_global_frame": p.copy(), "heading_global_frame": heading, "other_agents_states": oas.copy()}
        pose = np.append(p, heading)
        targets = pol.find_targets_in_obs(obs)
        action = pol.find_next_action(obs, None, 0, None)
        after = pol.targetMap.map
        ratio = after / belief0.astype(np.float64)
        vis = ~np.isclose(ratio, 1.0)
        occ = np.isclose(ratio, 1.5)
        assert np.all(np.isclose(ratio[vis & ~occ], 0.66))
        cand = []
        for v in (0.0, 2.0, 4.0):
            for w in (-np.pi, 0, np.pi):
                nxt = pol.get_next_pose(pose, np.array([v, w]))
                cand.append(-1.0 if nxt is None else pol.targetMap.get_reward_from_pose(nxt))
        cand = np.array(cand)
        if isinstance(action, np.ndarray):
            act = action.astype(np.float64)
            best = cand.max()
        else:
            act = np.array([-1.0, -1.0])
            best = -1.0
        rec["map_id"].append(mid); rec["pose"].append(pose); rec["radius"].append(radius); rec["oas"].append(oas)
        rec["belief0"].append(belief0); rec["vis"].append(np.packbits(vis)); rec["occ"].append(np.packbits(occ))
        rec["action"].append(act); rec["cand_mi"].append(cand); rec["best_mi"].append(best)
        rec["n_targets"].append(len(targets))
    for k, v in rec.items():
        out[k] = np.array(v)
    # SURVEY Appendix B known answers (crossing map, pose (-5, 0, 0))
    pol = make_policy(ns, ref_maps[0], 0.5)
    pose = np.array([-5.0, 0.0, 0.0])
    cells = pol.targetMap.getVisibleCells(pose)
    out["kb_num_visible"] = np.array(len(cells))
    out["kb_reward_prior"] = np.array(pol.targetMap.get_reward_from_pose(pose))
    np.savez_compressed(os.path.join(GOLDEN, "explore.npz"), **out)
    print("wrote explore.npz: %d cases, targets detected in %d, infeasible candidates in %d, kb %d cells %.16f" % (
        ncases, sum(1 for n in rec["n_targets"] if n > 0), sum(1 for c in rec["cand_mi"] if (c < 0).any()),
        len(cells), float(out["kb_reward_prior"])))


def main_rollouts(ns, ncases=24, K=6, H=10, seed=777):
    """tests/golden/rollouts.npz: Dec-MCTS simulation rollouts through the UNMODIFIED reference --
    pydecmcts.DecMCTS.Tree._simulate from the root node with ig_mcts's own callbacks
    (mcts_sim_state_storer / get_next_pose / mcts_reward, ig_mcts.py:173-265).  The one substitution is the
    rollout's random choice (mcts_sim_selection_func, :255-257), replaced by a scripted choice so that the
    sequence is known."""
    from gym_collision_avoidance.envs.policies.ig_mcts import ig_mcts, MCTS_state
    from gym_collision_avoidance.envs.policies.pydecmcts.DecMCTS import Tree
    rng = np.random.default_rng(seed)
    ref_maps = [ns.Map(30, 30, 0.1, obstacles_of(r)) for r in MAPS]
    out = {"static_%d" % k: np.packbits(m.static_map.astype(np.uint8)) for k, m in enumerate(ref_maps)}
    rec = {k: [] for k in ("map_id", "pose", "radius", "belief", "actions", "exclude", "rewards", "end_pose", "obsvd",
                           "n_infeasible")}
    for case in range(ncases):
        mid = case % len(MAPS)
        radius = float(np.float32(rng.choice([0.2, 0.35, 0.5])))
        pol = ig_mcts()
        ego = types.SimpleNamespace(radius=radius)
        pol.set_param(ego, ref_maps[mid], (30, 30), 0.1, 60.0, 5.0, dt=0.1, xdt=1, mcts_horizon=H)
        p = free_pose(rng, pol.edfMap, radius)
        if case % 3 == 1:      # start next to a wall or the map border: infeasible steps
            for _ in range(400):
                q = free_pose(rng, pol.edfMap, radius)
                if pol.edfMap.get_edf_value_from_pose(q) < radius + 0.4 or np.abs(q).max() > 13.0:
                    p = q
                    break
        heading = float(np.float32(rng.uniform(-np.pi, np.pi)))
        pose = np.append(p, heading)
        a = rng.integers(0, 4, (60, 60)) * (rng.random((60, 60)) < 0.3)
        b = rng.integers(0, 6, (60, 60)) * (rng.random((60, 60)) < 0.5)
        belief = np.float32(1.5 ** a * 0.66 ** b)
        if case % 5 == 0:
            belief[:] = 1.0
        pol.targetMap.map = belief.astype(np.float64).copy()
        # the other robots' claimed cells: what they would see from a pose near ours (or nothing)
        other_cells = set()
        if case % 2 == 0:
            op = np.append(p + rng.uniform(-2, 2, 2), heading + rng.uniform(-1, 1))
            other_cells = set(pol.targetMap.getVisibleCells(op))
        exclude = np.zeros((60, 60), bool)
        for (i, j) in other_cells:
            exclude[j, i] = True
        acts = rng.integers(0, 9, (K, H)).astype(np.uint8)
        acts[0, :] = 7                                   # straight ahead at full speed until something stops it
        acts[1, :] = rng.choice([6, 7, 8], H)            # always moving
        if case % 4 == 2:
            acts[2, H // 2:] = 255                       # a shorter sequence (a node rollout cut by the horizon)
        options = ig_mcts.mcts_avail_actions(None, None, 0)
        rewards, ends, obs_k, ninf = [], [], [], 0
        for k in range(K):
            seq = [int(v) for v in acts[k] if v != 255]
            sel = lambda data, opts, st, seq=seq: opts[seq[len(st.action_seq)]]
            tree = Tree({"current_pose": pose}, pol.mcts_reward, pol.mcts_avail_actions, pol.mcts_state_storer, sel,
                        pol.mcts_avail_actions, pol.mcts_sim_state_storer, 5, robot_id=0, horizon=len(seq))
            res = tree._simulate(1, {1: MCTS_state([], [], set(), set(other_cells), 0)})
            st = res["temp_state"]
            assert len(st.action_seq) == len(seq)
            ninf += sum(1 for t, aa in enumerate(st.action_seq)
                        if options[seq[t]][0] != 0.0 and aa[0] == 0.0)
            m = np.zeros((60, 60), bool)
            for (i, j) in st.obsvd_cells:
                m[j, i] = True
            rewards.append(float(res["reward"])); ends.append(np.asarray(st.pose_seq[-1], np.float64))
            obs_k.append(np.packbits(m))
        rec["map_id"].append(mid); rec["pose"].append(pose); rec["radius"].append(radius); rec["belief"].append(belief)
        rec["actions"].append(acts); rec["exclude"].append(np.packbits(exclude)); rec["rewards"].append(rewards)
        rec["end_pose"].append(ends); rec["obsvd"].append(obs_k); rec["n_infeasible"].append(ninf)
    for k, v in rec.items():
        out[k] = np.array(v)
    np.savez_compressed(os.path.join(GOLDEN, "rollouts.npz"), **out)
    print("wrote rollouts.npz: %d cases x %d rollouts x %d steps, %d infeasible steps, reward range %.3f..%.3f" % (
        ncases, K, H, int(np.sum(rec["n_infeasible"])), np.min(rec["rewards"]), np.max(rec["rewards"])))


def main_mcts(ns, seed=4242, H=6, nsims=4, ntree=40, gamma=0.9):
    """tests/golden/mcts.npz: two Dec-MCTS trees grown by the UNMODIFIED reference (pydecmcts Tree + ig_mcts callbacks,
    np.random seeded): tree A alone; tree B after receiving A's action distribution (receive_comms).  Recorded per
    tree: every node's parent / stage / action path / mu / N, the per-iteration average reward and the final action."""
    from gym_collision_avoidance.envs.policies.ig_mcts import ig_mcts
    from gym_collision_avoidance.envs.policies.pydecmcts.DecMCTS import Tree
    rng = np.random.default_rng(seed)
    ref_map = ns.Map(30, 30, 0.1, obstacles_of(CROSSING))
    out = {"static": np.packbits(ref_map.static_map.astype(np.uint8)), "H": H, "nsims": nsims, "ntree": ntree,
           "gamma": gamma, "seed": seed}
    options = ig_mcts.mcts_avail_actions(None, None, 0)

    def idx_of(a):
        return [k for k, o in enumerate(options) if (o == a).all()][0]

    radius = 0.5
    a = rng.integers(0, 4, (60, 60)) * (rng.random((60, 60)) < 0.3)
    b = rng.integers(0, 6, (60, 60)) * (rng.random((60, 60)) < 0.5)
    belief = np.float32(1.5 ** a * 0.66 ** b)
    poses = [np.array([-6.0, 0.5, 0.2]), np.array([-4.5, 1.2, 1.3])]        # B faces a wall: infeasible expansions
    out.update(belief=belief, poses=np.array(poses), radius=radius)
    trees = []
    for k, pose in enumerate(poses):
        pol = ig_mcts()
        pol.set_param(types.SimpleNamespace(radius=radius), ref_map, (30, 30), 0.1, 60.0, 5.0, dt=0.1, xdt=1, mcts_horizon=H)
        pol.targetMap.map = belief.astype(np.float64).copy()
        np.random.seed(seed + k)
        tree = Tree({"current_pose": pose}, pol.mcts_reward, pol.mcts_avail_actions, pol.mcts_state_storer,
                    pol.mcts_sim_selection_func, pol.mcts_avail_actions, pol.mcts_sim_state_storer, 5, robot_id=k,
                    horizon=H, c_p=1.0)
        if k == 1:
            tree.receive_comms(trees[0].send_comms(), 0)
        avgs = [tree.grow(nsims=nsims, gamma=gamma) for _ in range(ntree)]
        g = tree.graph
        n = len(g)
        parent = np.array([-1] + [list(g.predecessors(i))[0] - 1 for i in range(2, n + 1)])
        stage = np.array([g.nodes[i]["stage"] for i in range(1, n + 1)])
        mu = np.array([g.nodes[i]["mu"] for i in range(1, n + 1)])
        N = np.array([g.nodes[i]["N"] for i in range(1, n + 1)], np.float64)
        path = np.full((n, H), 255, np.uint8)
        for i in range(1, n + 1):
            for d, act in enumerate(g.nodes[i]["state"].action_seq):
                path[i - 1, d] = idx_of(act)
        dist = tree.send_comms()
        out.update({"parent_%d" % k: parent, "stage_%d" % k: stage, "mu_%d" % k: mu, "N_%d" % k: N, "path_%d" % k: path,
                    "avg_%d" % k: np.array(avgs), "q_%d" % k: np.array(dist.q),
                    "action_%d" % k: idx_of(dist.X[0].action_seq[0])})
        trees.append(tree)
        print("tree %d: %d nodes, max stage %d, avg reward %.4f -> %.4f, action %d" % (
            k, n, stage.max(), avgs[0], avgs[-1], out["action_%d" % k]))
    np.savez_compressed(os.path.join(GOLDEN, "mcts.npz"), **out)


def main_team(ns, seed=99, nsteps=3):
    """tests/golden/team.npz: ig_mcts.update_belief (ig_mcts.py:117-137) of the UNMODIFIED reference for small teams
    over several steps, called in the order collision_avoidance_env._take_action_dmcts does (ascending index, done
    agents skipped).  Scenes: 3 explorers + 2 static targets; explorer 1 is done from step 1 on in scene 1."""
    from gym_collision_avoidance.envs.policies.ig_mcts import ig_mcts
    rng = np.random.default_rng(seed)
    ref_map = ns.Map(30, 30, 0.1, obstacles_of(CROSSING))
    out = {"static": np.packbits(ref_map.static_map.astype(np.uint8))}
    nscenes, A, n_ig = 4, 5, 3
    rec = {k: [] for k in ("pose", "oas", "done", "belief_after", "team_mask", "team_reward", "n_targets", "belief0")}
    for sc in range(nscenes):
        pols = []
        a = rng.integers(0, 4, (60, 60)) * (rng.random((60, 60)) < 0.3)
        b = rng.integers(0, 6, (60, 60)) * (rng.random((60, 60)) < 0.5)
        belief0 = np.float32(1.5 ** a * 0.66 ** b)
        for i in range(n_ig):
            pol = ig_mcts()
            pol.set_param(types.SimpleNamespace(radius=0.5), ref_map, (30, 30), 0.1, 60.0, 5.0, dt=0.1, xdt=1)
            pol.targetMap.map = belief0.astype(np.float64).copy()
            pols.append(pol)
        agents = [types.SimpleNamespace(policy=pols[i] if i < n_ig else ns.StaticPolicy(), pos_global_frame=None,
                                        heading_global_frame=None) for i in range(A)]
        centre = free_pose(rng, pols[0].edfMap, 1.5)
        srec = {k: [] for k in rec}
        srec["belief0"] = belief0
        for step in range(nsteps):
            poses = np.zeros((A, 3))
            for i in range(A):       # a loose cluster so that the explorers see each other's cells and the targets
                for _ in range(1000):
                    p = np.float32(np.clip(centre + rng.uniform(-3, 3, 2), -13.5, 13.5)).astype(np.float64)
                    if pols[0].edfMap.get_edf_value_from_pose(p) > 0.6:
                        break
                poses[i, :2] = p
                poses[i, 2] = float(np.float32(rng.uniform(-np.pi, np.pi)))
                agents[i].pos_global_frame = poses[i, :2].copy()
                agents[i].heading_global_frame = poses[i, 2]
            for i in range(n_ig):    # look at a target, so that detections happen
                tgt = poses[n_ig + (i % 2), :2]
                if rng.random() < 0.8:
                    d = tgt - poses[i, :2]
                    poses[i, 2] = float(np.float32(np.arctan2(d[1], d[0]) + rng.uniform(-0.3, 0.3)))
                    agents[i].heading_global_frame = poses[i, 2]
            done = np.zeros(A, bool)
            if sc == 1 and step >= 1:
                done[1] = True
            oas = np.zeros((A, 9, 10))
            for i in range(n_ig):
                others = [j for j in range(A) if j != i]
                for r, j in enumerate(others):
                    oas[i, r, 0:2] = np.float32(poses[j, :2] - poses[i, :2])
                    oas[i, r, 2:9] = np.float32(rng.uniform(-1, 1, 7))
                    oas[i, r, 9] = 1.0 if j >= n_ig else 2.0
            ba, tm, tr, nt = [], [], [], []
            for i in range(n_ig):
                if done[i]:
                    ba.append(pols[i].targetMap.map.copy()); tm.append(np.packbits(np.zeros((60, 60), bool)))
                    tr.append(0.0); nt.append(-1)
                    continue
                obs = {"pos_global_frame": poses[i, :2].copy(), "heading_global_frame": poses[i, 2],
                       "other_agents_states": oas[i].copy()}
                dm = [j for j in range(A) if "ig_" in str(type(agents[j].policy)) and j != i]
                pols[i].update_belief(dm, np.append(poses[i, :2], poses[i, 2]), agents, obs)
                m = np.zeros((60, 60), bool)
                for (ci, cj) in pols[i].team_obsv_cells:
                    m[cj, ci] = True
                ba.append(pols[i].targetMap.map.copy()); tm.append(np.packbits(m)); tr.append(float(pols[i].team_reward))
                nt.append(len(pols[i].obsvd_targets))
            srec["pose"].append(poses); srec["oas"].append(oas); srec["done"].append(done)
            srec["belief_after"].append(ba); srec["team_mask"].append(tm); srec["team_reward"].append(tr)
            srec["n_targets"].append(nt)
        for k in rec:
            rec[k].append(srec[k])
    for k, v in rec.items():
        out[k] = np.array(v)
    out["n_ig"] = n_ig
    np.savez_compressed(os.path.join(GOLDEN, "team.npz"), **out)
    nt = out["n_targets"]
    print("wrote team.npz: %d scenes x %d steps x %d explorers; detections %d; mean team cells %.1f" % (
        nscenes, nsteps, n_ig, int(nt[nt > 0].sum()), np.mean([np.unpackbits(m).sum() for m in out["team_mask"].reshape(-1, out["team_mask"].shape[-1])])))


def main_mcts_env(ns, seed=31, nsteps=3, Ntree=6, Nsims=3, H=4, Ncycles=2):
    """tests/golden/mcts_env.npz: the UNMODIFIED reference CollisionAvoidanceEnv.step on the fork's own exploration
    scenario tc.IG_agent_crossing (test_cases.py:3209-3239: three ig_mcts explorers, two static targets, crossing map),
    planner parameters set through ig_mcts.set_param, np.random seeded before every step.  Recorded per step: every
    agent's pose and action, every explorer's target map, detections and planned first action."""
    from oracle import ref_harness as rh
    with rh.quiet():
        from gym_collision_avoidance.envs import test_cases as tc
        agents, obstacles = tc.IG_agent_crossing()
    env = rh.make_env(ns, agents, obstacles, ['dist_to_goal', 'rel_goal', 'radius', 'heading_ego_frame',
                                              'heading_global_frame', 'pos_global_frame', 'pref_speed',
                                              'other_agents_states', 'laserscan'])
    ig = [i for i, a in enumerate(agents) if "ig_mcts" in str(type(a.policy))]
    for i in ig:
        agents[i].policy.set_param(agents[i], env.map, (30, 30), 0.1, 60.0, 5.0, dt=0.1, xdt=1, Ntree=Ntree, Nsims=Nsims,
                                   mcts_horizon=H, mcts_gamma=1, Ncycles=Ncycles)
    A = len(agents)
    out = {"static": np.packbits(env.map.static_map.astype(np.uint8)), "seed": seed, "Ntree": Ntree, "Nsims": Nsims,
           "H": H, "Ncycles": Ncycles, "n_ig": len(ig),
           "init": np.array([[a.pos_global_frame[0], a.pos_global_frame[1], a.goal_global_frame[0], a.goal_global_frame[1],
                              a.radius, a.pref_speed, a.heading_global_frame] for a in agents])}
    rec = {k: [] for k in ("pose", "belief", "action", "n_targets", "rewards", "done")}
    for step in range(nsteps):
        np.random.seed(seed + step)
        obs, rewards, game_over, info = rh.step(env)
        rec["pose"].append([[a.pos_global_frame[0], a.pos_global_frame[1], a.heading_global_frame] for a in agents])
        rec["belief"].append([agents[i].policy.targetMap.map.copy() for i in ig])
        rec["action"].append([np.asarray(agents[i].policy.best_paths.X[0].action_seq[0], np.float64) for i in ig])
        rec["n_targets"].append([len(agents[i].policy.obsvd_targets) for i in ig])
        rec["rewards"].append(np.asarray(rewards, np.float64).reshape(-1)[:A])
        rec["done"].append([bool(a.is_done) for a in agents])
    for k, v in rec.items():
        out[k] = np.array(v)
    np.savez_compressed(os.path.join(GOLDEN, "mcts_env.npz"), **out)
    print("wrote mcts_env.npz: %d steps, actions %s, moved %s" % (
        nsteps, out["action"].tolist(), np.round(out["pose"][-1][:, :2] - out["init"][:, :2], 3).tolist()))
