"""GPU parity of the Dec-MCTS rollout evaluator (casim_explore_rollouts) against (a) the golden vectors recorded
from the reference's pydecmcts Tree._simulate driven with ig_mcts's callbacks and (b) the CPU oracle on seeded
batches.  Observed-cell sets and feasibility bit-exact; end poses to fp32 rounding; rewards within 1e-5 relative
(north_star)."""
import numpy as np
import pytest

import helpers as H
from casim import layout as L

pytestmark = pytest.mark.gpu


def _cfg(**kw):
    from casim import lib
    c = lib.default_config()
    c.game_over_mode = 0
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def _sim(E, A, cfg):
    from casim.sim import BatchedSim
    sim = BatchedSim(E, A, cfg=cfg, sensors=("other_agents_states",))
    sim.enable_exploration()
    return sim


def _state_of(cases):
    E = len(cases)
    state = np.zeros((E, 1, L.STATE_WORDS))
    for k, cs in enumerate(cases):
        state[k, 0, L.PX], state[k, 0, L.PY], state[k, 0, L.HEADING] = cs["pose"]
        state[k, 0, L.RADIUS] = cs["radius"]
    state[:, 0, L.PREF_SPEED] = 1.0
    state[:, 0, L.TYPE] = L.encode_type(L.POLICY_IG_GREEDY, L.DYN_FIRST_ORDER, L.SENSOR_OTHER_AGENTS)
    return state


def test_rollouts_golden_vs_reference():
    import torch
    from casim.sim import pack_cell_mask, unpack_cell_mask
    z = H.load("rollouts.npz")
    maps, cases = H.rollout_cases(z)
    cfg = _cfg()
    for mid in range(len(maps)):
        sel = [cs for cs in cases if cs["map_id"] == mid]
        E = len(sel)
        sim = _sim(E, 1, cfg)
        sim.set_map(maps[mid])
        state = _state_of(sel)
        sim.bind(types=state[:, :, L.TYPE].astype(np.int32), n_agents=np.ones((E,), np.int32))
        sim.load_state_numpy(state)
        sim.belief.copy_(torch.from_numpy(np.stack([cs["belief"] for cs in sel])[:, None]))
        actions = np.stack([cs["actions"] for cs in sel])[:, None]           # [E,1,K,H]
        exclude = pack_cell_mask(np.stack([cs["exclude"] for cs in sel])[:, None])
        out = sim.explore_rollouts(actions, exclude=exclude)
        torch.cuda.synchronize()
        obs = unpack_cell_mask(out["obsvd"].cpu().numpy(), 60)[:, 0]
        rew = out["rewards"].cpu().numpy()[:, 0]
        end = out["end_pose"].cpu().numpy()[:, 0]
        lens = (actions != 255).sum(-1)[:, 0]
        n_inf = (lens - out["n_feasible"].cpu().numpy()[:, 0]).sum(-1)
        assert np.array_equal(n_inf, [int(z["n_infeasible"][cs["index"]]) for cs in sel])
        for k, cs in enumerate(sel):
            assert np.array_equal(obs[k], cs["obsvd"]), "map %d case %d observed cells" % (mid, k)
            H.assert_close(rew[k], cs["rewards"], "map %d case %d rewards" % (mid, k), rtol=1e-5, atol=1e-7)
            H.assert_close(end[k], cs["end_pose"], "map %d case %d end pose" % (mid, k), rtol=1e-6, atol=1e-6)
        # per-rollout masks give the same answer as the shared mask
        ex_k = np.repeat(exclude[:, :, None], actions.shape[2], axis=2)
        out2 = sim.explore_rollouts(actions, exclude=ex_k)
        assert torch.equal(out2["rewards"], out["rewards"]) and torch.equal(out2["obsvd"], out["obsvd"])
        # without exclusion the reward is the MI of the whole observed set
        out3 = sim.explore_rollouts(actions)
        assert bool((out3["rewards"] >= out["rewards"] - 1e-6).all())
        sim.close()


def test_rollouts_batch_vs_oracle(oracle_lib):
    """C4 scene (2 explorers + 2 static targets per env, crossing map) after a few steps of exploration: 8
    rollouts x 10 steps per agent against the oracle, and 0 reward for the non-explorers."""
    import torch
    from casim import scenarios as S
    from casim.sim import unpack_cell_mask
    oracle = oracle_lib
    E, K, Hn = 24, 8, 10
    init, types, n, rects = S.exploration(E, seed=11)
    cfg = _cfg()
    smap = oracle.rasterize(cfg, rects.astype(np.float64))
    sim = _sim(E, 4, cfg)
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    for _ in range(5):
        sim.explore_step()
        sim.step(policy_actions=sim.buf["policy_actions"])
    rng = np.random.default_rng(5)
    actions = rng.integers(0, 9, (E, 4, K, Hn)).astype(np.uint8)
    actions[:, :, 0, :] = 7
    actions[:, :, 1, 6:] = 255
    out = sim.explore_rollouts(actions)
    torch.cuda.synchronize()
    state = sim.state_numpy()
    belief = sim.belief.cpu().numpy().astype(np.float64)
    rew = out["rewards"].cpu().numpy()
    end = out["end_pose"].cpu().numpy()
    nfe = out["n_feasible"].cpu().numpy()
    obs = unpack_cell_mask(out["obsvd"].cpu().numpy(), cfg.ig_map_dim)
    ig = (types & 0xff) == L.POLICY_IG_GREEDY
    assert ig.sum() == 2 * E and not rew[~ig].any() and not obs[~ig].any()
    for e in range(E):
        for a in np.nonzero(ig[e])[0]:
            pose = state[e, a, [L.PX, L.PY, L.HEADING]]
            o = oracle.explore_rollouts(cfg, smap, pose, state[e, a, L.RADIUS], belief[e, a], actions[e, a])
            assert np.array_equal(obs[e, a], o["obsvd"].astype(bool)), (e, a)
            H.assert_close(rew[e, a], o["rewards"], "rewards %d %d" % (e, a), rtol=1e-5, atol=1e-7)
            H.assert_close(end[e, a], o["end_pose"], "end pose %d %d" % (e, a), rtol=1e-6, atol=1e-5)
            assert np.array_equal(nfe[e, a], o["n_feasible"]), (e, a)
    sim.close()


def test_rollouts_full_size_properties():
    """C4 size (1024 envs x 4 agents) x 10 rollouts x horizon 10: determinism, shard invariance, monotonicity
    (a sequence's reward never drops when it is extended), and a zero-speed sequence sees only what turning in
    place sees."""
    import torch
    from casim import scenarios as S
    E, K, Hn = 1024, 10, 10
    init, types, n, rects = S.exploration(E, seed=99)
    cfg = _cfg()
    sim = _sim(E, 4, cfg)
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    for _ in range(2):
        sim.explore_step()
        sim.step(policy_actions=sim.buf["policy_actions"])
    rng = np.random.default_rng(3)
    actions = rng.integers(0, 9, (E, 4, K, Hn)).astype(np.uint8)
    a = sim.explore_rollouts(actions)
    b = sim.explore_rollouts(actions)
    assert torch.equal(a["rewards"], b["rewards"]) and torch.equal(a["obsvd"], b["obsvd"])
    short = actions.copy()
    short[..., 5:] = 255
    s = sim.explore_rollouts(short)
    assert bool((s["rewards"] <= a["rewards"] + 1e-6).all())
    assert bool(((s["obsvd"] & ~a["obsvd"]) == 0).all())
    assert float(a["rewards"].max()) > 0.5
    lo, hi = 256, 384
    s2 = _sim(hi - lo, 4, cfg)
    s2.set_obstacles(rects)
    s2.bind(types=types[lo:hi], n_agents=n[lo:hi])
    s2.state.copy_(sim.state[:, lo:hi])
    s2.belief.copy_(sim.belief[lo:hi])
    c = s2.explore_rollouts(actions[lo:hi])
    assert torch.equal(c["rewards"], a["rewards"][lo:hi]) and torch.equal(c["end_pose"], a["end_pose"][lo:hi])
    s2.close()
    # tabulated mutual information (casim_explore_mi_table / casim_explore_set_mi_table): the same rewards, bit for bit,
    # with and without the other robots' claimed cells; and the probe marker gives n_feasible without a reward
    from casim import lib as _lib
    table = torch.zeros(tuple(sim.belief.shape), dtype=torch.int64, device=sim.belief.device)
    _lib.check(sim.lib.casim_explore_mi_table(sim.handle, _lib.ptr(sim.state), _lib.ptr(sim.belief), _lib.ptr(table), None),
               "casim_explore_mi_table")
    ex = rng.integers(-2**31, 2**31, (E, 4, 128)).astype(np.int32)
    plain_ex = sim.explore_rollouts(actions, exclude=ex)
    _lib.check(sim.lib.casim_explore_set_mi_table(sim.handle, _lib.ptr(table)), "casim_explore_set_mi_table")
    t0, t1 = sim.explore_rollouts(actions), sim.explore_rollouts(actions, exclude=ex)
    _lib.check(sim.lib.casim_explore_set_mi_table(sim.handle, None), "casim_explore_set_mi_table")
    assert torch.equal(t0["rewards"], a["rewards"]) and torch.equal(t0["obsvd"], a["obsvd"])
    assert torch.equal(t1["rewards"], plain_ex["rewards"]) and torch.equal(t1["n_feasible"], a["n_feasible"])
    assert float((plain_ex["rewards"] - a["rewards"]).abs().max()) > 0.1
    # slot list (casim_explore_set_rollout_slots): only the listed agents are evaluated, at their own [slot][k] positions
    ig_slots = np.nonzero(((types & 0xff) == L.POLICY_IG_GREEDY).ravel())[0]
    pick = torch.from_numpy(ig_slots[::3].astype(np.int32)).to(sim.belief.device)
    _lib.check(sim.lib.casim_explore_set_rollout_slots(sim.handle, _lib.ptr(pick), len(pick)), "casim_explore_set_rollout_slots")
    part = sim.explore_rollouts(actions)              # (fresh torch.empty outputs: the unlisted slots hold whatever was there)
    _lib.check(sim.lib.casim_explore_set_rollout_slots(sim.handle, None, 0), "casim_explore_set_rollout_slots")
    pl = pick.long()
    for k in ("rewards", "n_feasible", "obsvd", "end_pose"):
        assert torch.equal(part[k].flatten(0, 1)[pl], a[k].flatten(0, 1)[pl]), k
    assert sim.lib.casim_explore_set_rollout_slots(sim.handle, _lib.ptr(pick), E * 4 + 1) == 10001
    probe = actions.copy()
    probe[..., 4] = 254                                # CASIM_ROLLOUT_PROBE after four steps
    pr = sim.explore_rollouts(probe)
    four = actions.copy()
    four[..., 4:] = 255
    f4 = sim.explore_rollouts(four)
    assert torch.equal(pr["n_feasible"], f4["n_feasible"]) and float(pr["rewards"].abs().max()) == 0.0
    assert float(f4["rewards"].max()) > 0.1 and int(pr["obsvd"].abs().max()) == 0
    still = np.full((E, 4, 1, Hn), 1, np.uint8)      # v = 0, dphi = 0: the pose never changes
    st = sim.explore_rollouts(still)
    end = st["end_pose"][:, :, 0]
    ig = torch.from_numpy((types & 0xff) == L.POLICY_IG_GREEDY).to(end.device)
    pos = torch.stack([sim.state[0, :, :, 0], sim.state[0, :, :, 1], sim.state[0, :, :, 2]], -1)
    assert torch.equal(end[ig], pos[ig])
    sim.close()
