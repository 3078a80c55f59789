"""Shared helpers for the parity tests."""
import os

import numpy as np

from casim import layout as L

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# continuous tolerance stated by BASELINE.json north_star: 1e-5 relative (fp32 vs the reference's fp64)
RTOL = 1e-5
ATOL = 1e-5     # absolute floor for values near zero (angles, ego-frame projections)

CONT_WORDS = [k for k in range(L.STATE_WORDS) if k not in L.INT_WORDS]


def load(name):
    return np.load(os.path.join(GOLDEN, name))


def unpack_maps():
    z = load("maps.npz")
    maps = {0: None}
    rects = {0: np.zeros((0, 4))}
    for mid in (1, 2):
        maps[mid] = np.unpackbits(z["map_%d" % mid])[:300 * 300].reshape(300, 300).astype(np.uint8)
        rects[mid] = z["rects_%d" % mid]
    return maps, rects


def assert_close(a, b, what, rtol=RTOL, atol=ATOL):
    a = np.atleast_1d(np.asarray(a, dtype=np.float64))
    b = np.atleast_1d(np.asarray(b, dtype=np.float64))
    assert a.shape == b.shape, (what, a.shape, b.shape)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    err = np.abs(a - b)
    err[both_inf] = 0
    tol = atol + rtol * np.abs(b)
    tol[both_inf] = 1
    bad = ~(err <= tol)
    if bad.any():
        i = np.argwhere(bad)[0]
        raise AssertionError("%s: %d/%d mismatches, first at %s: %r vs %r" %
                             (what, bad.sum(), bad.size, tuple(i), a[tuple(i)], b[tuple(i)]))


def assert_state_close(s, ref, n_agents, what):
    """continuous words within tolerance, integer words (step_num, flags, type) exact"""
    for e in range(s.shape[0]):
        n = int(n_agents[e])
        for k in L.INT_WORDS:
            assert np.array_equal(s[e, :n, k], ref[e, :n, k]), (what, "word", k, "env", e, s[e, :n, k], ref[e, :n, k])
        a, b = s[e, :n][:, CONT_WORDS].copy(), ref[e, :n][:, CONT_WORDS]
        # heading_ego is reported in fp32: at the +-pi branch cut it may land on the other side (same angle)
        k = CONT_WORDS.index(L.HEADING_EGO)
        d = a[:, k] - b[:, k]
        a[:, k] = b[:, k] + (d + np.pi) % (2 * np.pi) - np.pi
        assert_close(a, b, "%s env %d" % (what, e))


# ---------------------------------------------------------------------------------------------
# random mid-episode batches (vectorised variant of oracle/gen_golden.py:random_case) for GPU-vs-oracle tests
def f32(x):
    return np.asarray(x, dtype=np.float32).astype(np.float64)


def random_batch(rng, E, A, sensor_choices=(0, 1), policies=None, dynamics=None, spread=None, n_min=2, T=6):
    """Returns dict(init, types, state, ext, traj, n_agents): fp32-representable doubles."""
    policies = policies or [L.POLICY_STATIC, L.POLICY_NONCOOPERATIVE, L.POLICY_NONCOOPERATIVE, L.POLICY_NONCOOPERATIVE,
                            L.POLICY_EXTERNAL, L.POLICY_CARRL, L.POLICY_PEDESTRIAN, L.POLICY_LEARNING]
    dynamics = dynamics or [L.DYN_UNICYCLE, L.DYN_UNICYCLE, L.DYN_FIRST_ORDER, L.DYN_UNICYCLE_MAX_TURN_RATE]
    n_agents = rng.integers(min(n_min, A), A + 1, size=E).astype(np.int32)
    pol = rng.choice(policies, size=(E, A))
    dyn = rng.choice(dynamics, size=(E, A))
    dyn = np.where(pol == L.POLICY_PEDESTRIAN, L.DYN_EXTERNAL, dyn)
    mask = rng.choice(list(sensor_choices), size=(E, A))
    types = (pol | (dyn << 8) | (mask << 16)).astype(np.int32)
    sp = rng.choice([2.0, 4.0, 8.0, 14.0], size=(E, 1, 1)) if spread is None else np.full((E, 1, 1), spread)
    start = f32(rng.uniform(-1, 1, (E, A, 2)) * sp)
    goal = f32(rng.uniform(-1, 1, (E, A, 2)) * sp)
    radius = f32(rng.uniform(0.2, 0.8, (E, A)))
    pref = f32(rng.uniform(0.5, 1.5, (E, A)))
    head0 = f32(rng.uniform(-np.pi, np.pi, (E, A)))
    init = np.concatenate([start, goal, radius[..., None], pref[..., None], head0[..., None]], -1)
    s = np.zeros((E, A, L.STATE_WORDS))
    along = start + (goal - start) * rng.uniform(0, 1.05, (E, A, 1)) + rng.normal(0, 0.3, (E, A, 2))
    clustered = rng.normal(0, 1.2, (E, A, 2))
    pos = f32(np.where(rng.random((E, A, 1)) < 0.5, along, clustered))
    fl = np.zeros((E, A), np.int64)
    u = rng.random((E, A))
    at_goal = u < 0.12
    pos = np.where(at_goal[..., None], f32(goal + rng.uniform(-0.3, 0.3, (E, A, 2))), pos)
    fl = np.where(at_goal, L.F_IS_AT_GOAL | L.F_IS_DONE | np.where(rng.random((E, A)) < 0.5, L.F_WAS_AT_GOAL_ALREADY, 0), fl)
    coll = (u >= 0.12) & (u < 0.24)
    fl = np.where(coll, L.F_IN_COLLISION | L.F_IS_DONE | np.where(rng.random((E, A)) < 0.6, L.F_WAS_IN_COLLISION_ALREADY, 0)
                  | np.where(rng.random((E, A)) < 0.3, L.F_RAN_OUT_OF_TIME, 0), fl)
    tout = (u >= 0.24) & (u < 0.32)
    fl = np.where(tout, L.F_RAN_OUT_OF_TIME | L.F_IS_DONE, fl)
    s[..., L.PX], s[..., L.PY] = pos[..., 0], pos[..., 1]
    s[..., L.GX], s[..., L.GY] = goal[..., 0], goal[..., 1]
    s[..., L.RADIUS], s[..., L.PREF_SPEED] = radius, pref
    s[..., L.HEADING] = f32(rng.uniform(-np.pi, np.pi, (E, A)))
    spd = f32(rng.uniform(0, 1, (E, A)) * pref)
    s[..., L.SPEED] = spd
    s[..., L.VX] = f32(spd * np.cos(s[..., L.HEADING]))
    s[..., L.VY] = f32(spd * np.sin(s[..., L.HEADING]))
    s[..., L.DELTA_HEADING] = f32(rng.uniform(-0.5, 0.5, (E, A)))
    s[..., L.TURNING_DIR] = f32(rng.choice([0.0, 0.11, -0.11, 0.05, -0.3, 1.0], size=(E, A)) * rng.uniform(0.5, 1, (E, A)))
    step_num = rng.integers(0, 60, (E, A))
    step_num = np.where(rng.random((E, A)) < 0.15, 0, step_num)
    s[..., L.STEP_NUM] = step_num
    s[..., L.T] = f32(step_num * 0.1)
    s[..., L.TIME_REMAINING] = f32(np.where(rng.random((E, A)) < 0.5, rng.uniform(0.5, 30, (E, A)), rng.uniform(-0.05, 0.25, (E, A))))
    gd = goal - pos
    s[..., L.DIST_TO_GOAL] = f32(np.hypot(gd[..., 0], gd[..., 1]))
    s[..., L.PAST_DIST_TO_GOAL] = f32(s[..., L.DIST_TO_GOAL] + rng.uniform(0, 0.15, (E, A)))
    for k, lo, hi in ((L.PAST_ACT0_V, 0, 1.5), (L.PAST_ACT0_W, -1, 1), (L.PAST_ACT1_V, 0, 1.5), (L.PAST_ACT1_W, -1, 1)):
        s[..., k] = f32(rng.uniform(lo, hi, (E, A)))
    he = s[..., L.HEADING] - np.arctan2(gd[..., 1], gd[..., 0])
    s[..., L.HEADING_EGO] = f32((he + np.pi) % (2 * np.pi) - np.pi)
    s[..., L.FLAGS] = fl
    s[..., L.TYPE] = types
    ext = f32(np.stack([rng.uniform(0, 1.5, (E, A)), rng.uniform(-1.2, 1.2, (E, A))], -1))
    ext[..., 0] = np.where(pol == L.POLICY_CARRL, rng.integers(0, 11, (E, A)), ext[..., 0])
    ext[..., 1] = np.where(pol == L.POLICY_CARRL, 0, ext[..., 1])
    ext[..., 1] = np.where(pol == L.POLICY_LEARNING, f32(rng.uniform(0, 1, (E, A))), ext[..., 1])
    p0 = pos + rng.normal(0, 0.2, (E, A, 2))
    v = rng.uniform(-1, 1, (E, A, T, 2))
    kk = np.arange(T)[None, None, :, None]
    traj = f32(np.concatenate([p0[:, :, None, :] + 0.1 * kk * v, v], -1))
    # zero the padded slots
    pad = np.arange(A)[None, :] >= n_agents[:, None]
    s[pad] = 0
    types = np.where(pad, 0, types).astype(np.int32)
    return dict(init=init, types=types, state=s, ext=ext, traj=traj, n_agents=n_agents)


def compare_step(what, n_agents, gstate, gbuf, ostate, oobs, check_sensors=True):
    """GPU (gstate float64 [E,A,24] decoded, gbuf dict of numpy arrays) vs oracle (ostate, ObsBuffers).
    Flags / done / ordering / laser hit index / grid cells exact; continuous values within RTOL/ATOL."""
    assert_state_close(gstate, ostate, n_agents, what)
    E, A = gstate.shape[:2]
    valid = np.arange(A)[None, :] < np.asarray(n_agents)[:, None]
    assert np.array_equal(gbuf["done"][valid], oobs.done[valid]), what + " done"
    assert np.array_equal(gbuf["game_over"], oobs.game_over), what + " game_over"
    if gbuf.get("collision_flags") is not None:
        assert np.array_equal(gbuf["collision_flags"][valid], oobs.collision_flags[valid]), what + " collision flags"
        assert_close(gbuf["nearest_dist"][valid], oobs.nearest_dist[valid], what + " nearest")
    assert_close(gbuf["rewards"][valid], oobs.rewards[valid], what + " rewards", atol=1e-7)
    compare_obs(what, n_agents, gbuf, oobs, check_sensors)


def compare_obs(what, n_agents, gbuf, oobs, check_sensors=True):
    E, A = gbuf["obs_misc"].shape[:2]
    valid = np.arange(A)[None, :] < np.asarray(n_agents)[:, None]
    assert_close(gbuf["obs_misc"][valid], oobs.obs_misc[valid], what + " obs_misc")
    if not check_sensors:
        return
    if gbuf.get("other_agents_states") is not None:
        g, o = gbuf["other_agents_states"][valid], oobs.other_agents_states[valid]
        # closest-N ordering: type and radius columns pin which agent sits in which row
        assert np.array_equal(g[..., 9], o[..., 9]), what + " other_agents ordering (type column)"
        assert np.array_equal(g[..., 6], o[..., 6].astype(np.float32)), what + " other_agents ordering (radius column)"
        assert_close(g, o, what + " other_agents_states")
    if gbuf.get("laserscan") is not None:
        assert_close(gbuf["laserscan"][valid], oobs.laserscan[valid], what + " laserscan", rtol=0, atol=1e-6)
    if gbuf.get("local_grid") is not None:
        assert np.array_equal(gbuf["local_grid"][valid], oobs.local_grid[valid]), what + " local_grid"


# ------------------------------------------------------------------------------------------------
# exploration fixtures (tests/golden/explore.npz, generated from the reference by oracle/gen_golden_explore.py)
def explore_maps(z):
    """-> list of [300,300] uint8 static maps, list of exact squared EDTs"""
    maps, d2 = [], []
    k = 0
    while "static_%d" % k in z.files:
        maps.append(np.unpackbits(z["static_%d" % k])[:300 * 300].reshape(300, 300).astype(np.uint8))
        d2.append(z["edf_d2_%d" % k])
        k += 1
    return maps, d2


def explore_case_batch(z, cases):
    """Stack golden explore cases (all sharing one map) into a batch of E = len(cases) envs x A = 1 agent.
    Returns dict(state [E,1,24] f64, oas [E,1,9,10], belief0 [E,1,60,60] f64, vis, occ [E,60,60] bool)."""
    E = len(cases)
    state = np.zeros((E, 1, L.STATE_WORDS))
    state[:, 0, L.PX] = z["pose"][cases, 0]
    state[:, 0, L.PY] = z["pose"][cases, 1]
    state[:, 0, L.HEADING] = z["pose"][cases, 2]
    state[:, 0, L.RADIUS] = z["radius"][cases]
    state[:, 0, L.PREF_SPEED] = 1.0
    state[:, 0, L.TYPE] = L.encode_type(L.POLICY_IG_GREEDY, L.DYN_FIRST_ORDER, L.SENSOR_OTHER_AGENTS)
    oas = z["oas"][cases][:, None]
    belief0 = z["belief0"][cases].astype(np.float64)[:, None]
    vis = np.unpackbits(z["vis"][cases], axis=-1)[:, :3600].reshape(E, 60, 60).astype(bool)
    occ = np.unpackbits(z["occ"][cases], axis=-1)[:, :3600].reshape(E, 60, 60).astype(bool)
    return dict(state=state, oas=np.ascontiguousarray(oas), belief0=np.ascontiguousarray(belief0), vis=vis, occ=occ)


def check_explore_vs_golden(what, z, cases, cfg, belief_after, out, rtol=RTOL):
    """belief_after [E,1,60,60], out = dict(policy_actions, best_mi, cand_mi, visible) vs the reference's results:
    visible-cell set, occupied/empty choice and the chosen action exact; MI scores within rtol."""
    b = explore_case_batch(z, cases)
    for n, k in enumerate(cases):
        vis = np.asarray(out["visible"][n, 0]).astype(bool)
        assert np.array_equal(vis, b["vis"][n]), (what, "visible set", k, int(vis.sum()), int(b["vis"][n].sum()))
        expect = b["belief0"][n, 0] * np.where(b["occ"][n], cfg.ig_rocc, np.where(b["vis"][n], cfg.ig_remp, 1.0))
        assert_close(belief_after[n, 0], expect, "%s belief case %d" % (what, k), rtol=rtol, atol=0)
        cand_ref = z["cand_mi"][k]
        cand = np.asarray(out["cand_mi"][n, 0], dtype=np.float64)
        assert np.array_equal(cand < 0, cand_ref < 0), (what, "feasibility", k, cand, cand_ref)
        assert_close(cand, cand_ref, "%s cand_mi case %d" % (what, k), rtol=rtol, atol=1e-9)
        assert_close(out["best_mi"][n, 0], z["best_mi"][k], "%s best_mi case %d" % (what, k), rtol=rtol, atol=1e-9)
        # the argmax is exact unless the reference's own top two candidates are closer than rounding
        srt = np.sort(cand_ref[cand_ref >= 0])
        decisive = len(srt) < 2 or (srt[-1] - srt[-2]) > 1e-9 * max(1.0, srt[-1]) or srt[-1] == srt[-2]
        if decisive:
            assert_close(out["policy_actions"][n, 0], z["action"][k], "%s action case %d" % (what, k), rtol=1e-6, atol=0)


def rollout_cases(z):
    """tests/golden/rollouts.npz -> list of dicts (one per case) with unpacked masks."""
    maps = []
    k = 0
    while "static_%d" % k in z.files:
        maps.append(np.unpackbits(z["static_%d" % k])[:300 * 300].reshape(300, 300).astype(np.uint8))
        k += 1
    cases = []
    for n in range(len(z["map_id"])):
        K = z["actions"][n].shape[0]
        cases.append(dict(
            index=n, map_id=int(z["map_id"][n]), pose=z["pose"][n], radius=float(z["radius"][n]), belief=z["belief"][n],
            actions=z["actions"][n], exclude=np.unpackbits(z["exclude"][n])[:3600].reshape(60, 60).astype(bool),
            rewards=z["rewards"][n], end_pose=z["end_pose"][n],
            obsvd=np.stack([np.unpackbits(z["obsvd"][n][k])[:3600].reshape(60, 60).astype(bool) for k in range(K)])))
    return maps, cases


def team_scene(z, sc):
    """tests/golden/team.npz scene -> per-step batch inputs (E = 1, A = 5: 3 explorers + 2 static targets)."""
    from casim import layout as L
    n_ig = int(z["n_ig"])
    steps = []
    for st in range(z["pose"].shape[1]):
        poses, done = z["pose"][sc, st], z["done"][sc, st]
        A = poses.shape[0]
        state = np.zeros((1, A, L.STATE_WORDS))
        state[0, :, L.PX], state[0, :, L.PY], state[0, :, L.HEADING] = poses[:, 0], poses[:, 1], poses[:, 2]
        state[0, :, L.RADIUS] = 0.5
        state[0, :, L.PREF_SPEED] = 1.0
        for i in range(A):
            pol = L.POLICY_IG_GREEDY if i < n_ig else L.POLICY_STATIC
            state[0, i, L.TYPE] = L.encode_type(pol, L.DYN_FIRST_ORDER, L.SENSOR_OTHER_AGENTS)
            state[0, i, L.FLAGS] = L.F_IS_DONE if done[i] else 0
        steps.append(dict(state=state, oas=z["oas"][sc, st][None], done=done,
                          belief_after=z["belief_after"][sc, st], team_reward=z["team_reward"][sc, st],
                          n_targets=z["n_targets"][sc, st],
                          team_mask=np.stack([np.unpackbits(m)[:3600].reshape(60, 60).astype(bool)
                                              for m in z["team_mask"][sc, st]])))
    return steps
