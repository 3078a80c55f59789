"""GPU parity of the exploration path (casim_explore_step) against (a) the golden vectors recorded from the
reference's edfMap / targetMap / ig_greedy classes and (b) the CPU oracle in lock-step over an episode.
Visible-cell sets, occupied/empty choices, feasibility and the greedy action bit-exact; odds and MI scores
within 1e-5 relative (north_star)."""
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


def test_explore_golden_vs_reference(oracle_lib):
    import torch
    z = H.load("explore.npz")
    maps, _ = H.explore_maps(z)
    cfg = _cfg()
    ncases = len(z["map_id"])
    for mid in range(len(maps)):
        cases = [k for k in range(ncases) if int(z["map_id"][k]) == mid]
        b = H.explore_case_batch(z, cases)
        E = len(cases)
        sim = _sim(E, 1, cfg)
        if mid == 2:
            sim.set_obstacles(z["rects_2"].astype(np.float32))       # the rasteriser + EDT kernels
        else:
            sim.set_map(maps[mid])                                    # ready-made raster + EDT kernel
        types = b["state"][:, :, L.TYPE].astype(np.int32)
        sim.bind(types=types, n_agents=np.ones((E,), np.int32))
        sim.load_state_numpy(b["state"])
        sim.buf["other_agents_states"].copy_(torch.from_numpy(b["oas"].astype(np.float32)))
        sim.belief.copy_(torch.from_numpy(b["belief0"].astype(np.float32)))
        sim.explore_step()
        torch.cuda.synchronize()
        after = sim.belief.cpu().numpy().astype(np.float64)
        out = {k: sim.buf[k].cpu().numpy() for k in ("policy_actions", "best_mi", "cand_mi")}
        out["visible"] = after != b["belief0"]
        H.check_explore_vs_golden("gpu map %d" % mid, z, cases, cfg, after, out, rtol=1e-5)
        cnt = sim.buf["explore_counters"].cpu().numpy()
        assert cnt[3] == 0 and cnt[0] > cnt[2] > 0 and cnt[1] > cnt[2]
        sim.close()


def test_explore_lockstep_episode_vs_oracle(oracle_lib):
    """Config C4 shape (2 ig_greedy + 2 Static targets per env, crossing map), 40 steps with auto-reset:
    every step the GPU explore step is checked against the oracle run on the GPU's own state / belief."""
    import torch
    from casim import scenarios as S
    oracle = oracle_lib
    E = 48
    init, types, n, rects = S.exploration(E, seed=7)
    cfg = _cfg(auto_reset=1)
    smap = oracle.rasterize(cfg, rects.astype(np.float64))
    sim = _sim(E, 4, cfg)
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    torch.cuda.synchronize()
    prev_go = np.zeros((E,), np.uint8)
    n_restarts = n_updates_occ = n_indecisive = 0
    for t in range(40):
        state = sim.state_numpy()
        oas = sim.buf["other_agents_states"].cpu().numpy().astype(np.float64)
        belief = sim.belief.cpu().numpy().astype(np.float64)
        before = belief.copy()
        o = oracle.explore_step(cfg, state, oas, belief, smap, n_agents=n, init=init.astype(np.float64),
                                prev_game_over=prev_go)
        sim.explore_step()
        torch.cuda.synchronize()
        gb = sim.belief.cpu().numpy().astype(np.float64)
        ig = (types & 0xff) == L.POLICY_IG_GREEDY
        H.assert_close(gb[ig], belief[ig], "belief t=%d" % t, rtol=1e-6, atol=0)
        n_updates_occ += int((belief[ig] > before[ig]).sum())
        gc, oc = sim.buf["cand_mi"].cpu().numpy()[ig], o["cand_mi"][ig]
        assert np.array_equal(gc < 0, oc < 0), "feasibility t=%d" % t
        H.assert_close(gc, oc, "cand_mi t=%d" % t, rtol=1e-5, atol=1e-9)
        ga, oa = sim.buf["policy_actions"].cpu().numpy()[ig], o["policy_actions"][ig]
        srt = np.sort(oc, axis=1)
        decisive = (srt[:, -1] - srt[:, -2] > 1e-9 * np.maximum(1.0, srt[:, -1])) | (srt[:, -1] == srt[:, -2])
        n_indecisive += int((~decisive).sum())
        H.assert_close(ga[decisive], oa[decisive], "action t=%d" % t, rtol=1e-6, atol=0)
        # instrumented counts (cells tested, EDF samples, visible cells) are part of the measurement contract
        assert np.array_equal(sim.buf["explore_counters"].cpu().numpy().astype(np.uint64)[:3], o["counters"][:3]), t
        n_restarts += int(prev_go.sum())
        sim.step(policy_actions=sim.buf["policy_actions"])
        torch.cuda.synchronize()
        prev_go = sim.buf["game_over"].cpu().numpy().copy()
    assert n_restarts > 0 and n_updates_occ > 0
    assert n_indecisive < 10
    sim.close()


def test_explore_c4_full_size_properties(oracle_lib):
    """Config C4 size (1024 envs x 4 agents): determinism, shard invariance, odds only ever multiplied by
    rOcc / rEmp, and agreement with the oracle on a sample of environments."""
    import torch
    from casim import scenarios as S
    oracle = oracle_lib
    E = 1024
    init, types, n, rects = S.exploration(E, seed=1236)
    cfg = _cfg()
    sim = _sim(E, 4, cfg)
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    for _ in range(3):
        sim.explore_step()
        sim.step(policy_actions=sim.buf["policy_actions"])
    torch.cuda.synchronize()
    ref_belief, ref_state = sim.belief.clone(), sim.state.clone()
    lo, hi = 512, 768
    s2 = _sim(hi - lo, 4, cfg)
    s2.set_obstacles(rects)
    s2.reset(init[lo:hi], types[lo:hi], n[lo:hi])
    for _ in range(3):
        s2.explore_step()
        s2.step(policy_actions=s2.buf["policy_actions"])
    torch.cuda.synchronize()
    assert torch.equal(s2.belief, ref_belief[lo:hi]) and torch.equal(s2.state, ref_state[:, lo:hi])
    s2.close()
    # every cell is prior * rOcc^a * rEmp^b with a + b <= 3
    b = ref_belief.cpu().numpy().astype(np.float64)
    allowed = np.array(sorted({1.5 ** a * 0.66 ** c for a in range(4) for c in range(4) if a + c <= 3}))
    d = np.abs(b[..., None] - allowed).min(-1) / b
    assert d.max() < 1e-6
    # oracle on a sample of environments for one more step
    sel = np.sort(np.random.default_rng(3).choice(E, 64, replace=False))
    state = sim.state_numpy()[sel]
    oas = sim.buf["other_agents_states"].cpu().numpy().astype(np.float64)[sel]
    belief = np.ascontiguousarray(b[sel])
    smap = oracle.rasterize(cfg, rects.astype(np.float64))
    o = oracle.explore_step(cfg, np.ascontiguousarray(state), oas, belief, smap, n_agents=n[sel])
    sim.explore_step()
    torch.cuda.synchronize()
    ig = (types[sel] & 0xff) == L.POLICY_IG_GREEDY
    H.assert_close(sim.belief.cpu().numpy()[sel][ig], belief[ig], "C4 belief", rtol=1e-6, atol=0)
    H.assert_close(sim.buf["cand_mi"].cpu().numpy()[sel][ig], o["cand_mi"][ig], "C4 cand_mi", rtol=1e-5, atol=1e-9)
    sim.close()


def test_explore_and_rollouts_per_env_maps():
    """n_maps == E: every environment samples ITS OWN distance field (the `n_maps == 1 ? 0 : e` indexing of k_explore,
    k_rollouts and k_team_update).  Environments alternate between two different maps; each must behave exactly like the
    same environment in a batch that shares its map."""
    import torch
    from casim import scenarios as S
    E = 16
    init, types, n, rects = S.exploration(E, seed=77)
    z = H.load("explore.npz")
    maps, _ = H.explore_maps(z)
    cfg = _cfg()
    which = np.arange(E) % 2                                       # map 0 / map 1 of the exploration fixtures
    rng = np.random.default_rng(2)
    actions = rng.integers(0, 9, (E, 4, 6, 8)).astype(np.uint8)

    def run(map_arg, lo=0, hi=E, idx=None):
        idx = np.arange(lo, hi) if idx is None else idx
        sim = _sim(len(idx), 4, cfg)
        sim.set_map(map_arg)
        sim.reset(init[idx], types[idx], n[idx])
        outs = []
        for _ in range(3):
            sim.explore_step()
            outs.append(sim.buf["policy_actions"].clone())
            sim.step(policy_actions=sim.buf["policy_actions"])
        r = sim.explore_rollouts(actions[idx])
        torch.cuda.synchronize()
        res = (sim.belief.clone(), sim.state.clone(), torch.stack(outs), r["rewards"].clone(), r["n_feasible"].clone(), r["obsvd"].clone())
        sim.close()
        return res

    per_env = run(np.stack([maps[int(w)] for w in which]))
    differ = False
    for m in (0, 1):
        idx = np.nonzero(which == m)[0]
        shared = run(maps[m], idx=idx)
        ti = torch.from_numpy(idx).cuda()
        assert torch.equal(per_env[0][ti], shared[0]) and torch.equal(per_env[1][:, ti], shared[1])
        assert torch.equal(per_env[2][:, ti], shared[2])
        for k in (3, 4, 5):
            assert torch.equal(per_env[k][ti], shared[k]), k
        other = run(maps[1 - m], idx=idx)                          # the wrong map does change the outcome
        differ |= not torch.equal(other[0], shared[0]) or not torch.equal(other[3], shared[3])
    assert differ
