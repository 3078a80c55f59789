"""GPU parity of casim_explore_team_update (ig_mcts.update_belief) against the golden vectors recorded from the
reference's ig_mcts / targetMap classes and against the CPU oracle on a C4-shaped batch over several steps."""
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


def test_team_update_golden_vs_reference():
    import torch
    from casim.sim import unpack_cell_mask
    z = H.load("team.npz")
    cfg = _cfg()
    smap = np.unpackbits(z["static"])[:300 * 300].reshape(300, 300).astype(np.uint8)
    n_ig = int(z["n_ig"])
    for sc in range(z["pose"].shape[0]):
        steps = H.team_scene(z, sc)
        A = steps[0]["state"].shape[1]
        sim = _sim(1, A, cfg)
        sim.set_map(smap)
        sim.bind(types=steps[0]["state"][:, :, L.TYPE].astype(np.int32), n_agents=np.full((1,), A, np.int32))
        sim.belief.copy_(torch.from_numpy(np.repeat(z["belief0"][sc][None, None], A, axis=1)))
        for st, s in enumerate(steps):
            sim.load_state_numpy(s["state"])
            sim.buf["other_agents_states"].copy_(torch.from_numpy(s["oas"].astype(np.float32)))
            mask, reward = sim.explore_team_update()
            torch.cuda.synchronize()
            b = sim.belief.cpu().numpy().astype(np.float64)
            m = unpack_cell_mask(mask.cpu().numpy(), 60)
            for i in range(n_ig):
                H.assert_close(b[0, i], s["belief_after"][i], "scene %d step %d agent %d belief" % (sc, st, i),
                               rtol=1e-6, atol=0)
                if s["done"][i]:
                    continue
                assert np.array_equal(m[0, i], s["team_mask"][i]), (sc, st, i)
                H.assert_close(float(reward[0, i]), s["team_reward"][i], "team reward", rtol=1e-5, atol=1e-7)
                assert int(sim.buf["team_n"][0][0, i]) == int(s["n_targets"][i])
        sim.close()


def test_team_update_lockstep_vs_oracle(oracle_lib):
    """C4 scene (2 explorers + 2 targets per env, crossing map), 12 steps: team update + greedy action each step
    (explorers move), GPU against the oracle run on the GPU's own state."""
    import torch
    from casim import scenarios as S
    from casim.sim import unpack_cell_mask
    oracle = oracle_lib
    E = 32
    init, types, n, rects = S.exploration(E, seed=21)
    cfg = _cfg()
    smap = oracle.rasterize(cfg, rects.astype(np.float64))
    sim = _sim(E, 4, cfg)
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    tp, npv = np.zeros((E, 4, 32, 2)), np.zeros((E, 4), np.int32)
    ig = (types & 0xff) == L.POLICY_IG_GREEDY
    seen_two = 0
    for t in range(12):
        state = sim.state_numpy()
        oas = sim.buf["other_agents_states"].cpu().numpy().astype(np.float64)
        belief = sim.belief.cpu().numpy().astype(np.float64)
        o = oracle.team_update(cfg, state, oas, belief, smap, tp, npv, n_agents=n)
        tp, npv = o["targets"], o["n_targets"]
        mask, reward = sim.explore_team_update()
        torch.cuda.synchronize()
        gb = sim.belief.cpu().numpy().astype(np.float64)
        H.assert_close(gb[ig], belief[ig], "belief t=%d" % t, rtol=1e-6, atol=0)
        assert np.array_equal(unpack_cell_mask(mask.cpu().numpy(), cfg.ig_map_dim)[ig], o["team_mask"][ig].astype(bool)), t
        H.assert_close(reward.cpu().numpy()[ig], o["team_reward"][ig], "team reward t=%d" % t, rtol=1e-5, atol=1e-7)
        assert np.array_equal(sim.buf["team_n"][0].cpu().numpy()[ig], npv[ig])
        live = (np.arange(32)[None, None, :] < npv[:, :, None]) & ig[:, :, None]      # entries past the count are unspecified
        H.assert_close(sim.buf["team_targets"][0].cpu().numpy()[live], tp[live], "targets t=%d" % t, rtol=1e-12, atol=1e-12)
        seen_two += int((o["team_mask"][ig].sum((-1, -2)) > 60).sum())
        sim.explore_step()          # greedy action on the team-updated belief (moves the explorers)
        sim.step(policy_actions=sim.buf["policy_actions"])
    assert seen_two > 0
    sim.close()
