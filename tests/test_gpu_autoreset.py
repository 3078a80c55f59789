"""auto_reset with policy kernels: an environment whose episode ended at step k restarts inside step k+1
from its initial condition.  The action applied in step k+1 must be the policy's answer on the FRESH state (what
env.reset() followed by env.step() computes in the reference, collision_avoidance_env.py:234-266 then :287-340), not on
the finished episode's terminal state.  Oracle = reset state substituted for the restarted envs -> policy -> step."""
import numpy as np
import pytest

import helpers as H
from casim import layout as L

pytestmark = pytest.mark.gpu


def _cfg(**kw):
    from casim import lib
    c = lib.default_config()
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def _np(sim):
    return {k: (None if v is None or isinstance(v, list) else v.detach().cpu().numpy()) for k, v in sim.buf.items()}


def _crowd(rng, E, A, policy, spread):
    """Dense crowds heading through each other: plenty of collisions / goals -> many restarts within a few steps."""
    ang = rng.uniform(-np.pi, np.pi, (E, A))
    rad = rng.uniform(0.8, spread, (E, A))
    start = np.stack([rad * np.cos(ang), rad * np.sin(ang)], -1)
    goal = -start * rng.uniform(0.3, 1.0, (E, A, 1))
    init = np.concatenate([start, goal, rng.uniform(0.2, 0.5, (E, A, 1)), rng.uniform(0.6, 1.4, (E, A, 1)),
                           np.full((E, A, 1), np.nan)], -1)
    init = H.f32(init)
    types = np.full((E, A), L.encode_type(policy, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS), np.int32)
    n = rng.integers(2, A + 1, E).astype(np.int32)
    return init, types, n


@pytest.mark.parametrize("policy", ["ga3c", "rvo"])
def test_policy_on_restart_step_vs_oracle(oracle_lib, policy):
    import torch
    from casim.sim import BatchedSim
    oracle = oracle_lib
    rng = np.random.default_rng(17)
    E, A = 48, 8
    pid = L.POLICY_GA3C_CADRL if policy == "ga3c" else L.POLICY_RVO
    init, types, n = _crowd(rng, E, A, pid, 3.0)
    cfg = _cfg(game_over_mode=0, auto_reset=1)
    w = oracle.ga3c_random_weights(2)
    sim = BatchedSim(E, A, cfg=cfg, sensors=("other_agents_states",))
    sim.reset(init, types, n)
    if policy == "ga3c":
        sim.ga3c_set_weights(w)
    torch.cuda.synchronize()
    state0, _ = oracle.reset(cfg, init, types, n_agents=n)
    prev_go = np.zeros((E,), np.uint8)
    n_restart_agents = 0
    for t in range(40):
        ostate = sim.state_numpy()
        eff = np.where(prev_go[:, None, None].astype(bool), state0, ostate)       # what the policy must look at
        if policy == "ga3c":
            o = oracle.ga3c_forward(cfg, np.ascontiguousarray(eff), w, n_agents=n)
            opol = o["policy_actions"]
            pol = sim.ga3c_forward(use_tensor_cores=False)
        else:
            opol = oracle.rvo_actions(cfg, np.ascontiguousarray(eff), n_agents=n)
            pol = sim.rvo_step()
        torch.cuda.synchronize()
        gpol = pol.cpu().numpy()
        valid = np.arange(A)[None, :] < n[:, None]
        live = valid & ((eff[..., L.FLAGS].astype(np.int64) & L.F_IS_DONE) == 0)
        if policy == "ga3c":
            same = o["action_idx"] == sim.buf["action_idx"].cpu().numpy()
            srt = np.sort(o["probs"], axis=-1)
            decisive = (srt[..., -1] - srt[..., -2]) > 1e-4
            assert same[live & decisive].all(), t
            cmp_ = live & same
        else:
            cmp_ = live
        np.testing.assert_allclose(gpol[cmp_], opol[cmp_].astype(np.float32), rtol=3e-7, atol=1e-7, err_msg="t=%d" % t)
        fresh = prev_go.astype(bool)[:, None] & valid
        n_restart_agents += int(fresh.sum())
        if fresh.any():          # the restarted agents did get an action (not the finished-agent zeros)
            assert (np.abs(gpol[fresh]).sum(-1) > 0).mean() > 0.5
        # the step itself, fed with the ORACLE's policy output where the GPU agreed (identical inputs on both sides)
        gp = torch.from_numpy(np.where(cmp_[..., None], opol, gpol).astype(np.float32)).cuda()
        oobs = oracle.step(cfg, ostate, policy_actions=np.where(cmp_[..., None], opol, gpol).astype(np.float32).astype(np.float64),
                           n_agents=n, init=init, types0=types, prev_game_over=prev_go)
        sim.step(policy_actions=gp)
        torch.cuda.synchronize()
        H.compare_step("%s auto-reset t=%d" % (policy, t), n, sim.state_numpy(), _np(sim), ostate, oobs)
        prev_go = oobs.game_over.copy()
    assert n_restart_agents > 20
    sim.close()
