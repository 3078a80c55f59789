"""GPU parity of the RVO policy kernel (casim_rvo_step) against (a) actions recorded from the
UNMODIFIED reference RVOPolicy.find_next_action running over the reference's own C++ RVO2 library
(tests/golden/rvo.npz, oracle/gen_golden_rvo.py) and (b) the CPU oracle in lock-step on crowds at C2 shape.
The simulator arithmetic is binary32 in the reference's operation order: actions must agree to fp32 rounding of the
same fp64 value, and the pi/6 turn-limit decision (speed zeroed) must be identical."""
import numpy as np
import pytest

import helpers as H
from casim import layout as L

pytestmark = pytest.mark.gpu


def _cfg(**kw):
    from casim import lib
    c = lib.default_config()
    c.game_over_mode = 1
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def _check_actions(got, want, what):
    want32 = want.astype(np.float32)
    # the kernel rounds the same fp64 expressions to fp32; CUDA and glibc atan2/sqrt may differ in the last fp64 bit
    assert np.array_equal((got[..., 0] == 0), (want32[..., 0] == 0)), what + ": turn-limit decisions differ"
    np.testing.assert_allclose(got, want32, rtol=2e-7, atol=1e-7, err_msg=what)
    assert (got == want32).mean() > 0.95, what + ": more than 5% of the actions are not bit-identical"


def test_rvo_golden_actions_from_reference_policy(oracle_lib):
    import torch
    from casim.sim import BatchedSim
    z = H.load("rvo.npz")
    st = np.ascontiguousarray(z["state"])
    E, A = st.shape[:2]
    sim = BatchedSim(E, A, cfg=_cfg(), sensors=())
    sim.bind(types=st[:, :, L.TYPE].astype(np.int32), n_agents=z["n_agents"])
    sim.load_state_numpy(st)
    pol = sim.rvo_step(coop=z["coop"].astype(np.float32))
    torch.cuda.synchronize()
    got = pol.cpu().numpy()
    for e in range(E):
        n = int(z["n_agents"][e])
        _check_actions(got[e, :n], z["actions"][e, :n], "golden case %d" % e)
    sim.close()


def test_rvo_golden_actions_with_obstacles(oracle_lib):
    """The crossing-map obstacle set: actions recorded from the reference RVOPolicy called WITH obstacles."""
    import torch
    from casim.sim import BatchedSim
    z = H.load("rvo.npz")
    st = np.ascontiguousarray(z["obst_state"])
    E, A = st.shape[:2]
    sim = BatchedSim(E, A, cfg=_cfg(), sensors=())
    sim.bind(types=st[:, :, L.TYPE].astype(np.int32), n_agents=z["obst_n_agents"])
    sim.load_state_numpy(st)
    sim.rvo_set_obstacles(list(z["obst_polygons"]))
    pol = sim.rvo_step(coop=z["obst_coop"].astype(np.float32))
    torch.cuda.synchronize()
    got = pol.cpu().numpy()
    for e in range(E):
        n = int(z["obst_n_agents"][e])
        _check_actions(got[e, :n], z["obst_actions"][e, :n], "obstacle golden case %d" % e)
    sim.close()


def test_rvo_obstacles_lockstep_vs_oracle(oracle_lib):
    """Random polygon sets (non-convex stars that force edge splitting in the obstacle tree, 2-vertex walls, the
    crossing map): kernel == oracle on the actions over free-running steps, with the env's static map set too."""
    import torch
    from casim.sim import BatchedSim
    rng = np.random.default_rng(11)
    cross = [np.array(p, float) for p in z_cross()]
    for trial in range(6):
        if trial % 3 == 0:
            polys = cross
        elif trial % 3 == 1:
            polys = []
            for k in range(4):
                m = int(rng.integers(3, 7))
                ang, rad = np.sort(rng.uniform(0, 2 * np.pi, m)), rng.uniform(0.8, 2.5, m)
                polys.append(rng.uniform(-7, 7, 2) + np.stack([rad * np.cos(ang), rad * np.sin(ang)], 1))
        else:
            polys = [np.array([[x, y], [x + w, y + h]]) for x, y, w, h in rng.uniform(-6, 6, (4, 4))]
        E, A = 48, 6
        init = np.zeros((E, A, 7), np.float32)
        init[..., 0:2] = rng.uniform(-8, 8, (E, A, 2))
        init[..., 2:4] = rng.uniform(-8, 8, (E, A, 2))
        init[..., 4] = rng.uniform(0.2, 0.6, (E, A))
        init[..., 5] = rng.uniform(0.5, 1.5, (E, A))
        init[..., 6] = np.nan
        types = np.full((E, A), L.encode_type(L.POLICY_RVO, L.DYN_UNICYCLE, 0), np.int32)
        n = np.full((E,), A, np.int32)
        n[::5] = 3
        cfg = _cfg()
        sim = BatchedSim(E, A, cfg=cfg, sensors=())
        sim.reset(init, types, n)
        sim.rvo_set_obstacles(polys)
        for t in range(12):
            ostate = sim.state_numpy()
            want = oracle_lib.rvo_actions(cfg, ostate, None, n, polygons=polys)
            pol = sim.rvo_step()
            torch.cuda.synchronize()
            got = pol.cpu().numpy()
            valid = np.arange(A)[None, :] < n[:, None]
            _check_actions(got[valid], want[valid], "trial %d step %d" % (trial, t))
            oracle_lib.step(cfg, ostate, policy_actions=want, n_agents=n)
            sim.step(policy_actions=pol)
        sim.close()


def z_cross():
    return [[(10, 10), (2, 10), (2, 2), (10, 2)], [(-2, 10), (-10, 10), (-10, 2), (-2, 2)],
            [(10, -2), (2, -2), (2, -10), (10, -10)], [(-2, -2), (-10, -2), (-10, -10), (-2, -10)]]


@pytest.mark.parametrize("A,max_neighbors", [(10, 10), (20, 10), (20, 32), (6, 3)])
def test_rvo_lockstep_vs_oracle(oracle_lib, A, max_neighbors):
    """Free-running episodes of all-RVO crowds: every step the kernel's actions equal the oracle's, then both step."""
    import torch
    from casim import lib
    from casim import scenarios as S
    from casim.sim import BatchedSim
    E = 64
    init, types, n = S.open_space(E, A, seed=5, sensors=0) if A == 10 else S.formations(E, A, seed=5) if A == 20 else \
        S.open_space(E, 10, seed=6, sensors=0)
    init, types, n = init[:, :A].copy(), np.array(types)[:, :A].copy(), np.minimum(n, A).astype(np.int32)
    types[:] = L.encode_type(L.POLICY_RVO, L.DYN_UNICYCLE, 0)
    types[:, 0] = L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, 0)       # a non-RVO agent among them
    n = n.copy()
    n[::9] = max(2, A - 3)                                                          # ragged environments
    rng = np.random.default_rng(1)
    coop = rng.choice([1.0, 1.0, 0.5, 0.0, -0.5], (E, A)).astype(np.float32)
    cfg = _cfg()
    prm = lib.rvo_default_params()
    prm.max_neighbors = max_neighbors
    oprm = oracle_lib.rvo_default_params()
    oprm.max_neighbors = max_neighbors
    sim = BatchedSim(E, A, cfg=cfg, sensors=())
    sim.reset(init, types, n)
    n_moving = 0
    for t in range(40):
        ostate = sim.state_numpy()
        want = oracle_lib.rvo_actions(cfg, ostate, coop.astype(np.float64), n, params=oprm)
        pol = sim.rvo_step(coop=coop, params=prm)
        torch.cuda.synchronize()
        got = pol.cpu().numpy()
        rvo = (ostate[:, :, L.TYPE].astype(np.int64) & 0xff) == L.POLICY_RVO
        valid = rvo & (np.arange(A)[None, :] < n[:, None])
        _check_actions(got[valid], want[valid], "A=%d step %d" % (A, t))
        n_moving += int((got[valid][:, 0] > 0).sum())
        oobs = oracle_lib.step(cfg, ostate, policy_actions=want, n_agents=n)
        sim.step(policy_actions=pol)
        torch.cuda.synchronize()
        H.assert_state_close(sim.state_numpy(), ostate, n, "state A=%d step %d" % (A, t))
        assert np.array_equal(sim.buf["done"].cpu().numpy().astype(bool), oobs.done.astype(bool))
    assert n_moving > 100
    sim.close()


def test_rvo_through_env_api_and_obstacles():
    import torch
    from casim import plugins as P
    from casim.env import BatchedCollisionAvoidanceEnv
    from casim.plugins import Agent
    agents = [Agent(-3, 0, 3, 0, 0.5, 1.0, 0.0, P.RVOPolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 0),
              Agent(3, 0.1, -3, 0, 0.5, 1.0, np.pi, P.RVOPolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 1,
                    cooperation_coef=0.5)]
    env = BatchedCollisionAvoidanceEnv(num_envs=1, max_agents=2)
    env.set_agents(agents)
    env.reset()
    for t in range(80):
        obs, rewards, game_over, info = env.step({})
    torch.cuda.synchronize()
    # two reciprocal agents swap sides without colliding (the reference example's RVO pair does the same)
    assert not agents[0].in_collision and not agents[1].in_collision
    assert agents[0].pos_global_frame[0] > 2.0 and agents[1].pos_global_frame[0] < -2.0
    env.close()
    # with the crossing map: an RVO agent sent straight at a block is held outside it by the obstacle ORCA lines
    through = [Agent(0.0, -9.0, 6.0, 9.0, 0.4, 1.0, None, P.RVOPolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 0)]
    env = BatchedCollisionAvoidanceEnv(num_envs=1, max_agents=2)
    env.set_agents(through)
    env.set_static_map(z_cross())
    env.reset()
    for t in range(300):
        env.step({})
    torch.cuda.synchronize()
    x, y = through[0].pos_global_frame
    assert np.isfinite([x, y]).all() and y > -9.0 and not (x > 2.0 and y > 2.0)
    env.close()
    env = BatchedCollisionAvoidanceEnv(num_envs=1, max_agents=2)
    env.set_agents(through)
    env.set_static_map(np.array([[2, 10, 10, 2]], np.float32))     # rectangles only: no polygons for the simulator
    with pytest.raises(RuntimeError):
        env.reset()
    env.close()


def test_rvo_c2_full_size_properties(oracle_lib):
    """4096 envs x 10 all-RVO agents (the rvo bench workload), 60 free-running steps: actions stay inside the
    policy's envelope (0 <= speed <= pref_speed up to the binary32 slack of the reference's linear programs, ~0.3 %, |delta heading| <= pi/6), no NaN, the result does not
    depend on how the batch is cut (first 64 envs alone give bit-identical states), and reciprocal avoidance keeps
    the collision rate far below the non-cooperative crowd's on the same scenarios."""
    import torch
    from casim import scenarios as S
    from casim.sim import BatchedSim
    E, A = 4096, 10
    cfg = _cfg()
    init, types, n = S.open_space(E, A, seed=1234, sensors=0, static_every=0, policy=L.POLICY_RVO)
    sim = BatchedSim(E, A, cfg=cfg, sensors=())
    sim.reset(init, types, n)
    small = BatchedSim(64, A, cfg=cfg, sensors=())
    small.reset(init[:64], types[:64], n[:64])
    pref = torch.as_tensor(init[:, :, 5], device=sim.device)
    for t in range(60):
        pol = sim.rvo_step()
        alive = (sim.state[5, :, :, 2].view(torch.int32) & L.F_IS_DONE) == 0
        assert torch.isfinite(pol).all()
        assert (pol[..., 0] >= 0).all() and (pol[..., 0][alive] <= pref[alive] * 1.01 + 1e-5).all()
        assert (pol[..., 1].abs() <= np.float32(np.pi / 6) * (1 + 1e-6)).all()
        sim.step(policy_actions=pol)
        small.step(policy_actions=small.rvo_step())
    torch.cuda.synchronize()
    assert torch.equal(sim.state[:, :64], small.state)
    rvo_coll = float(((sim.state[5, :, :, 2].view(torch.int32) & L.F_IN_COLLISION) != 0).float().mean())
    init2, types2, n2 = S.open_space(E, A, seed=1234, sensors=0, static_every=0, policy=L.POLICY_NONCOOPERATIVE)
    nc = BatchedSim(E, A, cfg=cfg, sensors=())
    nc.reset(init2, types2, n2)
    for t in range(60):
        nc.step()
    torch.cuda.synchronize()
    nc_coll = float(((nc.state[5, :, :, 2].view(torch.int32) & L.F_IN_COLLISION) != 0).float().mean())
    assert rvo_coll < 0.5 * nc_coll, (rvo_coll, nc_coll)
    for s in (sim, small, nc):
        s.close()
