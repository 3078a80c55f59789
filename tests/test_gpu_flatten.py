"""GPU: the flattened observation wire format (casim_flatten_obs, casim.wrappers) against the array the reference's
MultiagentFlattenDictWrapper.observation produced for the same agent states (tests/golden/flatten.npz, recorded by
oracle/gen_golden.py:gen_flatten from the unmodified reference env + wrapper)."""
import numpy as np
import pytest

import helpers as H
from casim import layout as L

pytestmark = pytest.mark.gpu


def test_flatten_matches_reference_wrapper():
    import torch
    from casim import lib
    from casim.sim import BatchedSim
    z = H.load("flatten.npz")
    keys = [str(k) for k in z["keys"]]
    maps, rects = H.unpack_maps()
    T, n = z["states"].shape[:2]
    A = 10                                           # Config.MAX_NUM_AGENTS_IN_ENVIRONMENT agent slots
    cfg = lib.default_config()
    sim = BatchedSim(T, A, cfg=cfg, sensors=("other_agents_states", "laserscan", "local_grid"))
    sim.set_obstacles(rects[int(z["map_id"])].astype(np.float32))
    st = np.zeros((T, A, L.STATE_WORDS))
    st[:, :n] = z["states"]                          # one environment per recorded step
    types = np.zeros((T, A), np.int32)
    types[:, :n] = z["types"][None, :]
    sim.bind(types=types, n_agents=np.full((T,), n, np.int32))
    sim.load_state_numpy(st)
    sim.observe()
    flat = sim.flatten_obs(keys)
    torch.cuda.synchronize()
    got = flat.cpu().numpy()
    want = z["flat"]
    assert got.shape == want.shape == (T, A * sim.flat_row_size(keys))
    S = sim.flat_row_size(keys)
    off = 0
    for k in keys:
        w = sim.flat_row_size([k])
        g = got.reshape(T, A, S)[:, :, off:off + w]
        r = want.reshape(T, A, S)[:, :, off:off + w]
        if k in ("local_grid", "num_other_agents", "use_ppo"):
            assert np.array_equal(g, r), k                                   # flags / cells / counts: exact
        else:
            H.assert_close(g, r, "flatten key " + k)
        off += w
    assert (got.reshape(T, A, S)[:, n:] == 0).all()                          # absent agent slots stay zero
    sim.close()


def test_wrapper_and_vecenv_api():
    import torch
    from casim import plugins as P
    from casim.env import BatchedCollisionAvoidanceEnv
    from casim.plugins import Agent
    from casim.wrappers import BatchedVecEnv, FlattenDictWrapper, MultiagentFlattenDictWrapper
    E, A = 6, 3
    keys = ['dist_to_goal', 'rel_goal', 'radius', 'heading_ego_frame', 'pref_speed', 'other_agents_states', 'use_ppo']
    envs_agents = [[Agent(-3 - e, -3, 3, 3, 0.5, 1.0, None, P.LearningPolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 0),
                    Agent(3, 3 + e, -3, -3, 0.5, 1.0, None, P.NonCooperativePolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 1)]
                   for e in range(E)]
    env = BatchedCollisionAvoidanceEnv(num_envs=E, max_agents=A)
    env.set_agents(envs_agents)
    w = MultiagentFlattenDictWrapper(env, keys, max_num_agents=A)
    venv = BatchedVecEnv(w)
    obs = venv.reset()
    S = 1 + 2 + 1 + 1 + 1 + 90 + 1
    assert obs.shape == (E, A * S) and obs.dtype == torch.float32
    assert w.observation_indices[1]['rel_goal'] == [S + 1, S + 3] and w.singleAgentObservationInds(2) == [2 * S, 3 * S]
    assert w.keyToArrayInds('use_ppo') == [[a * S + S - 1, a * S + S] for a in range(A)]
    # the dict view of one env's row gives back the per-key observations of the batched env
    d = w.observationArrayToDict(obs[2])
    bo = env.step(torch.full((E, A, 2), 0.5, device=obs.device))[0]
    obs2, rews, dones, infos = venv.step(torch.full((E, A, 2), 0.5, device=obs.device))
    torch.cuda.synchronize()
    assert float(d[0]['use_ppo']) == 1.0 and float(d[1]['use_ppo']) == 0.0 and float(d[2]['radius']) == 0.0
    d2 = w.observationArrayToDict(obs2[4])
    np.testing.assert_allclose(d2[1]['rel_goal'].cpu().numpy(), env.sim.buf["obs_misc"][4, 1, 0:2].cpu().numpy())
    np.testing.assert_allclose(d2[0]['other_agents_states'].cpu().numpy().reshape(9, 10),
                               env.sim.buf["other_agents_states"][4, 0].cpu().numpy())
    f0 = FlattenDictWrapper(env, keys)
    assert f0.observation().shape == (E, S)
    # DummyVecEnv semantics: finished environments restart on their own
    for t in range(400):
        obs2, rews, dones, infos = venv.step(torch.full((E, A, 2), 0.5, device=obs.device))
    torch.cuda.synchronize()
    assert int(env.sim.state[5, :, 0, 1].max().item()) < 400              # step_num was reset at least once
    venv.close()


def test_flatten_full_size_equals_concatenation():
    """BASELINE full size (4096 envs x 10 agents, all three sensors): the kernel's [E, A*S] array equals the
    concatenation of the per-key device buffers (size-independent property; exact)."""
    import torch
    from casim import lib, scenarios as S
    from casim.sim import BatchedSim
    E, A = 4096, 10
    cfg = lib.default_config()
    init, types, n, rects = S.crossing(E, A, seed=3, sensors=L.SENSOR_OTHER_AGENTS | L.SENSOR_LASERSCAN | L.SENSOR_OCCUPANCY_GRID)
    n = n.copy()
    n[::5] = 7                                             # ragged environments: absent slots must be zero
    sim = BatchedSim(E, A, cfg=cfg, sensors=("other_agents_states", "laserscan", "local_grid"))
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    for _ in range(3):
        sim.step()
    keys = ['dist_to_goal', 'rel_goal', 'radius', 'heading_ego_frame', 'heading_global_frame', 'pos_global_frame',
            'pref_speed', 'num_other_agents', 'other_agent_states', 'other_agents_states', 'laserscan', 'local_grid', 'use_ppo']
    flat = sim.flatten_obs(keys).view(E, A, -1)
    st, b = sim.state, sim.buf
    w = lambda k: st[k // 4, :, :, k % 4].unsqueeze(-1)    # noqa: E731
    parts = [w(L.DIST_TO_GOAL), b["obs_misc"][..., 0:2], w(L.RADIUS), w(L.HEADING_EGO), w(L.HEADING), st[0, :, :, 0:2],
             w(L.PREF_SPEED), b["obs_misc"][..., 2:3], b["other_agents_states"][:, :, 0, :],
             b["other_agents_states"].reshape(E, A, -1), b["laserscan"], b["local_grid"].reshape(E, A, -1).float(),
             b["obs_misc"][..., 3:4]]
    want = torch.cat(parts, dim=-1)
    valid = (torch.arange(A, device=want.device)[None, :] < torch.as_tensor(n, device=want.device)[:, None]).unsqueeze(-1)
    want = torch.where(valid, want, torch.zeros_like(want))
    torch.cuda.synchronize()
    assert flat.shape == want.shape == (E, A, sim.flat_row_size(keys))
    assert torch.equal(flat, want)
    sim.close()
