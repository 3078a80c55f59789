"""TEST INFRASTRUCTURE ONLY -- ctypes wrapper around oracle/libcasim_oracle.so.

The oracle is the CPU restatement (plain C, fp64) of the reference stepping path.  It may be
imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs, as the checker or the timed CPU baseline -- never by the product package.
"""
import ctypes
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_PKG = os.path.join(_ROOT, "gym-exploration-2d_b200")
if _PKG not in sys.path:
    sys.path.insert(0, _PKG)

from casim.config import SimConfig, MAP_DIM, NUM_STATS  # noqa: E402  (struct layout only)
from casim import layout as L  # noqa: E402

_LIB_PATH = os.path.join(_HERE, "libcasim_oracle.so")
_lib = None


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith(".c")]
    if force or not os.path.exists(_LIB_PATH) or \
            os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE, "libcasim_oracle.so"])
    return _LIB_PATH


class _Obs(ctypes.Structure):
    _fields_ = [("obs_misc", ctypes.c_void_p), ("other_agents_states", ctypes.c_void_p),
                ("laserscan", ctypes.c_void_p), ("local_grid", ctypes.c_void_p),
                ("rewards", ctypes.c_void_p), ("done", ctypes.c_void_p), ("game_over", ctypes.c_void_p),
                ("collision_flags", ctypes.c_void_p), ("nearest_dist", ctypes.c_void_p),
                ("stats", ctypes.c_void_p)]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.casim_oracle_default_config.argtypes = [ctypes.POINTER(SimConfig)]
        _lib.casim_oracle_default_config.restype = None
    return _lib


def default_config():
    c = SimConfig()
    lib().casim_oracle_default_config(ctypes.byref(c))
    return c


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def rasterize(cfg, rects):
    """Map.get_occupancy_grid: rects [n,4] (x1,y1,x3,y3) -> [300,300] uint8."""
    rects = np.ascontiguousarray(rects, dtype=np.float64).reshape(-1, 4)
    out = np.zeros((MAP_DIM, MAP_DIM), np.uint8)
    lib().casim_oracle_rasterize(ctypes.byref(cfg), _p(rects), ctypes.c_int(len(rects)), _p(out))
    return out


class ObsBuffers:
    """Output arrays of one oracle step (doubles / uint8)."""

    def __init__(self, cfg, E, A):
        rows = cfg.max_other_agents_rows
        self.obs_misc = np.zeros((E, A, 4))
        self.other_agents_states = np.zeros((E, A, rows, 10))
        self.laserscan = np.zeros((E, A, cfg.laser_num_beams))
        self.local_grid = np.zeros((E, A, cfg.submap_w, cfg.submap_h), np.uint8)
        self.rewards = np.zeros((E, A))
        self.done = np.zeros((E, A), np.uint8)
        self.game_over = np.zeros((E,), np.uint8)
        self.collision_flags = np.zeros((E, A), np.uint8)
        self.nearest_dist = np.zeros((E, A))
        self.stats = np.zeros((NUM_STATS,))

    def struct(self):
        s = _Obs()
        for name, _ in _Obs._fields_:
            setattr(s, name, _p(getattr(self, name)))
        return s


def maps_arg(maps):
    if maps is None:
        return None, 0
    maps = np.ascontiguousarray(maps, dtype=np.uint8).reshape(-1, MAP_DIM, MAP_DIM)
    return maps, maps.shape[0]


def reset(cfg, init, types, n_agents=None, maps=None, obs=None):
    """Agent.__init__ for every agent + first _get_obs.  init [E,A,7] f64, types [E,A] i32.
    Returns (state [E,A,24] f64, ObsBuffers)."""
    init = np.ascontiguousarray(init, dtype=np.float64)
    E, A = init.shape[:2]
    types = np.ascontiguousarray(types, dtype=np.int32).reshape(E, A)
    n_agents = None if n_agents is None else np.ascontiguousarray(n_agents, dtype=np.int32)
    state = np.zeros((E, A, L.STATE_WORDS))
    obs = obs or ObsBuffers(cfg, E, A)
    maps, n_maps = maps_arg(maps)
    o = obs.struct()
    rc = lib().casim_oracle_reset(ctypes.byref(cfg), E, A, _p(n_agents), _p(state), _p(init), _p(types),
                                  _p(maps), n_maps, ctypes.byref(o))
    assert rc == 0
    return state, obs


def observe(cfg, state, n_agents=None, maps=None, obs=None):
    state = np.ascontiguousarray(state, dtype=np.float64)
    E, A = state.shape[:2]
    n_agents = None if n_agents is None else np.ascontiguousarray(n_agents, dtype=np.int32)
    obs = obs or ObsBuffers(cfg, E, A)
    maps, n_maps = maps_arg(maps)
    o = obs.struct()
    rc = lib().casim_oracle_observe(ctypes.byref(cfg), E, A, _p(n_agents), _p(state), _p(maps), n_maps,
                                    ctypes.byref(o))
    assert rc == 0
    return obs


def step(cfg, state, ext_actions=None, policy_actions=None, dt=None, n_agents=None, maps=None,
         traj=None, init=None, types0=None, prev_game_over=None, obs=None, n_threads=1):
    """One CollisionAvoidanceEnv.step per environment, in place on ``state`` [E,A,24] f64."""
    assert state.dtype == np.float64 and state.flags.c_contiguous
    E, A = state.shape[:2]
    dt = cfg.dt if dt is None else float(dt)
    f64 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64)  # noqa: E731
    ext_actions, policy_actions, traj, init = f64(ext_actions), f64(policy_actions), f64(traj), f64(init)
    T = 0 if traj is None else traj.shape[2]
    n_agents = None if n_agents is None else np.ascontiguousarray(n_agents, dtype=np.int32)
    types0 = None if types0 is None else np.ascontiguousarray(types0, dtype=np.int32)
    prev_game_over = None if prev_game_over is None else np.ascontiguousarray(prev_game_over, dtype=np.uint8)
    obs = obs or ObsBuffers(cfg, E, A)
    maps, n_maps = maps_arg(maps)
    o = obs.struct()
    rc = lib().casim_oracle_step(ctypes.byref(cfg), E, A, _p(n_agents), _p(state), _p(ext_actions),
                                 _p(policy_actions), ctypes.c_double(dt), _p(maps), n_maps, _p(traj), T,
                                 _p(init), _p(types0), _p(prev_game_over), ctypes.byref(o), int(n_threads))
    assert rc == 0, rc
    return obs


# ------------------------------------------------------------------------------------------
# exploration (casim_oracle_explore.c)
def edt(static_map):
    """edfMap.update: exact squared EDT (cells^2) of one [300,300] map; edfMap.map == sqrt(d2) * 0.1."""
    m = np.ascontiguousarray(static_map, dtype=np.uint8).reshape(MAP_DIM, MAP_DIM)
    d2 = np.zeros((MAP_DIM, MAP_DIM), np.int32)
    lib().casim_oracle_edt(_p(m), _p(d2))
    return d2


def visible_cells(cfg, static_map, pose):
    """targetMap.getVisibleCells -> uint8 mask [dim,dim] indexed [j (y cell), i (x cell)]."""
    m = np.ascontiguousarray(static_map, dtype=np.uint8).reshape(MAP_DIM, MAP_DIM)
    pose = np.ascontiguousarray(pose, dtype=np.float64)
    vis = np.zeros((cfg.ig_map_dim, cfg.ig_map_dim), np.uint8)
    lib().casim_oracle_visible_cells(ctypes.byref(cfg), _p(m), _p(pose), _p(vis))
    return vis


def reward_from_pose(cfg, static_map, pose, belief):
    m = np.ascontiguousarray(static_map, dtype=np.uint8).reshape(MAP_DIM, MAP_DIM)
    pose = np.ascontiguousarray(pose, dtype=np.float64)
    belief = np.ascontiguousarray(belief, dtype=np.float64)
    f = lib().casim_oracle_reward_from_pose
    f.restype = ctypes.c_double
    return float(f(ctypes.byref(cfg), _p(m), _p(pose), _p(belief)))


def explore_rollouts(cfg, static_map, pose, radius, belief, actions, exclude=None, xdt=1):
    """Dec-MCTS simulation rollouts of ONE agent (ig_mcts.py:173-253 driven like DecMCTS._simulate).
    actions [K,H] uint8 (3*vel_idx + dphi_idx, 255 = ended); exclude [dim,dim] bool or None.
    Returns dict(rewards [K], end_pose [K,3], obsvd [K,dim,dim] uint8)."""
    m = np.ascontiguousarray(static_map, dtype=np.uint8).reshape(MAP_DIM, MAP_DIM)
    pose = np.ascontiguousarray(pose, dtype=np.float64)
    belief = np.ascontiguousarray(belief, dtype=np.float64)
    actions = np.ascontiguousarray(actions, dtype=np.uint8)
    K, H = actions.shape
    dim = cfg.ig_map_dim
    ex = None if exclude is None else np.ascontiguousarray(exclude, dtype=np.uint8).reshape(dim, dim)
    rewards = np.zeros(K, np.float64)
    end_pose = np.zeros((K, 3), np.float64)
    obsvd = np.zeros((K, dim, dim), np.uint8)
    n_feasible = np.zeros(K, np.int32)
    f = lib().casim_oracle_explore_rollouts
    f.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double, ctypes.c_void_p,
                  ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                  ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    f.restype = ctypes.c_int
    rc = f(ctypes.addressof(cfg), _p(m), _p(pose), float(radius), _p(belief), _p(actions), K, H, int(xdt),
           None if ex is None else _p(ex), _p(rewards), _p(end_pose), _p(obsvd), _p(n_feasible))
    assert rc == 0
    return dict(rewards=rewards, end_pose=end_pose, obsvd=obsvd, n_feasible=n_feasible)


def team_update(cfg, state, oas, belief, static_map, targets_prev, n_prev, n_agents=None):
    """ig_mcts.update_belief for every explorer (sequenced like collision_avoidance_env._take_action_dmcts).
    belief [E,A,dim,dim] f64 is updated in place.  targets_prev [E,A,32,2] f64, n_prev [E,A] int32.
    Returns dict(targets [E,A,32,2], n_targets [E,A], team_mask [E,A,dim,dim] uint8, team_reward [E,A])."""
    assert state.dtype == np.float64 and belief.dtype == np.float64 and belief.flags.c_contiguous
    E, A = state.shape[:2]
    dim = cfg.ig_map_dim
    m = np.ascontiguousarray(static_map, dtype=np.uint8).reshape(MAP_DIM, MAP_DIM)
    state = np.ascontiguousarray(state)
    oas = np.ascontiguousarray(oas, dtype=np.float64)
    targets_prev = np.ascontiguousarray(targets_prev, dtype=np.float64)
    n_prev = np.ascontiguousarray(n_prev, dtype=np.int32)
    n_agents = None if n_agents is None else np.ascontiguousarray(n_agents, dtype=np.int32)
    targets = np.zeros((E, A, 32, 2), np.float64)
    n_t = np.zeros((E, A), np.int32)
    mask = np.zeros((E, A, dim, dim), np.uint8)
    reward = np.zeros((E, A), np.float64)
    f = lib().casim_oracle_team_update
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 11
    f.restype = ctypes.c_int
    rc = f(ctypes.addressof(cfg), E, A, _p(n_agents), _p(state), _p(oas), _p(belief), _p(m), _p(targets_prev),
           _p(n_prev), _p(targets), _p(n_t), _p(mask), _p(reward))
    assert rc == 0
    return dict(targets=targets, n_targets=n_t, team_mask=mask, team_reward=reward)


def explore_step(cfg, state, oas, belief, maps, n_agents=None, init=None, prev_game_over=None, n_threads=1):
    """ig_greedy.find_next_action for every ig_greedy agent.  belief [E,A,dim,dim] f64 is updated in place.
    Returns dict(policy_actions [E,A,2], best_mi [E,A], cand_mi [E,A,9], counters [4], visible [E,A,dim,dim])."""
    assert state.dtype == np.float64 and belief.dtype == np.float64 and belief.flags.c_contiguous
    E, A = state.shape[:2]
    dim = cfg.ig_map_dim
    oas = np.ascontiguousarray(oas, dtype=np.float64)
    n_agents = None if n_agents is None else np.ascontiguousarray(n_agents, dtype=np.int32)
    init = None if init is None else np.ascontiguousarray(init, dtype=np.float64)
    prev_game_over = None if prev_game_over is None else np.ascontiguousarray(prev_game_over, dtype=np.uint8)
    maps, n_maps = maps_arg(maps)
    out = dict(policy_actions=np.zeros((E, A, 2)), best_mi=np.zeros((E, A)), cand_mi=np.zeros((E, A, 9)),
               counters=np.zeros((4,), np.uint64), visible=np.zeros((E, A, dim, dim), np.uint8))
    rc = lib().casim_oracle_explore_step(ctypes.byref(cfg), E, A, _p(n_agents), _p(np.ascontiguousarray(state)),
                                         _p(oas), _p(belief), _p(maps), n_maps, _p(out["policy_actions"]),
                                         _p(out["best_mi"]), _p(out["cand_mi"]), _p(out["counters"]),
                                         _p(out["visible"]), _p(init), _p(prev_game_over), int(n_threads))
    assert rc == 0, rc
    return out


# ------------------------------------------------------------------------------------------
# GA3C-CADRL (casim_oracle_ga3c.c)
GA3C_NUM_PARAMS = 170764
# (name, shape) of the parameter blob in order; TF variable names of GA3C_CADRL/network.py
GA3C_LAYOUT = [("rnn/lstm_cell/kernel", (71, 256)), ("rnn/lstm_cell/bias", (256,)),
               ("layer1/kernel", (68, 256)), ("layer1/bias", (256,)),
               ("layer2/kernel", (256, 256)), ("layer2/bias", (256,)),
               ("fullyconnected1/kernel", (256, 256)), ("fullyconnected1/bias", (256,)),
               ("logits_p/kernel", (256, 11)), ("logits_p/bias", (11,)),
               ("logits_v/kernel", (256, 1)), ("logits_v/bias", (1,))]


def ga3c_random_weights(seed=0):
    """Random-init weights of the architecture: Glorot-uniform kernels, zero biases (the TF1 defaults of
    tf.layers.dense / LSTMCell), numpy Generator(seed)."""
    rng = np.random.default_rng(seed)
    parts = []
    for name, shape in GA3C_LAYOUT:
        if len(shape) == 2:
            lim = np.sqrt(6.0 / (shape[0] + shape[1]))
            parts.append(rng.uniform(-lim, lim, shape).astype(np.float32).ravel())
        else:
            parts.append(np.zeros(shape, np.float32))
    w = np.concatenate(parts)
    assert w.size == GA3C_NUM_PARAMS
    return w


def ga3c_actions():
    t = np.zeros((11, 2))
    lib().casim_oracle_ga3c_actions(_p(t))
    return t


def ga3c_forward(cfg, state, weights, n_agents=None, n_threads=1):
    """GA3CCADRLPolicy.find_next_action for every GA3C agent.  Returns dict(policy_actions [E,A,2] f64,
    action_idx [E,A] i32, probs [E,A,11] f32, nn_input [E,A,75] f64)."""
    state = np.ascontiguousarray(state, dtype=np.float64)
    E, A = state.shape[:2]
    weights = np.ascontiguousarray(weights, dtype=np.float32)
    n_agents = None if n_agents is None else np.ascontiguousarray(n_agents, dtype=np.int32)
    out = dict(policy_actions=np.zeros((E, A, 2)), action_idx=np.zeros((E, A), np.int32),
               probs=np.zeros((E, A, 11), np.float32), nn_input=np.zeros((E, A, 75)))
    rc = lib().casim_oracle_ga3c_forward(ctypes.byref(cfg), E, A, _p(n_agents), _p(state), _p(weights),
                                         int(weights.size), _p(out["policy_actions"]), _p(out["action_idx"]),
                                         _p(out["probs"]), _p(out["nn_input"]), int(n_threads))
    assert rc == 0, rc
    return out


class RvoParams(ctypes.Structure):
    """casim_rvo_params (include/casim.h): RVOPolicy.py:13-28 defaults."""
    _fields_ = [("neighbor_dist", ctypes.c_double), ("time_horizon", ctypes.c_double), ("radius_scale", ctypes.c_double),
                ("max_delta_heading", ctypes.c_double), ("max_neighbors", ctypes.c_int32), ("reserved", ctypes.c_int32 * 3)]


def rvo_default_params():
    return RvoParams(neighbor_dist=float("inf"), time_horizon=5.0, radius_scale=1 + 15e-2, max_delta_heading=np.pi / 6,
                     max_neighbors=10)


def rvo_actions(cfg, state, coop=None, n_agents=None, params=None, n_threads=1, polygons=None):
    """RVOPolicy.find_next_action for every RVO agent: state [E,A,24] f64 -> actions [E,A,2] f64.
    polygons: the obstacle vertex lists handed to sim.addObstacle (shared by all environments)."""
    state = np.ascontiguousarray(state, dtype=np.float64)
    E, A = state.shape[:2]
    coop = None if coop is None else np.ascontiguousarray(coop, dtype=np.float64)
    n_agents = None if n_agents is None else np.ascontiguousarray(n_agents, dtype=np.int32)
    prm = params or rvo_default_params()
    out = np.zeros((E, A, 2))
    polygons = [] if polygons is None else [np.asarray(p, np.float64).reshape(-1, 2) for p in polygons]
    sizes = np.array([len(p) for p in polygons] or [0], np.int32)
    verts = np.ascontiguousarray(np.concatenate(polygons) if polygons else np.zeros((1, 2)), np.float64)
    rc = lib().casim_oracle_rvo_actions_obst(ctypes.byref(cfg), ctypes.byref(prm), E, A, _p(n_agents), _p(state), _p(coop),
                                             _p(out), int(n_threads), len(polygons), _p(sizes), _p(verts))
    assert rc == 0, rc
    return out


def ga3c_network(weights, x):
    """Network forward alone: x [B,75] f32 -> probs [B,11] f32."""
    weights = np.ascontiguousarray(weights, dtype=np.float32)
    x = np.ascontiguousarray(x, dtype=np.float32)
    probs = np.zeros((x.shape[0], 11), np.float32)
    lib().casim_oracle_ga3c_network(_p(weights), _p(x), int(x.shape[0]), _p(probs))
    return probs


# ------------------------------------------------------------------------------------------
# conversions between the GPU layout (state [6,E,A,4] fp32 with int32 words) and the oracle's
# [E,A,24] doubles
def state_from_packs(packs):
    """packs: numpy float32 [6,E,A,4] (int words bit-cast) -> float64 [E,A,24]."""
    packs = np.ascontiguousarray(packs, dtype=np.float32)
    P, E, A, _ = packs.shape
    w = packs.transpose(1, 2, 0, 3).reshape(E, A, P * 4)
    out = w.astype(np.float64)
    iw = np.ascontiguousarray(w).view(np.int32)
    for k in L.INT_WORDS:
        out[..., k] = iw[..., k]
    return np.ascontiguousarray(out)


def packs_from_state(state):
    """float64 [E,A,24] -> float32 [6,E,A,4] with int32 words bit-cast (values rounded to fp32)."""
    E, A, W = state.shape
    w = state.astype(np.float32)
    iw = w.view(np.int32)
    for k in L.INT_WORDS:
        iw[..., k] = state[..., k].astype(np.int32)
    return np.ascontiguousarray(w.reshape(E, A, W // 4, 4).transpose(2, 0, 1, 3))


# scenario generators (casim_oracle_scenario.c)
SCENARIO_SWAP_CIRCLE, SCENARIO_RANDOM_POSITIONS, SCENARIO_PAIRWISE_SWAP = 0, 1, 2


def scenario(kind, E, A, number_of_agents, seeds, ego_policy, ego_dynamics, other_dynamics=0):
    """train_agents_swap_circle / train_agents_random_positions(seed=seeds[e]) for E environments.
    Returns dict(init [E,A,7] f64 (heading NaN = towards the goal), types [E,A] i32, n_agents [E] i32, coop [E,A])."""
    seeds = np.ascontiguousarray(seeds, dtype=np.uint32)
    assert seeds.shape == (E,) and (seeds != 0).all()
    init = np.zeros((E, A, 7), np.float64)
    types = np.zeros((E, A), np.int32)
    n = np.zeros((E,), np.int32)
    coop = np.zeros((E, A), np.float64)
    f = lib().casim_oracle_scenario
    f.argtypes = [ctypes.c_int] * 4 + [ctypes.c_void_p] + [ctypes.c_int] * 3 + [ctypes.c_void_p] * 4
    f.restype = ctypes.c_int
    rc = f(int(kind), E, A, int(number_of_agents), _p(seeds), int(ego_policy), int(ego_dynamics), int(other_dynamics),
           _p(init), _p(types), _p(n), _p(coop))
    if rc:
        raise ValueError("casim_oracle_scenario: invalid arguments")
    return dict(init=init, types=types, n_agents=n, coop=coop)
