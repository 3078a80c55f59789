"""Synthetic scenario generators (host side, numpy; once per episode, not on the step path).

They produce the tensors ``BatchedSim.reset`` consumes -- ``init [E,A,7]``, ``types [E,A]``, ``n_agents [E]`` --
plus obstacle rectangles, with the shapes and value ranges of the reference's scenario fixtures
(envs/test_cases.py; ``*_agents_500_cases.p`` hold [start_x, start_y, goal_x, goal_y, pref_speed, radius] with
positions within +-16.7 m, speeds 0.53-1.5, radii 0.2-0.8) and of BASELINE.json's configs.
The fixtures themselves live in the reference tree and are not copied; data here is random of that shape.
"""
import numpy as np

from . import layout as L

# envs/test_cases.py:3081-3085 (crossing) and :3150-3154 (hallway): four rectangles each, as
# (x of corner 1, y of corner 1, x of corner 3, y of corner 3) = top-left, bottom-right (Map.py:117-118).
# The last hallway rectangle is listed right-to-left in the reference, so it rasterises to nothing there too.
CROSSING_RECTS = np.array([[2, 10, 10, 2], [-10, 10, -2, 2], [2, -2, 10, -10], [-10, -2, -2, -10]], np.float32)
HALLWAY_RECTS = np.array([[3, 7, 10, -7], [-10, 7, -3, -7], [-10.5, 10, -10, -10], [10.5, 10, 10, -10]], np.float32)


def _f32(x):
    return np.asarray(x, dtype=np.float32)


def open_space(E, A, seed=1234, static_every=5, sensors=0, extent=8.0, policy=L.POLICY_NONCOOPERATIVE,
               dynamics=L.DYN_UNICYCLE):
    """Config C2: E envs x A agents in open space; agent i with i % static_every == static_every-1 is a
    StaticPolicy agent, the rest NonCooperativePolicy, all UnicycleDynamics; heading towards the goal."""
    rng = np.random.default_rng(seed)
    start = rng.uniform(-extent, extent, (E, A, 2))
    ang = rng.uniform(-np.pi, np.pi, (E, A))
    dist = rng.uniform(4.0, 2 * extent, (E, A))
    goal = np.clip(start + dist[..., None] * np.stack([np.cos(ang), np.sin(ang)], -1), -2 * extent, 2 * extent)
    radius = rng.uniform(0.2, 0.8, (E, A))
    pref = rng.uniform(0.53, 1.5, (E, A))
    init = np.concatenate([start, goal, radius[..., None], pref[..., None], np.full((E, A, 1), np.nan)], -1)
    idx = np.arange(A)[None, :].repeat(E, 0)
    is_static = (idx % static_every) == (static_every - 1) if static_every else np.zeros((E, A), bool)
    pol = np.where(is_static, L.POLICY_STATIC, policy)
    types = (pol | (dynamics << 8) | (int(sensors) << 16)).astype(np.int32)
    n_agents = np.full((E,), A, np.int32)
    return _f32(init), types, n_agents


def crossing(E, A, seed=1235, static_every=5, sensors=L.SENSOR_OTHER_AGENTS | L.SENSOR_LASERSCAN,
             policy=L.POLICY_NONCOOPERATIVE, dynamics=L.DYN_UNICYCLE):
    """Config C3: agents start in the free cross of the crossing map (four 8x8 m blocks leaving a +-2 m cross),
    arm uniform in {+x,-x,+y,-y}, along-arm U(3,12), across U(-1.4,1.4), goal = -start."""
    rng = np.random.default_rng(seed)
    arm = rng.integers(0, 4, (E, A))
    along = rng.uniform(3.0, 12.0, (E, A))
    across = rng.uniform(-1.4, 1.4, (E, A))
    sx = np.where(arm == 0, along, np.where(arm == 1, -along, across))
    sy = np.where(arm == 2, along, np.where(arm == 3, -along, across))
    start = np.stack([sx, sy], -1)
    goal = -start
    radius = rng.uniform(0.2, 0.5, (E, A))
    pref = rng.uniform(0.53, 1.5, (E, A))
    init = np.concatenate([start, goal, radius[..., None], pref[..., None], np.full((E, A, 1), np.nan)], -1)
    idx = np.arange(A)[None, :].repeat(E, 0)
    is_static = (idx % static_every) == (static_every - 1) if static_every else np.zeros((E, A), bool)
    pol = np.where(is_static, L.POLICY_STATIC, policy)
    types = (pol | (dynamics << 8) | (int(sensors) << 16)).astype(np.int32)
    n_agents = np.full((E,), A, np.int32)
    return _f32(init), types, n_agents, CROSSING_RECTS.copy()


def exploration(E, seed=1236):
    """Config C4: 4 agents per env: agents 0,1 ig_greedy + FirstOrderDynamics (r = 0.5) on the cross, heading
    along the arm towards the centre; agents 2,3 StaticPolicy targets (r = 0.2) in free space (mirrors
    IG_agent_crossing, envs/test_cases.py:3226-3237)."""
    rng = np.random.default_rng(seed)
    A = 4
    init = np.zeros((E, A, 7))
    types = np.zeros((E, A), np.int32)
    arm = rng.integers(0, 4, (E, 2))
    along = rng.uniform(3.0, 12.0, (E, 2))
    across = rng.uniform(-1.0, 1.0, (E, 2))
    sx = np.where(arm == 0, along, np.where(arm == 1, -along, across))
    sy = np.where(arm == 2, along, np.where(arm == 3, -along, across))
    heading = np.where(arm == 0, np.pi, np.where(arm == 1, 0.0, np.where(arm == 2, -np.pi / 2, np.pi / 2)))
    init[:, :2, 0], init[:, :2, 1] = sx, sy
    init[:, :2, 2], init[:, :2, 3] = -sx, -sy
    init[:, :2, 4], init[:, :2, 5], init[:, :2, 6] = 0.5, 1.0, heading
    types[:, :2] = L.encode_type(L.POLICY_IG_GREEDY, L.DYN_FIRST_ORDER, L.SENSOR_OTHER_AGENTS)
    # targets: uniform in the free cross
    tarm = rng.integers(0, 4, (E, 2))
    talong = rng.uniform(0.0, 12.0, (E, 2))
    tacross = rng.uniform(-1.7, 1.7, (E, 2))
    tx = np.where(tarm == 0, talong, np.where(tarm == 1, -talong, tacross))
    ty = np.where(tarm == 2, talong, np.where(tarm == 3, -talong, tacross))
    init[:, 2:, 0], init[:, 2:, 1] = tx, ty
    init[:, 2:, 2], init[:, 2:, 3] = 0.0, 0.0      # goal (0,0) like the reference's static targets
    init[:, 2:, 4], init[:, 2:, 5], init[:, 2:, 6] = 0.2, 1.0, 0.0
    types[:, 2:] = L.encode_type(L.POLICY_STATIC, L.DYN_UNICYCLE, 0)
    return _f32(init), types, np.full((E,), A, np.int32), CROSSING_RECTS.copy()


def formations(E, A=20, seed=1237):
    """Config C5: starts on a radius-10 circle with U(-0.5,0.5) jitter (cf. gen_circle_test_case,
    envs/test_cases.py:2207-2218), goals on a letter-like polyline layout, all GA3C-CADRL + Unicycle + OtherAgents."""
    rng = np.random.default_rng(seed)
    th = np.linspace(0, 2 * np.pi, A, endpoint=False)[None, :] + rng.uniform(-0.1, 0.1, (E, A))
    start = 10.0 * np.stack([np.cos(th), np.sin(th)], -1) + rng.uniform(-0.5, 0.5, (E, A, 2))
    # goals: A points spread along a per-env zig-zag stroke inside a 12 x 8 m box (letter strokes)
    u = np.linspace(0, 1, A)[None, :]
    phase = rng.uniform(0, 2 * np.pi, (E, 1))
    goal = np.stack([-6 + 12 * u + 0 * phase, 4 * np.sin(3 * np.pi * u + phase)], -1)
    perm = np.argsort(rng.random((E, A)), axis=1)
    goal = np.take_along_axis(goal, perm[..., None], axis=1)
    init = np.concatenate([start, goal, np.full((E, A, 1), 0.5), np.full((E, A, 1), 1.0), np.full((E, A, 1), np.nan)], -1)
    types = np.full((E, A), L.encode_type(L.POLICY_GA3C_CADRL, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS), np.int32)
    return _f32(init), types, np.full((E,), A, np.int32)
