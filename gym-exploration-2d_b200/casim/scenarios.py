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


def _types(E, A, static_every, policy, dynamics, sensors):
    idx = np.arange(A)[None, :].repeat(E, 0)
    is_static = (idx % static_every) == (static_every - 1) if static_every else np.zeros((E, A), bool)
    pol = np.where(is_static, L.POLICY_STATIC, policy)
    return (pol | (dynamics << 8) | (int(sensors) << 16)).astype(np.int32)


def _separated(rng, E, A, draw, min_sep, tries=64):
    """Rejection sampling over the whole batch: agent i's start is redrawn (up to ``tries`` times) while it lies closer
    than ``min_sep`` to an earlier agent's start of the same environment.  ``draw(n)`` returns n candidate points."""
    start = np.zeros((E, A, 2))
    for i in range(A):
        pts = draw(E)
        for _ in range(tries):
            if i == 0:
                break
            d = np.linalg.norm(start[:, :i] - pts[:, None, :], axis=-1).min(axis=1)
            bad = d < min_sep
            if not bad.any():
                break
            pts[bad] = draw(int(bad.sum()))
        start[:, i] = pts
    return start


def from_cases(cases, E, A=None, seed=1234, static_every=5, sensors=0, jitter=0.05, policy=L.POLICY_NONCOOPERATIVE,
               dynamics=L.DYN_UNICYCLE):
    """Config C2 as SURVEY.md 8(d) specifies it: environment e takes case ``e mod len(cases)`` of the reference's
    10-agent scenario suite (``cases`` [n,A,6] = start_x, start_y, goal_x, goal_y, pref_speed, radius: the array of
    envs/test_cases/10_agents_500_cases.p, shipped as tests/golden/c2_10_agents_500_cases.npz) plus N(0, jitter^2)
    position jitter; heading towards the goal (agent.py:29-31); agent i with i % 5 == 4 StaticPolicy, the rest
    NonCooperativePolicy."""
    cases = np.asarray(cases, np.float64)
    A = cases.shape[1] if A is None else A
    rng = np.random.default_rng(seed)
    c = cases[np.arange(E) % len(cases), :A]
    start = c[..., 0:2] + rng.normal(0.0, jitter, (E, A, 2))
    goal = c[..., 2:4] + rng.normal(0.0, jitter, (E, A, 2))
    init = np.concatenate([start, goal, c[..., 5:6], c[..., 4:5], np.full((E, A, 1), np.nan)], -1)
    return _f32(init), _types(E, A, static_every, policy, dynamics, sensors), np.full((E,), A, np.int32)


def open_space(E, A, seed=1234, static_every=5, sensors=0, extent=8.0, policy=L.POLICY_NONCOOPERATIVE,
               dynamics=L.DYN_UNICYCLE, min_sep=1.7):
    """C2-shaped synthetic stand-in for when the reference's case suite is not at hand: value ranges of the fixture
    (speeds 0.53-1.5, radii 0.2-0.8), starts at least ``min_sep`` apart (the fixture's closest pair of starts leaves a
    0.2 m gap), goals 4-16 m away."""
    rng = np.random.default_rng(seed)
    start = _separated(rng, E, A, lambda k: rng.uniform(-extent, extent, (k, 2)), min_sep)
    ang = rng.uniform(-np.pi, np.pi, (E, A))
    dist = rng.uniform(4.0, 2 * extent, (E, A))
    goal = np.clip(start + dist[..., None] * np.stack([np.cos(ang), np.sin(ang)], -1), -2 * extent, 2 * extent)
    radius = rng.uniform(0.2, 0.8, (E, A))
    pref = rng.uniform(0.53, 1.5, (E, A))
    init = np.concatenate([start, goal, radius[..., None], pref[..., None], np.full((E, A, 1), np.nan)], -1)
    return _f32(init), _types(E, A, static_every, policy, dynamics, sensors), np.full((E,), A, np.int32)


def crossing(E, A, seed=1235, static_every=5, sensors=L.SENSOR_OTHER_AGENTS | L.SENSOR_LASERSCAN,
             policy=L.POLICY_NONCOOPERATIVE, dynamics=L.DYN_UNICYCLE, min_sep=1.2):
    """Config C3: agents start in the free cross of the crossing map (four 8x8 m blocks leaving a
    +-2 m cross), arm uniform in {+x,-x,+y,-y}, along-arm U(3,12), across U(-1.4,1.4), goal = -start, a start is
    redrawn while it lies < 1.2 m from an earlier start of its environment."""
    rng = np.random.default_rng(seed)

    def draw(k):
        arm = rng.integers(0, 4, k)
        along = rng.uniform(3.0, 12.0, k)
        across = rng.uniform(-1.4, 1.4, k)
        sx = np.where(arm == 0, along, np.where(arm == 1, -along, across))
        sy = np.where(arm == 2, along, np.where(arm == 3, -along, across))
        return np.stack([sx, sy], -1)

    start = _separated(rng, E, A, draw, min_sep)
    goal = -start
    radius = rng.uniform(0.2, 0.5, (E, A))
    pref = rng.uniform(0.53, 1.5, (E, A))
    init = np.concatenate([start, goal, radius[..., None], pref[..., None], np.full((E, A, 1), np.nan)], -1)
    return _f32(init), _types(E, A, static_every, policy, dynamics, sensors), np.full((E,), A, np.int32), CROSSING_RECTS.copy()


def per_env_crossing_rects(E, seed=77):
    """C3's second data point ("and, as a second data point, per-env maps"): one crossing map per
    environment, corridor half-width drawn from U(1.9, 2.6) m and block size from U(7, 9) m -> rects [E,4,4]."""
    rng = np.random.default_rng(seed)
    hw = rng.uniform(1.9, 2.6, (E, 1))
    sz = rng.uniform(7.0, 9.0, (E, 1))
    lo, hi = hw, hw + sz
    z = np.concatenate
    r = np.stack([z([lo, hi, hi, lo], 1), z([-hi, hi, -lo, lo], 1), z([lo, -lo, hi, -hi], 1), z([-hi, -lo, -lo, -hi], 1)], 1)
    return r.astype(np.float32)


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


# Goal points of the reference's formation() letters (envs/test_cases.py:1916-1957), joined into pen strokes.
_LETTER_STROKES = {
    "C": [[(1.5, 3.0), (0.0, 3.0), (-0.5, 2.0), (-0.5, 1.0), (0.0, 0.0), (1.5, 0.0)]],
    "A": [[(-1.5, 0.0), (-0.75, 1.5), (0.0, 3.0), (0.75, 1.5), (1.5, 0.0)], [(-0.75, 1.5), (0.0, 1.5), (0.75, 1.5)]],
    "D": [[(0.0, 0.0), (0.0, 1.5), (0.0, 3.0), (1.2, 2.5), (1.5, 1.5), (1.2, 0.5), (0.0, 0.0)]],
    "R": [[(0.0, 0.0), (0.0, 1.5), (0.0, 3.0), (1.3, 2.8), (1.2, 1.7), (0.0, 1.5)], [(1.2, 1.7), (1.7, 0.0)]],
    "L": [[(0.0, 3.0), (0.0, 2.0), (0.0, 1.0), (0.0, 0.0), (0.75, 0.0), (1.5, 0.0)]],
}


def letter_goals(letter, n, min_sep=1.05, max_extent=8.0):
    """``n`` points at equal arc-length steps along the strokes of one formation() letter, centred on the origin.  The
    reference's letters are 2 x the table above and hold 6 goals; for n agents of radius 0.5 the letter is scaled up
    until no two goals are closer than ``min_sep`` -- as far as it still fits inside the radius-10 start circle
    (``max_extent``); beyond that neighbouring goals overlap and those agents end in a collision or a timeout."""
    strokes = [2.0 * np.asarray(st, np.float64) for st in _LETTER_STROKES[letter]]
    seg = [(a, b) for st in strokes for a, b in zip(st[:-1], st[1:])]
    lens = np.array([np.linalg.norm(b - a) for a, b in seg])
    ends = np.cumsum(lens)
    pts = []
    for u in (np.arange(n) + 0.5) / n * ends[-1]:
        k = min(int(np.searchsorted(ends, u)), len(seg) - 1)
        t = (u - (ends[k] - lens[k])) / lens[k]
        pts.append(seg[k][0] + t * (seg[k][1] - seg[k][0]))
    pts = np.array(pts)
    pts -= 0.5 * (pts.min(0) + pts.max(0))
    d = np.linalg.norm(pts[:, None] - pts[None], axis=-1) + 1e9 * np.eye(n)
    scale = max(1.0, min_sep / d.min())
    scale = min(scale, max_extent / np.abs(pts).max())
    return pts * scale


def formations(E, A=20, seed=1237):
    """Config C5: starts on a radius-10 circle with U(-0.5,0.5) jitter (cf. gen_circle_test_case,
    envs/test_cases.py:2207-2218), goals = A points along the strokes of the letter "CADRL"[e mod 5] of formation()
    (envs/test_cases.py:1916-1957) assigned in a random order (np.random.shuffle(agent_inds), :1960-1961); r = 0.5,
    v_pref = 1.0, all GA3C-CADRL + Unicycle + OtherAgentsStates."""
    rng = np.random.default_rng(seed)
    th = np.linspace(0, 2 * np.pi, A, endpoint=False)[None, :] + rng.uniform(-0.1, 0.1, (E, A))
    start = 10.0 * np.stack([np.cos(th), np.sin(th)], -1) + rng.uniform(-0.5, 0.5, (E, A, 2))
    table = np.stack([letter_goals(ch, A) for ch in "CADRL"])            # [5, A, 2]
    goal = table[np.arange(E) % 5]
    perm = np.argsort(rng.random((E, A)), axis=1)
    goal = np.take_along_axis(goal, perm[..., None], axis=1)
    init = np.concatenate([start, goal, np.full((E, A, 1), 0.5), np.full((E, A, 1), 1.0), np.full((E, A, 1), np.nan)], -1)
    types = np.full((E, A), L.encode_type(L.POLICY_GA3C_CADRL, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS), np.int32)
    return _f32(init), types, np.full((E,), A, np.int32)
