"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by executing the UNMODIFIED reference
(/root/reference, behind oracle/ref_harness.py) in the build container.

    python oracle/gen_golden.py [core|sensors|learning|c1|reset|maps|explore|rollouts|mcts|mcts_env|team|scenarios|ga3c|all]

The fixtures are committed; this script is the provenance.  Inputs are float32-representable
doubles (the GPU state is fp32) and headings / dt are passed as np.float64 so that numpy 2
(NEP 50) keeps the reference's arithmetic in float64 like the numpy 1.x it was written for
(see DESIGN.md "NEP-50 note").
"""
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
sys.path.insert(0, _ROOT)
sys.path.insert(0, os.path.join(_ROOT, "gym-exploration-2d_b200"))

from oracle import ref_harness as rh  # noqa: E402
from casim import layout as L  # noqa: E402

GOLDEN = os.path.join(_ROOT, "tests", "golden")
A_MAX = 10

CROSSING = [[(10, 10), (2, 10), (2, 2), (10, 2)], [(-2, 10), (-10, 10), (-10, 2), (-2, 2)],
            [(10, -2), (2, -2), (2, -10), (10, -10)], [(-2, -2), (-10, -2), (-10, -10), (-2, -10)]]
HALLWAY = [[(10, 7), (3, 7), (3, -7), (10, -7)], [(-3, 7), (-10, 7), (-10, -7), (-3, -7)],
           [(-10, 10), (-10.5, 10), (-10.5, -10), (-10, -10)], [(10, 10), (10.5, 10), (10.5, -10), (10, -10)]]
OBSTACLE_SETS = {0: [], 1: CROSSING, 2: HALLWAY}


def rects_of(obstacles):
    """reference obstacle (4 corners; corner 1 and 3 used, Map.py:117-118) -> [n,4] x1,y1,x3,y3"""
    return np.array([[o[1][0], o[1][1], o[3][0], o[3][1]] for o in obstacles], dtype=np.float64).reshape(-1, 4)


def f32(x):
    return np.asarray(x, dtype=np.float32).astype(np.float64)


class _Traj:
    def __init__(self, pose_vec, vel_vec):
        self.pose_vec = pose_vec
        self.vel_vec = vel_vec


def policy_classes(ns):
    class PassThroughExternalPolicy(ns.ExternalPolicy):
        """A user plugin on the reference's ExternalPolicy API: the caller's action is used as is."""

        def convert_to_action(self, a):
            return np.asarray(a)

    from gym_collision_avoidance.envs.policies.PedestrianDatasetPolicy import PedestrianDatasetPolicy
    from gym_collision_avoidance.envs.policies.LearningPolicy import LearningPolicy

    class CallSiteLearningPolicy(LearningPolicy):
        """The fork's call site passes (agent_index, agents, actions) (collision_avoidance_env.py:304) while the
        class it ships still takes (agent, network_output) (LearningPolicy.py:11): the plain class raises TypeError
        inside env.step.  This adapter only re-routes the arguments; the arithmetic is the reference's own method."""

        def network_output_to_action(self, agent_index, agents, actions):
            return LearningPolicy.network_output_to_action(self, agents[agent_index], actions[agent_index])

    from gym_collision_avoidance.envs.dynamics.UnicycleDynamicsMaxTurnRate import UnicycleDynamicsMaxTurnRate
    from gym_collision_avoidance.envs.dynamics.UnicycleDynamicsMaxAcc import UnicycleDynamicsMaxAcc
    from gym_collision_avoidance.envs.dynamics.UnicycleSecondOrderEulerDynamics import UnicycleSecondOrderEulerDynamics
    pol = {L.POLICY_STATIC: ns.StaticPolicy, L.POLICY_NONCOOPERATIVE: ns.NonCooperativePolicy,
           L.POLICY_EXTERNAL: PassThroughExternalPolicy, L.POLICY_CARRL: ns.CARRLPolicy,
           L.POLICY_PEDESTRIAN: PedestrianDatasetPolicy, L.POLICY_LEARNING: CallSiteLearningPolicy}
    dyn = {L.DYN_UNICYCLE: ns.UnicycleDynamics, L.DYN_FIRST_ORDER: ns.FirstOrderDynamics,
           L.DYN_EXTERNAL: ns.ExternalDynamics, L.DYN_UNICYCLE_MAX_TURN_RATE: UnicycleDynamicsMaxTurnRate,
           L.DYN_UNICYCLE_MAX_ACC: UnicycleDynamicsMaxAcc, L.DYN_UNICYCLE_SECOND_ORDER_EULER: UnicycleSecondOrderEulerDynamics}
    return pol, dyn


def sensor_list(ns, mask):
    s = []
    if mask & L.SENSOR_OTHER_AGENTS:
        s.append(ns.OtherAgentsStatesSensor)
    if mask & L.SENSOR_LASERSCAN:
        s.append(ns.LaserScanSensor)
    if mask & L.SENSOR_OCCUPANCY_GRID:
        s.append(ns.OccupancyGridSensor)
    return s


def build_agents(ns, init, types, traj=None):
    pol, dyn = policy_classes(ns)
    agents = []
    for i in range(len(init)):
        p, d, m = L.decode_type(types[i])
        h = init[i, 6]
        with rh.quiet():
            ag = ns.Agent(float(init[i, 0]), float(init[i, 1]), float(init[i, 2]), float(init[i, 3]),
                          float(init[i, 4]), float(init[i, 5]), None if np.isnan(h) else np.float64(h),
                          pol[p], dyn[d], sensor_list(ns, m), i)
        ag.max_heading_change = 4          # collision_avoidance_env.py:97,467 (set by _init_agents, which the harness bypasses)
        if d == L.DYN_EXTERNAL and traj is not None:
            tr = traj[i]
            ag.policy.trajectory = _Traj(np.stack([tr[:, 0], tr[:, 1], np.zeros(len(tr))], 1),
                                         np.stack([tr[:, 2], tr[:, 3], np.zeros(len(tr))], 1))
        agents.append(ag)
    return agents


def inject_state(ag, s):
    """Overwrite the attributes the stepping path reads with the 24-word record ``s``."""
    ag.pos_global_frame = np.array([s[L.PX], s[L.PY]], dtype=np.float64)
    ag.heading_global_frame = np.float64(s[L.HEADING])
    ag.speed_global_frame = np.float64(s[L.SPEED])
    ag.vel_global_frame = np.array([s[L.VX], s[L.VY]], dtype=np.float64)
    ag.delta_heading_global_frame = np.float64(s[L.DELTA_HEADING])
    ag.turning_dir = np.float64(s[L.TURNING_DIR])
    # the TURNING_DIR / SPEED words double as the extra dynamics state (only one dynamics model owns an agent):
    # UnicycleDynamicsMaxAcc.current_turning_rate / current_speed, Agent.angular_speed_global_frame
    if hasattr(ag.dynamics_model, "current_turning_rate"):
        ag.dynamics_model.current_turning_rate = np.float64(s[L.TURNING_DIR])
        ag.dynamics_model.current_speed = np.float64(s[L.SPEED])
    if type(ag.dynamics_model).__name__ == "UnicycleSecondOrderEulerDynamics":
        ag.angular_speed_global_frame = np.float64(s[L.TURNING_DIR])
    ag.goal_global_frame = np.array([s[L.GX], s[L.GY]], dtype=np.float64)
    ag.radius = float(s[L.RADIUS])
    ag.pref_speed = float(s[L.PREF_SPEED])
    ag.t = np.float64(s[L.T])
    ag.time_remaining_to_reach_goal = np.float64(s[L.TIME_REMAINING])
    ag.dist_to_goal = np.float64(s[L.DIST_TO_GOAL])
    ag.past_dist_to_goal = np.float64(s[L.PAST_DIST_TO_GOAL])
    ag.past_actions = np.array([[s[L.PAST_ACT0_V], s[L.PAST_ACT0_W]], [s[L.PAST_ACT1_V], s[L.PAST_ACT1_W]]])
    ag.heading_ego_frame = np.float64(s[L.HEADING_EGO])
    ag.step_num = int(s[L.STEP_NUM])
    fl = int(s[L.FLAGS])
    ag.is_at_goal = bool(fl & L.F_IS_AT_GOAL)
    ag.was_at_goal_already = bool(fl & L.F_WAS_AT_GOAL_ALREADY)
    ag.in_collision = bool(fl & L.F_IN_COLLISION)
    ag.was_in_collision_already = bool(fl & L.F_WAS_IN_COLLISION_ALREADY)
    ag.ran_out_of_time = bool(fl & L.F_RAN_OUT_OF_TIME)
    ag.is_done = bool(fl & L.F_IS_DONE)
    ag.is_infeasible = bool(fl & L.F_IS_INFEASIBLE)
    # derived, not stored: ego axes as left by the last update_ego_frame (agent.py:250-269)
    gd = ag.goal_global_frame - ag.pos_global_frame
    d = np.sqrt(gd[0] ** 2 + gd[1] ** 2)
    ag.ref_prll = gd / d if d > 1e-8 else gd
    ag.ref_orth = np.array([-ag.ref_prll[1], ag.ref_prll[0]])
    ag.rel_goal = gd


def extract_state(ag, type_word):
    s = np.zeros(L.STATE_WORDS)
    s[L.PX], s[L.PY] = ag.pos_global_frame
    s[L.HEADING] = ag.heading_global_frame
    s[L.SPEED] = ag.speed_global_frame
    s[L.VX], s[L.VY] = ag.vel_global_frame
    s[L.DELTA_HEADING] = ag.delta_heading_global_frame
    s[L.TURNING_DIR] = ag.turning_dir
    if hasattr(ag.dynamics_model, "current_turning_rate"):
        s[L.TURNING_DIR] = ag.dynamics_model.current_turning_rate
        assert ag.dynamics_model.current_speed == ag.speed_global_frame
    if type(ag.dynamics_model).__name__ == "UnicycleSecondOrderEulerDynamics":
        s[L.TURNING_DIR] = ag.angular_speed_global_frame
    s[L.GX], s[L.GY] = ag.goal_global_frame
    s[L.RADIUS] = ag.radius
    s[L.PREF_SPEED] = ag.pref_speed
    s[L.T] = ag.t
    s[L.TIME_REMAINING] = ag.time_remaining_to_reach_goal
    s[L.DIST_TO_GOAL] = ag.dist_to_goal
    s[L.PAST_DIST_TO_GOAL] = ag.past_dist_to_goal
    s[L.PAST_ACT0_V], s[L.PAST_ACT0_W] = ag.past_actions[0]
    s[L.PAST_ACT1_V], s[L.PAST_ACT1_W] = ag.past_actions[1]
    s[L.HEADING_EGO] = ag.heading_ego_frame
    s[L.STEP_NUM] = ag.step_num
    fl = 0
    fl |= L.F_IS_AT_GOAL if ag.is_at_goal else 0
    fl |= L.F_WAS_AT_GOAL_ALREADY if ag.was_at_goal_already else 0
    fl |= L.F_IN_COLLISION if ag.in_collision else 0
    fl |= L.F_WAS_IN_COLLISION_ALREADY if ag.was_in_collision_already else 0
    fl |= L.F_RAN_OUT_OF_TIME if ag.ran_out_of_time else 0
    fl |= L.F_IS_DONE if ag.is_done else 0
    fl |= L.F_IS_INFEASIBLE if ag.is_infeasible else 0
    s[L.FLAGS] = fl
    s[L.TYPE] = type_word
    return s


STATES_ALL = ['dist_to_goal', 'rel_goal', 'radius', 'heading_ego_frame', 'heading_global_frame',
              'pos_global_frame', 'pref_speed', 'num_other_agents', 'other_agents_states', 'laserscan',
              'local_grid']


def states_for(mask_union):
    st = ['dist_to_goal', 'rel_goal', 'radius', 'heading_ego_frame', 'heading_global_frame',
          'pos_global_frame', 'pref_speed']
    if mask_union & L.SENSOR_OTHER_AGENTS:
        st += ['other_agents_states']
    if mask_union & L.SENSOR_LASERSCAN:
        st += ['laserscan']
    if mask_union & L.SENSOR_OCCUPANCY_GRID:
        st += ['local_grid']
    return st


def set_mode(ns, mode):
    ns.Config.TRAIN_SINGLE_AGENT = False
    if mode == 0:
        ns.Config.EVALUATE_MODE, ns.Config.HOMOGENEOUS_TESTING = True, False
    elif mode == 1:
        ns.Config.EVALUATE_MODE, ns.Config.HOMOGENEOUS_TESTING = True, True
    else:
        ns.Config.EVALUATE_MODE, ns.Config.HOMOGENEOUS_TESTING = False, False


class Recorder:
    """Collects reference outputs of one step/observe into fixed-size arrays."""

    def __init__(self, ncases, A=A_MAX):
        z = lambda *s, dt=np.float64: np.zeros((ncases, A) + s, dt)  # noqa: E731
        self.state_out = z(L.STATE_WORDS)
        self.rewards = z()
        self.done = z(dt=np.uint8)
        self.game_over = np.zeros((ncases,), np.uint8)
        self.collision_flags = z(dt=np.uint8)
        self.nearest_dist = z()
        self.obs_misc = z(4)
        self.other_agents_states = z(9, 10)
        self.laserscan = z(16)
        self.local_grid = z(60, 60, dt=np.uint8)
        self.obs_dist_to_goal = z()
        self.obs_heading_ego = z()

    def record_obs(self, k, agents, obs, types):
        for i, ag in enumerate(agents):
            o = obs[i]
            p, d, m = L.decode_type(types[i])
            self.state_out[k, i] = extract_state(ag, types[i])
            self.obs_misc[k, i, 0:2] = o['rel_goal']
            self.obs_misc[k, i, 2] = getattr(ag, 'num_other_agents_observed', 0) if m & L.SENSOR_OTHER_AGENTS else 0
            self.obs_misc[k, i, 3] = float(ag.get_agent_data_equiv("policy.str", "learning"))   # config.py:209-214 use_ppo
            self.obs_dist_to_goal[k, i] = o['dist_to_goal']
            self.obs_heading_ego[k, i] = o['heading_ego_frame']
            if m & L.SENSOR_OTHER_AGENTS:
                self.other_agents_states[k, i] = o['other_agents_states']
            if m & L.SENSOR_LASERSCAN:
                self.laserscan[k, i] = o['laserscan']
            if m & L.SENSOR_OCCUPANCY_GRID:
                self.local_grid[k, i] = o['local_grid'].astype(np.uint8)

    def arrays(self):
        return dict(self.__dict__)


DYN_CHOICES = [L.DYN_UNICYCLE, L.DYN_UNICYCLE, L.DYN_FIRST_ORDER, L.DYN_UNICYCLE_MAX_TURN_RATE]
POLICY_CHOICES = [L.POLICY_STATIC, L.POLICY_NONCOOPERATIVE, L.POLICY_NONCOOPERATIVE, L.POLICY_NONCOOPERATIVE,
                  L.POLICY_EXTERNAL, L.POLICY_CARRL, L.POLICY_PEDESTRIAN]


def random_case(rng, sensors_mode):
    """One random mid-episode environment.  Returns init, types, state_in, ext, traj, n, map_id."""
    n = int(rng.integers(2, A_MAX + 1))
    map_id = int(rng.integers(0, 3)) if sensors_mode else int(rng.integers(0, 2))
    init = np.zeros((A_MAX, 7))
    types = np.zeros((A_MAX,), np.int32)
    state = np.zeros((A_MAX, L.STATE_WORDS))
    ext = np.zeros((A_MAX, 2))
    T = 6
    traj = np.zeros((A_MAX, T, 4))
    spread = float(rng.choice([2.0, 4.0, 8.0, 14.0]))
    for i in range(n):
        pol = int(rng.choice(POLICY_CHOICES))
        if pol == L.POLICY_PEDESTRIAN:
            dyn = L.DYN_EXTERNAL
        else:
            dyn = int(rng.choice(DYN_CHOICES))
        if sensors_mode:
            mask = int(rng.choice([1, 3, 5, 7, 2, 4, 0]))
        else:
            mask = int(rng.choice([0, 1, 1]))
        types[i] = L.encode_type(pol, dyn, mask)
        start = f32(rng.uniform(-spread, spread, 2))
        goal = f32(rng.uniform(-spread, spread, 2))
        radius = float(f32(rng.uniform(0.2, 0.8)))
        pref = float(f32(rng.uniform(0.5, 1.5)))
        init[i] = [start[0], start[1], goal[0], goal[1], radius, pref, float(f32(rng.uniform(-np.pi, np.pi)))]
        s = state[i]
        # current position: somewhere between start and goal, or clustered for collisions
        if rng.random() < 0.5:
            pos = f32(start + (goal - start) * rng.uniform(0, 1.05) + rng.normal(0, 0.3, 2))
        else:
            pos = f32(rng.normal(0, 1.2, 2))
        s[L.PX], s[L.PY] = pos
        s[L.GX], s[L.GY] = goal
        s[L.RADIUS], s[L.PREF_SPEED] = radius, pref
        s[L.HEADING] = float(f32(rng.uniform(-np.pi, np.pi)))
        sp = float(f32(rng.uniform(0, pref)))
        s[L.SPEED] = sp
        s[L.VX] = float(f32(sp * np.cos(s[L.HEADING])))
        s[L.VY] = float(f32(sp * np.sin(s[L.HEADING])))
        s[L.DELTA_HEADING] = float(f32(rng.uniform(-0.5, 0.5)))
        s[L.TURNING_DIR] = float(f32(rng.choice([0.0, 0.11, -0.11, 0.05, -0.3, 1.0]) * rng.uniform(0.5, 1)))
        step_num = int(rng.integers(0, 60))
        if rng.random() < 0.15:
            step_num = 0
        s[L.STEP_NUM] = step_num
        s[L.T] = float(f32(step_num * 0.1))
        s[L.TIME_REMAINING] = float(f32(rng.choice([rng.uniform(0.5, 30), rng.uniform(-0.05, 0.25)])))
        gd = goal - pos
        s[L.DIST_TO_GOAL] = float(f32(np.hypot(gd[0], gd[1])))
        s[L.PAST_DIST_TO_GOAL] = float(f32(s[L.DIST_TO_GOAL] + rng.uniform(0, 0.15)))
        s[L.PAST_ACT0_V], s[L.PAST_ACT0_W] = f32(rng.uniform(0, 1.5)), f32(rng.uniform(-1, 1))
        s[L.PAST_ACT1_V], s[L.PAST_ACT1_W] = f32(rng.uniform(0, 1.5)), f32(rng.uniform(-1, 1))
        ref = np.arctan2(gd[1], gd[0])
        he = s[L.HEADING] - ref
        he = (he + np.pi) % (2 * np.pi) - np.pi
        s[L.HEADING_EGO] = float(f32(he))
        fl = 0
        u = rng.random()
        if u < 0.12:
            fl = L.F_IS_AT_GOAL | L.F_IS_DONE
            if rng.random() < 0.5:
                fl |= L.F_WAS_AT_GOAL_ALREADY
            # put it at its goal
            pos = f32(goal + rng.uniform(-0.3, 0.3, 2))
            s[L.PX], s[L.PY] = pos
        elif u < 0.24:
            fl = L.F_IN_COLLISION | L.F_IS_DONE
            if rng.random() < 0.6:
                fl |= L.F_WAS_IN_COLLISION_ALREADY
            if rng.random() < 0.3:
                fl |= L.F_RAN_OUT_OF_TIME
        elif u < 0.32:
            fl = L.F_RAN_OUT_OF_TIME | L.F_IS_DONE
        s[L.FLAGS] = fl
        s[L.TYPE] = types[i]
        if pol == L.POLICY_CARRL:
            ext[i] = [int(rng.integers(0, 11)), 0]
        elif pol == L.POLICY_LEARNING:
            ext[i] = [float(f32(rng.uniform(0, 1))), float(f32(rng.uniform(0, 1)))]
        else:
            ext[i] = [float(f32(rng.uniform(0, 1.5))), float(f32(rng.uniform(-1.2, 1.2)))]
        p0 = pos + rng.normal(0, 0.2, 2)
        for k in range(T):
            v = rng.uniform(-1, 1, 2)
            traj[i, k] = f32([p0[0] + 0.1 * k * v[0], p0[1] + 0.1 * k * v[1], v[0], v[1]])
    return init, types, state, ext, traj, n, map_id


def gen_single_step(ns, name, ncases, seed, sensors_mode):
    rng = np.random.default_rng(seed)
    rec = Recorder(ncases)
    init_a = np.zeros((ncases, A_MAX, 7))
    types_a = np.zeros((ncases, A_MAX), np.int32)
    state_in = np.zeros((ncases, A_MAX, L.STATE_WORDS))
    ext_a = np.zeros((ncases, A_MAX, 2))
    traj_a = np.zeros((ncases, A_MAX, 6, 4))
    n_a = np.zeros((ncases,), np.int32)
    map_id = np.zeros((ncases,), np.int32)
    mode_a = np.zeros((ncases,), np.int32)
    for k in range(ncases):
        init, types, state, ext, traj, n, mid = random_case(rng, sensors_mode)
        mode = int(rng.integers(0, 3))
        set_mode(ns, mode)
        init_a[k], types_a[k], state_in[k], ext_a[k], traj_a[k], n_a[k], map_id[k], mode_a[k] = \
            init, types, state, ext, traj, n, mid, mode
        agents = build_agents(ns, init[:n], types[:n], traj[:n])
        union = 0
        for i in range(n):
            inject_state(agents[i], state[i])
            union |= L.decode_type(types[i])[2]
        env = rh.make_env(ns, agents, OBSTACLE_SETS[mid], states_for(union))
        captured = {}
        _cc = env._check_for_collisions

        def cc():
            r = _cc()
            captured['r'] = r
            return r

        env._check_for_collisions = cc
        actions = {}
        for i in range(n):
            p = L.decode_type(types[i])[0]
            if p == L.POLICY_CARRL:
                actions[i] = int(ext[i, 0])
            elif p in (L.POLICY_EXTERNAL, L.POLICY_LEARNING):
                actions[i] = np.array(ext[i])
        with rh.quiet():
            obs, rew, go, info = env.step(actions, dt=np.float64(0.1))
        rec.record_obs(k, agents, obs, types)
        rec.rewards[k, :n] = rew
        rec.done[k, :n] = [info['which_agents_done'][i] for i in range(n)]
        rec.game_over[k] = bool(go)
        ca, cw, _, near = captured['r']
        rec.collision_flags[k, :n] = np.array(ca, np.uint8) | (np.array(cw, np.uint8) << 1)
        rec.nearest_dist[k, :n] = near
    out = rec.arrays()
    out.update(init=init_a, types=types_a, state_in=state_in, ext_actions=ext_a, traj=traj_a, n_agents=n_a,
               map_id=map_id, game_over_mode=mode_a)
    np.savez_compressed(os.path.join(GOLDEN, name), **out)
    print("wrote", name, ncases, "cases")


def gen_reset(ns, ncases=40, seed=77):
    rng = np.random.default_rng(seed)
    rec = Recorder(ncases)
    init_a = np.zeros((ncases, A_MAX, 7))
    types_a = np.zeros((ncases, A_MAX), np.int32)
    n_a = np.zeros((ncases,), np.int32)
    map_id = np.zeros((ncases,), np.int32)
    set_mode(ns, 2)
    for k in range(ncases):
        init, types, _, _, _, n, mid = random_case(rng, True)
        for i in range(n):
            p, d, m = L.decode_type(types[i])
            if d == L.DYN_EXTERNAL:
                types[i] = L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, m)
            if rng.random() < 0.4:
                init[i, 6] = np.nan
            if rng.random() < 0.1:  # starts within the goal radius -> non-positive time budget
                init[i, 2:4] = f32(init[i, 0:2] + rng.uniform(-0.3, 0.3, 2))
        init_a[k], types_a[k], n_a[k], map_id[k] = init, types, n, mid
        agents = build_agents(ns, init[:n], types[:n])
        union = 0
        for i in range(n):
            union |= L.decode_type(types[i])[2]
        env = rh.make_env(ns, agents, OBSTACLE_SETS[mid], states_for(union))
        rec.record_obs(k, agents, env.observation, types)
    out = rec.arrays()
    out.update(init=init_a, types=types_a, n_agents=n_a, map_id=map_id)
    np.savez_compressed(os.path.join(GOLDEN, "reset.npz"), **out)
    print("wrote reset.npz")


def gen_c1(ns):
    """Config C1: get_testcase_two_agents geometry, 100 free-running steps."""
    for tag, heading in (("c1_two_agents", np.float64(0.5)), ("c1_two_agents_nep50", 0.5)):
        set_mode(ns, 2)
        with rh.quiet():
            agents = [ns.Agent(-3, -3, 3, 3, 0.5, 1.0, heading, ns.NonCooperativePolicy, ns.UnicycleDynamics,
                               [ns.OtherAgentsStatesSensor], 0),
                      ns.Agent(3, 3, -3, -3, 0.5, 1.0, heading, ns.NonCooperativePolicy, ns.UnicycleDynamics,
                               [ns.OtherAgentsStatesSensor], 1)]
        types = np.array([L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS)] * 2,
                         np.int32)
        env = rh.make_env(ns, agents, [], states_for(1))
        T = 100
        states = np.zeros((T + 1, 2, L.STATE_WORDS))
        rewards = np.zeros((T, 2))
        done = np.zeros((T, 2), np.uint8)
        game_over = np.zeros((T,), np.uint8)
        oas = np.zeros((T + 1, 2, 9, 10))
        for i in range(2):
            states[0, i] = extract_state(agents[i], types[i])
            oas[0, i] = env.observation[i]['other_agents_states']
        for t in range(T):
            dt = np.float64(0.1) if tag == "c1_two_agents" else None
            with rh.quiet():
                obs, rew, go, info = env.step({}, dt=dt)
            rewards[t] = rew
            done[t] = [info['which_agents_done'][i] for i in range(2)]
            game_over[t] = bool(go)
            for i in range(2):
                states[t + 1, i] = extract_state(agents[i], types[i])
                oas[t + 1, i] = obs[i]['other_agents_states']
        init = np.array([[-3, -3, 3, 3, 0.5, 1.0, 0.5], [3, 3, -3, -3, 0.5, 1.0, 0.5]], dtype=np.float64)
        np.savez_compressed(os.path.join(GOLDEN, tag + ".npz"), init=init, types=types, states=states,
                            rewards=rewards, done=done, game_over=game_over, other_agents_states=oas)
        print("wrote", tag)


def gen_maps(ns):
    out = {}
    for mid, obst in OBSTACLE_SETS.items():
        if not obst:
            continue
        with rh.quiet():
            m = ns.Map(30, 30, 0.1, obst)
        out["map_%d" % mid] = np.packbits(m.static_map.astype(np.uint8))
        out["rects_%d" % mid] = rects_of(obst)
    np.savez_compressed(os.path.join(GOLDEN, "maps.npz"), **out)
    print("wrote maps.npz")


def gen_occgrid_cv2(ns, n=400, seed=5):
    """Direct cv2 check of the OccupancyGridSensor restatement on many poses (incl. out-of-map ones)."""
    rng = np.random.default_rng(seed)
    with rh.quiet():
        maps = {mid: ns.Map(30, 30, 0.1, obst) for mid, obst in OBSTACLE_SETS.items() if obst}
    sensor = ns.OccupancyGridSensor()
    poses = np.zeros((n, 3))
    mids = np.zeros((n,), np.int32)
    grids = np.zeros((n, 60, 60), np.uint8)

    class _A:
        pass

    for k in range(n):
        mid = int(rng.choice([1, 2]))
        lim = 14.9 if rng.random() < 0.85 else 15.6
        x, y = f32(rng.uniform(-lim, lim, 2))
        h = float(f32(rng.uniform(-np.pi, np.pi))) if rng.random() < 0.9 else float(rng.choice([0.0, np.pi / 2, -np.pi / 2, np.pi]))
        a = _A()
        a.pos_global_frame = np.array([x, y])
        a.heading_global_frame = np.float64(h)
        with rh.quiet():
            g = sensor.sense([a], 0, maps[mid])
        poses[k] = [x, y, h]
        mids[k] = mid
        grids[k] = g.astype(np.uint8)
    np.savez_compressed(os.path.join(GOLDEN, "occgrid.npz"), poses=poses, map_id=mids,
                        grids=np.packbits(grids, axis=-1))
    print("wrote occgrid.npz")


FLAT_KEYS = ['dist_to_goal', 'rel_goal', 'radius', 'heading_ego_frame', 'heading_global_frame', 'pos_global_frame',
             'pref_speed', 'num_other_agents', 'other_agents_states', 'laserscan', 'local_grid', 'use_ppo']


def gen_flatten(ns, T=4):
    """Wire format: the reference's MultiagentFlattenDictWrapper.observation (envs/wrappers.py:37-45) applied to the
    observation dict of a 3-agent crossing-map environment with all three sensors, 10 agent slots."""
    from gym_collision_avoidance.envs.wrappers import MultiagentFlattenDictWrapper
    set_mode(ns, 2)
    mask = L.SENSOR_OTHER_AGENTS | L.SENSOR_LASERSCAN | L.SENSOR_OCCUPANCY_GRID
    init = f32(np.array([[-6.0, 0.3, 6.0, -0.5, 0.4, 1.0, 0.1], [-1.0, 5.5, 0.5, -7.0, 0.3, 1.2, -1.5],
                         [7.0, -0.6, -7.0, 0.4, 0.5, 0.8, 3.0]]))
    types = np.full((3,), L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, mask), np.int32)
    agents = build_agents(ns, init, types)
    env = rh.make_env(ns, agents, OBSTACLE_SETS[1], FLAT_KEYS)
    w = object.__new__(MultiagentFlattenDictWrapper)
    w.dict_keys, w.max_num_agents = FLAT_KEYS, ns.Config.MAX_NUM_AGENTS_IN_ENVIRONMENT
    states = np.zeros((T, 3, L.STATE_WORDS))
    flat = []
    for t in range(T):
        for i in range(3):
            states[t, i] = extract_state(agents[i], types[i])
        flat.append(np.asarray(w.observation(env.observation), np.float64))
        with rh.quiet():
            env.step({})
    np.savez_compressed(os.path.join(GOLDEN, "flatten.npz"), init=init, types=types, states=states, flat=np.stack(flat),
                        keys=np.array(FLAT_KEYS), map_id=1)
    print("wrote flatten.npz", np.stack(flat).shape)


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    os.makedirs(GOLDEN, exist_ok=True)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref"))   # real rvo2 binding if built
    ns = rh.modules()
    if what in ("c1", "all"):
        gen_c1(ns)
    if what in ("core", "all"):
        gen_single_step(ns, "core_single_step.npz", 300, 11, sensors_mode=False)
    if what in ("sensors", "all"):
        gen_single_step(ns, "sensors_single_step.npz", 120, 12, sensors_mode=True)
    if what in ("dynamics", "all"):
        # the two remaining dynamics models, in their own fixture so that the older ones keep
        # their random streams
        global DYN_CHOICES
        keep = DYN_CHOICES
        DYN_CHOICES = [L.DYN_UNICYCLE_MAX_ACC, L.DYN_UNICYCLE_SECOND_ORDER_EULER, L.DYN_UNICYCLE]
        gen_single_step(ns, "dynamics_extra_single_step.npz", 160, 13, sensors_mode=False)
        DYN_CHOICES = keep
    if what in ("learning", "all"):
        # LearningPolicy (id 3) + game_over mode "all learning agents done" (collision_avoidance_env.py:303-304,722-736),
        # again in a fixture of its own so that core_single_step.npz keeps its random stream
        global POLICY_CHOICES
        keepp = POLICY_CHOICES
        POLICY_CHOICES = [L.POLICY_LEARNING, L.POLICY_LEARNING, L.POLICY_LEARNING, L.POLICY_NONCOOPERATIVE, L.POLICY_STATIC,
                          L.POLICY_EXTERNAL, L.POLICY_CARRL]
        gen_single_step(ns, "learning_single_step.npz", 200, 14, sensors_mode=False)
        POLICY_CHOICES = keepp
    if what in ("reset", "all"):
        gen_reset(ns)
    if what in ("maps", "all"):
        gen_maps(ns)
        gen_occgrid_cv2(ns)
    if what in ("explore", "all"):
        from oracle import gen_golden_explore
        gen_golden_explore.main(ns)
    if what in ("rollouts", "all"):
        from oracle import gen_golden_explore
        gen_golden_explore.main_rollouts(ns)
    if what in ("mcts", "all"):
        from oracle import gen_golden_explore
        gen_golden_explore.main_mcts(ns)
    if what in ("scenarios", "all"):
        from oracle import gen_golden_scenarios
        gen_golden_scenarios.main(ns)
    if what in ("mcts_env", "all"):
        from oracle import gen_golden_explore
        gen_golden_explore.main_mcts_env(ns)
    if what in ("team", "all"):
        from oracle import gen_golden_explore
        gen_golden_explore.main_team(ns)
    if what in ("ga3c", "all"):
        from oracle import gen_golden_ga3c
        gen_golden_ga3c.main(ns)
        gen_golden_ga3c.gen_checkpoint_fixture()
    if what in ("flatten", "all"):
        gen_flatten(ns)
    if what in ("rvo", "all"):
        from oracle import gen_golden_rvo
        assert gen_golden_rvo.main() == 0
        gen_golden_rvo.main_policy(ns)
        gen_golden_rvo.main_policy_obstacles(ns)


if __name__ == "__main__":
    main()
