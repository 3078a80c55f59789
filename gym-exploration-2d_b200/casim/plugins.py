"""Plugin vocabulary of the reference, as descriptors.

The reference passes policy / dynamics / sensor CLASSES to ``Agent`` (envs/agent.py:9-18) and instantiates them per
agent; their Python methods are the per-agent hot loop.  Here the same names exist so that scenario code reads
the same, but a class is only a descriptor: it carries the integer id that selects the in-kernel implementation
(include/casim.h enums).  There is no Python/CPU implementation behind them.
"""
from . import layout as L


class Policy:
    """envs/policies/Policy.py:4-12"""
    policy_id = None
    str = "NoPolicy"
    is_still_learning = False
    is_external = False
    needs_actions = False       # the caller supplies this agent's action in step(actions)

    def __init__(self, str=None):
        if str is not None:
            self.str = str

    def find_next_action(self, obs, agents, i, obstacles=None):
        raise RuntimeError("%s is evaluated inside the CUDA step kernel (libcasim.so); there is no Python path"
                           % type(self).__name__)


class StaticPolicy(Policy):
    """envs/policies/StaticPolicy.py:9-12 -> action [0, 0]"""
    policy_id, str = L.POLICY_STATIC, "Static"


class NonCooperativePolicy(Policy):
    """envs/policies/NonCooperativePolicy.py:10-13 -> [pref_speed, -heading_ego_frame]"""
    policy_id, str = L.POLICY_NONCOOPERATIVE, "NonCooperativePolicy"


class ExternalPolicy(Policy):
    """envs/policies/ExternalPolicy.py: the caller's action [speed, delta_heading] is used as is"""
    policy_id, str, is_external, needs_actions = L.POLICY_EXTERNAL, "External", True, True


class LearningPolicy(Policy):
    """envs/policies/LearningPolicy.py:11-16: action in [0,1]^2 -> [pref_speed*u0, max_heading_change*(2u1-1)]"""
    policy_id, str, is_still_learning, needs_actions = L.POLICY_LEARNING, "learning", True, True


class CARRLPolicy(Policy):
    """envs/policies/CARRLPolicy.py:14-15: discrete action index -> 11-entry table"""
    policy_id, str, is_external, needs_actions = L.POLICY_CARRL, "CARRL", True, True


class ig_greedy(Policy):  # noqa: N801  (reference class name)
    """envs/policies/ig_greedy.py:31-93: belief update + greedy information-gain action (casim_explore_step)"""
    policy_id, str = L.POLICY_IG_GREEDY, "ig_greedy"


class ig_mcts(Policy):  # noqa: N801  (reference class name)
    """envs/policies/ig_mcts.py: team belief update (casim_explore_team_update) + Dec-MCTS planning (casim/mcts.py over
    casim_explore_rollouts).  An explorer like ig_greedy as far as the kernels are concerned (same policy id); the
    planner parameters are those of ``set_param`` (ig_mcts.py:54-77) with the reference's defaults (:45-52)."""
    policy_id, str = L.POLICY_IG_GREEDY, "ig_mcts"

    def __init__(self, Ntree=100, Nsims=10, mcts_cp=1.0, mcts_horizon=10, mcts_gamma=1.0, Ncycles=5, xdt=1, seed=0):
        self.Ntree, self.Nsims, self.mcts_cp, self.mcts_horizon = int(Ntree), int(Nsims), float(mcts_cp), int(mcts_horizon)
        self.mcts_gamma, self.Ncycles, self.xDT, self.seed = float(mcts_gamma), int(Ncycles), int(xdt), int(seed)
        self.reference_order = getattr(self, "reference_order", False)   # tests: draw from np.random like the reference

    def set_param(self, ego_agent=None, occ_map=None, map_size=None, map_res=None, detect_fov=None, detect_range=None,
                  dt=0.1, xdt=1, Ntree=100, Nsims=10, parallelize_sims=False, mcts_cp=1., mcts_horizon=10, mcts_gamma=1,
                  Ncycles=5, parallelize_agents=False):
        """ig_mcts.set_param (ig_mcts.py:54-77): the planner parameters; the map / sensor arguments are the batch's
        (Config + set_static_map) and are accepted for call compatibility only."""
        self.__init__(Ntree=Ntree, Nsims=Nsims, mcts_cp=mcts_cp, mcts_horizon=mcts_horizon, mcts_gamma=mcts_gamma,
                      Ncycles=Ncycles, xdt=xdt, seed=self.seed)


class GA3CCADRLPolicy(Policy):
    """envs/policies/GA3CCADRLPolicy.py:34-43 (casim_ga3c_forward)"""
    policy_id, str = L.POLICY_GA3C_CADRL, "GA3CCADRLPolicy"


class RVOPolicy(Policy):
    """envs/policies/RVOPolicy.py:53-117 + Python-RVO2 (casim_rvo_step); uses Agent.cooperation_coef (:85)"""
    policy_id, str = L.POLICY_RVO, "RVO"


class PedestrianDatasetPolicy(Policy):
    """envs/policies/PedestrianDatasetPolicy.py:15-16 -> zeros (motion comes from ExternalDynamics)"""
    policy_id, str = L.POLICY_PEDESTRIAN, "PedestrianDatasetPolicy"


class Dynamics:
    """envs/dynamics/Dynamics.py:6-14"""
    dynamics_id = None

    def __init__(self, agent=None):
        self.agent = agent


class UnicycleDynamics(Dynamics):
    """envs/dynamics/UnicycleDynamics.py:10-31"""
    dynamics_id = L.DYN_UNICYCLE


class FirstOrderDynamics(Dynamics):
    """envs/dynamics/FirstOrderDynamics.py:10-23"""
    dynamics_id = L.DYN_FIRST_ORDER


class ExternalDynamics(Dynamics):
    """envs/dynamics/ExternalDynamics.py:8-23 (trajectory playback; tables via set_trajectories)"""
    dynamics_id = L.DYN_EXTERNAL


class UnicycleDynamicsMaxAcc(Dynamics):
    """envs/dynamics/UnicycleDynamicsMaxAcc.py:17-41"""
    dynamics_id = L.DYN_UNICYCLE_MAX_ACC


class UnicycleSecondOrderEulerDynamics(Dynamics):
    """envs/dynamics/UnicycleSecondOrderEulerDynamics.py:12-31"""
    dynamics_id = L.DYN_UNICYCLE_SECOND_ORDER_EULER


class UnicycleDynamicsMaxTurnRate(Dynamics):
    """envs/dynamics/UnicycleDynamicsMaxTurnRate.py:11-26"""
    dynamics_id = L.DYN_UNICYCLE_MAX_TURN_RATE


class Sensor:
    """envs/sensors/Sensor.py:1-6"""
    name = "empty_obs"
    bit = 0

    def sense(self, agents, agent_index, top_down_map):
        raise RuntimeError("%s runs inside libcasim.so; there is no Python path" % type(self).__name__)


class OtherAgentsStatesSensor(Sensor):
    """envs/sensors/OtherAgentsStatesSensor.py:11-77"""
    name, bit = "other_agents_states", L.SENSOR_OTHER_AGENTS


class LaserScanSensor(Sensor):
    """envs/sensors/LaserScanSensor.py:27-58"""
    name, bit = "laserscan", L.SENSOR_LASERSCAN


class OccupancyGridSensor(Sensor):
    """envs/sensors/OccupancyGridSensor.py:70-98"""
    name, bit = "local_grid", L.SENSOR_OCCUPANCY_GRID


def _cls(x):
    return x if isinstance(x, type) else type(x)


class Agent:
    """envs/agent.py:9-112: same constructor; policy / dynamics_model / sensors are passed as classes.  After the
    agent is handed to an env, the kinematic attributes read through to the device state of its (env, slot)."""

    def __init__(self, start_x, start_y, goal_x, goal_y, radius, pref_speed, initial_heading, policy, dynamics_model,
                 sensors, id, cooperation_coef=1.0):
        self.start = (float(start_x), float(start_y))
        self.goal = (float(goal_x), float(goal_y))
        self.radius = float(radius)
        self.pref_speed = float(pref_speed)
        self.initial_heading = None if initial_heading is None else float(initial_heading)
        self.policy = _cls(policy)()
        self.dynamics_model = _cls(dynamics_model)(self)
        self.sensors = [_cls(s)() for s in sensors]
        self.id = id
        self.cooperation_coef = cooperation_coef
        self._env, self._e, self._i = None, 0, 0

    @property
    def type_word(self):
        mask = 0
        for s in self.sensors:
            mask |= s.bit
        return L.encode_type(self.policy.policy_id, self.dynamics_model.dynamics_id, mask)

    def init_row(self):
        h = float("nan") if self.initial_heading is None else self.initial_heading
        return [self.start[0], self.start[1], self.goal[0], self.goal[1], self.radius, self.pref_speed, h]

    # ---- read-through views of the device state (host copy on access; for drop-in use at small E) ----
    def _word(self, k):
        if self._env is None:
            raise RuntimeError("agent is not bound to an environment yet")
        return self._env._state_word(self._e, self._i, k)

    pos_global_frame = property(lambda self: (self._word(L.PX), self._word(L.PY)))
    vel_global_frame = property(lambda self: (self._word(L.VX), self._word(L.VY)))
    goal_global_frame = property(lambda self: (self._word(L.GX), self._word(L.GY)))
    heading_global_frame = property(lambda self: self._word(L.HEADING))
    speed_global_frame = property(lambda self: self._word(L.SPEED))
    heading_ego_frame = property(lambda self: self._word(L.HEADING_EGO))
    dist_to_goal = property(lambda self: self._word(L.DIST_TO_GOAL))
    t = property(lambda self: self._word(L.T))
    step_num = property(lambda self: int(self._word(L.STEP_NUM)))
    is_at_goal = property(lambda self: bool(int(self._word(L.FLAGS)) & L.F_IS_AT_GOAL))
    in_collision = property(lambda self: bool(int(self._word(L.FLAGS)) & L.F_IN_COLLISION))
    ran_out_of_time = property(lambda self: bool(int(self._word(L.FLAGS)) & L.F_RAN_OUT_OF_TIME))
    is_done = property(lambda self: bool(int(self._word(L.FLAGS)) & L.F_IS_DONE))
