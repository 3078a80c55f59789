"""The reference's FIXED named scenarios (gym_collision_avoidance/envs/test_cases.py) on casim's descriptor classes, so that
scripts written against the fork -- ``agents, obstacles = tc.IG_agent_crossing()`` then ``env.set_agents(agents)`` /
``env.set_static_map(obstacles)`` (experiments/src/dmcts.py:60-70) -- keep working.  Host side, once per episode, not on
the step path.  The seeded TRAINING generators (train_agents_swap_circle / pairwise_swap / random_positions,
test_cases.py:1192-1463) are not here: they run on the device with the reference's own random streams
(``env.set_scenario_generator``, csrc/casim_scenario.cu).  Every other generator of that 3 000-line file is scenario
authoring outside the hot path and is not restated.

``Config.SCENARIOS_FOR_TRAINING`` of the fork is ``["IG_agent_crossing"]`` (envs/config.py:91): an environment that is
reset without ``set_agents`` / ``set_scenario_tensors`` / ``set_scenario_generator`` starts that scenario in every
environment of the batch, like the reference's ``_init_agents`` (collision_avoidance_env.py:403-445)."""
import numpy as np

from . import plugins as P

# the four 8 m blocks of the crossing map (test_cases.py:3219-3223, :3081-3085): corners 1 and 3 are what Map reads
CROSSING_OBSTACLES = [[(10, 10), (2, 10), (2, 2), (10, 2)], [(-2, 10), (-10, 10), (-10, 2), (-2, 2)],
                      [(10, -2), (2, -2), (2, -10), (10, -10)], [(-2, -2), (-10, -2), (-10, -10), (-2, -10)]]


def IG_agent_crossing(number_of_agents=1, ego_agent_policy=None, other_agents_policy=None, ego_agent_dynamics=None,
                      other_agents_dynamics=None, agents_sensors=(), seed=None, obstacle=None):    # noqa: N802
    """test_cases.py:3209-3237: three ig_mcts explorers on the horizontal arm of the crossing, two static targets in
    the vertical arm; the arguments are accepted and ignored, as in the reference."""
    sens = [P.OtherAgentsStatesSensor, P.LaserScanSensor]
    agents = [P.Agent(x, 0, 16, 0, 0.5, 1.0, 0, P.ig_mcts, P.FirstOrderDynamics, sens, i) for i, x in enumerate((-5, 0, 5))]
    agents.append(P.Agent(6, 12, 0, 0, 0.2, 1.0, 0, P.StaticPolicy, P.FirstOrderDynamics, [], 3))
    agents.append(P.Agent(-6, -12, 0, 0, 0.2, 1.0, 0, P.StaticPolicy, P.FirstOrderDynamics, [], 4))
    return agents, [list(o) for o in CROSSING_OBSTACLES]


def get_testcase_two_agents():
    """test_cases.py:79-86: a LearningPolicy agent and a GA3C-CADRL agent swapping corners"""
    g = 3
    return [P.Agent(-g, -g, g, g, 0.5, 1.0, 0.5, P.LearningPolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 0),
            P.Agent(g, g, -g, -g, 0.5, 1.0, 0.5, P.GA3CCADRLPolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 1)]


def get_testcase_two_agents_external_rvo():
    """test_cases.py:88-97: a CARRL agent facing an RVO agent on the x axis"""
    g = 3
    return [P.Agent(-g, 0, g, 0, 0.5, 1.0, 0.0, P.CARRLPolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 0),
            P.Agent(g, 0, -g, 0, 0.5, 1.0, np.pi, P.RVOPolicy, P.UnicycleDynamics, [P.OtherAgentsStatesSensor], 1)]


def get_testcase_two_agents_laserscanners():
    """test_cases.py:99-106.  The fork names a PPOPolicy class there that it does not define (the call raises NameError);
    LearningPolicy -- actions supplied by an external learner -- stands in."""
    g = 3
    return [P.Agent(-g, -g, g, g, 0.5, 1.0, 0.5, P.LearningPolicy, P.UnicycleDynamics, [P.LaserScanSensor], 0),
            P.Agent(g, g, -g, -g, 0.5, 1.0, 0.5, P.LearningPolicy, P.UnicycleDynamics, [P.LaserScanSensor], 1)]


FIXED_SCENARIOS = {"IG_agent_crossing": IG_agent_crossing}


def default_scenario(name):
    """-> (agents, obstacles) of a scenario named in Config.SCENARIOS_FOR_TRAINING, or None if it is not a fixed one"""
    fn = FIXED_SCENARIOS.get(name)
    return None if fn is None else fn()
