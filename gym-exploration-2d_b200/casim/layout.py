"""State-record layout shared by the Python host code and the CUDA library (include/casim.h).

24 four-byte words per agent, stored on the GPU as 6 planes of float4: ``state[6, E, A, 4]``.
"""
NUM_PACKS = 6
STATE_WORDS = 24
(PX, PY, HEADING, SPEED,
 VX, VY, DELTA_HEADING, TURNING_DIR,
 GX, GY, RADIUS, PREF_SPEED,
 T, TIME_REMAINING, DIST_TO_GOAL, PAST_DIST_TO_GOAL,
 PAST_ACT0_V, PAST_ACT0_W, PAST_ACT1_V, PAST_ACT1_W,
 HEADING_EGO, STEP_NUM, FLAGS, TYPE) = range(STATE_WORDS)
INT_WORDS = (STEP_NUM, FLAGS, TYPE)

F_IS_AT_GOAL = 1
F_WAS_AT_GOAL_ALREADY = 2
F_IN_COLLISION = 4
F_WAS_IN_COLLISION_ALREADY = 8
F_RAN_OUT_OF_TIME = 16
F_IS_DONE = 32
F_IS_INFEASIBLE = 64

POLICY_STATIC, POLICY_NONCOOPERATIVE, POLICY_EXTERNAL, POLICY_LEARNING, POLICY_CARRL, \
    POLICY_IG_GREEDY, POLICY_GA3C_CADRL, POLICY_PEDESTRIAN, POLICY_RVO = range(9)
DYN_UNICYCLE, DYN_FIRST_ORDER, DYN_EXTERNAL, DYN_UNICYCLE_MAX_TURN_RATE, DYN_UNICYCLE_MAX_ACC, \
    DYN_UNICYCLE_SECOND_ORDER_EULER = range(6)
SENSOR_OTHER_AGENTS, SENSOR_LASERSCAN, SENSOR_OCCUPANCY_GRID = 1, 2, 4


def encode_type(policy_id, dynamics_id, sensor_mask):
    return int(policy_id) | (int(dynamics_id) << 8) | (int(sensor_mask) << 16)


def decode_type(t):
    t = int(t)
    return t & 0xff, (t >> 8) & 0xff, (t >> 16) & 0xff


# enum casim_obs_key (include/casim.h): names of Config.STATE_INFO_DICT, envs/config.py:104-215
OBS_KEYS = ("dist_to_goal", "rel_goal", "radius", "heading_ego_frame", "heading_global_frame", "pos_global_frame",
            "pref_speed", "num_other_agents", "other_agent_states", "other_agents_states", "local_grid", "laserscan",
            "use_ppo")
OBS_KEY_ID = {k: i for i, k in enumerate(OBS_KEYS)}

# enum casim_scenario (include/casim.h): envs/test_cases.py:1192-1281 / :1365-1463
SCENARIO_SWAP_CIRCLE, SCENARIO_RANDOM_POSITIONS, SCENARIO_PAIRWISE_SWAP = 0, 1, 2
