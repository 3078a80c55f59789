"""Configuration of the batched simulator.

``Config`` mirrors the names of the reference's global ``Config`` class
(gym_collision_avoidance/envs/config.py:3-221) for the attributes the stepping path reads;
scripts overwrite class attributes before creating an env exactly as with the reference
(experiments/src/run_cadrl_formations.py:15-24).  ``SimConfig`` is the frozen ctypes snapshot
(``casim_config`` in include/casim.h) handed to the CUDA library.
"""
import ctypes

import numpy as np

MAX_BEAMS = 32
MAX_RANGES = 32
MAX_AGENTS = 32
MAP_DIM = 300
NUM_STATS = 8


class Config:
    # envs/config.py:29-49
    DT = 0.1
    REWARD_AT_GOAL = 3.0
    REWARD_COLLISION_WITH_AGENT = -10.0
    REWARD_TIMEOUT = -10.0
    REWARD_INFEASIBLE = 0.0
    REWARD_COLLISION_WITH_WALL = -0.25
    REWARD_GETTING_CLOSE = 0.0
    REWARD_ENTERED_NORM_ZONE = 0.0
    REWARD_TIME_STEP = -0.01
    REWARD_DISTANCE_TO_GOAL = 0.0
    REWARD_WIGGLY_BEHAVIOR = 0.0
    WIGGLY_BEHAVIOR_THRESHOLD = 0.0
    COLLISION_DIST = 0.5
    GETTING_CLOSE_RANGE = 0.2
    COLLISION_AV_W_STATIC_AGENT = False
    # envs/config.py:15-16,46
    EVALUATE_MODE = False
    SCENARIOS_FOR_TRAINING = ["IG_agent_crossing"]   # config.py:91 of the fork (casim.test_cases holds the fixed ones)
    TRAIN_SINGLE_AGENT = True
    HOMOGENEOUS_TESTING = False
    # envs/config.py:55-72
    LASERSCAN_LENGTH = 16
    NEAR_GOAL_THRESHOLD = 0.75
    MAX_TIME_RATIO = 3.0
    SENSING_HORIZON = np.inf
    RVO_TIME_HORIZON = 5.0               # config.py:66
    MAX_NUM_AGENTS_IN_ENVIRONMENT = 10
    MAX_NUM_OTHER_AGENTS_IN_ENVIRONMENT = MAX_NUM_AGENTS_IN_ENVIRONMENT - 1
    MAX_NUM_OTHER_AGENTS_OBSERVED = MAX_NUM_AGENTS_IN_ENVIRONMENT - 1
    # envs/config.py:83-95
    SUBMAP_WIDTH = 60
    SUBMAP_HEIGHT = 60
    SUBMAP_RESOLUTION = 0.1
    MAP_WIDTH = 30
    MAP_HEIGHT = 30
    NUM_OF_SLICES = 16
    MAX_RANGE = 6
    # envs/config.py:98 (default); the batched env emits the entries whose sensors are enabled
    STATES_IN_OBS = ['radius', 'heading_global_frame', 'pos_global_frame', 'pref_speed',
                     'other_agents_states', 'local_grid']
    # collision_avoidance_env.py:97
    MAX_HEADING_CHANGE = 4.0
    # exploration (experiments/src/dmcts.py:73-78, policies/ig_greedy.py:18-29, targetMap.py:7-8)
    IG_DETECT_FOV = 60.0      # degrees
    IG_DETECT_RANGE = 5.0
    IG_MAP_RES = 0.1          # edfMap cell; the target map uses 5x this
    IG_R_OCC = 1.5
    IG_R_EMP = 0.66
    IG_TOLERANCE = 0.01
    IG_PRIOR = 1.0
    IG_P_FALSE_NEG = 0.1
    IG_P_FALSE_POS = 0.05
    # envs/config.py:7-8
    continuous, discrete = range(2)
    ACTION_SPACE_TYPE = continuous
    # envs/config.py:104-215: size / bounds / dtype of every observation entry (the 'attr' strings are replaced by the
    # SoA tensors of casim.env._obs_tensor); note other_agent_states is declared with size 9 but carries a 10-value row
    STATE_INFO_DICT = {
        'dist_to_goal': {'dtype': np.float32, 'size': 1, 'bounds': [-np.inf, np.inf]},
        'radius': {'dtype': np.float32, 'size': 1, 'bounds': [0, np.inf]},
        'rel_goal': {'dtype': np.float32, 'size': 2, 'bounds': [-np.inf, np.inf]},
        'heading_ego_frame': {'dtype': np.float32, 'size': 1, 'bounds': [-np.pi, np.pi]},
        'heading_global_frame': {'dtype': np.float32, 'size': 1, 'bounds': [-np.pi, np.pi]},
        'pos_global_frame': {'dtype': np.float32, 'size': 2, 'bounds': [-np.inf, np.inf]},
        'pref_speed': {'dtype': np.float32, 'size': 1, 'bounds': [0, np.inf]},
        'num_other_agents': {'dtype': np.float32, 'size': 1, 'bounds': [0, np.inf]},
        'other_agent_states': {'dtype': np.float32, 'size': 9, 'bounds': [-np.inf, np.inf]},
        'other_agents_states': {'dtype': np.float32, 'size': (MAX_NUM_OTHER_AGENTS_IN_ENVIRONMENT, 10),
                                'bounds': [-np.inf, np.inf]},
        'local_grid': {'dtype': np.float32, 'size': (SUBMAP_WIDTH, SUBMAP_HEIGHT), 'bounds': [-np.inf, np.inf]},
        'angular_map': {'dtype': np.float32, 'size': NUM_OF_SLICES, 'bounds': [0., 6.]},
        'laserscan': {'dtype': np.float32, 'size': LASERSCAN_LENGTH, 'bounds': [0., 6.]},
        'use_ppo': {'dtype': np.float32, 'size': 1, 'bounds': [0., 1.]},
    }
    # batched-env extension
    AUTO_RESET = False


class SimConfig(ctypes.Structure):
    """ctypes image of ``struct casim_config`` (include/casim.h)."""
    _fields_ = [
        ("dt", ctypes.c_double),
        ("reward_at_goal", ctypes.c_double),
        ("reward_collision_with_agent", ctypes.c_double),
        ("reward_timeout", ctypes.c_double),
        ("reward_infeasible", ctypes.c_double),
        ("reward_collision_with_wall", ctypes.c_double),
        ("reward_time_step", ctypes.c_double),
        ("reward_distance_to_goal", ctypes.c_double),
        ("reward_wiggly_behavior", ctypes.c_double),
        ("wiggly_behavior_threshold", ctypes.c_double),
        ("reward_min", ctypes.c_double),
        ("reward_max", ctypes.c_double),
        ("getting_close_range", ctypes.c_double),
        ("near_goal_threshold", ctypes.c_double),
        ("max_time_ratio", ctypes.c_double),
        ("sensing_horizon", ctypes.c_double),
        ("max_heading_change", ctypes.c_double),
        ("map_width", ctypes.c_double),
        ("map_height", ctypes.c_double),
        ("map_cell", ctypes.c_double),
        ("laser_max_range", ctypes.c_double),
        ("laser_norm_range", ctypes.c_double),
        ("laser_angles", ctypes.c_double * MAX_BEAMS),
        ("laser_ranges", ctypes.c_double * MAX_RANGES),
        ("ig_sens_fov", ctypes.c_double),
        ("ig_detect_fov", ctypes.c_double),
        ("ig_range", ctypes.c_double),
        ("ig_cell", ctypes.c_double),
        ("ig_rocc", ctypes.c_double),
        ("ig_remp", ctypes.c_double),
        ("ig_tolerance", ctypes.c_double),
        ("ig_p_false_neg", ctypes.c_double),
        ("ig_p_false_pos", ctypes.c_double),
        ("ig_prior", ctypes.c_double),
        ("ig_dt", ctypes.c_double),
        ("laser_num_beams", ctypes.c_int32),
        ("laser_num_ranges", ctypes.c_int32),
        ("collision_av_w_static_agent", ctypes.c_int32),
        ("max_other_agents_observed", ctypes.c_int32),
        ("max_other_agents_rows", ctypes.c_int32),
        ("submap_w", ctypes.c_int32),
        ("submap_h", ctypes.c_int32),
        ("game_over_mode", ctypes.c_int32),
        ("auto_reset", ctypes.c_int32),
        ("ig_map_dim", ctypes.c_int32),
        ("reserved", ctypes.c_int32 * 6),
    ]


GAME_OVER_AGENT0, GAME_OVER_ALL, GAME_OVER_ALL_LEARNING = 0, 1, 2


def make_sim_config(cfg=Config, **overrides):
    """Snapshot ``cfg`` (a Config-like class) into a SimConfig.

    Derived members follow the reference expressions literally:
    laser_angles = np.linspace(-pi, pi, LASERSCAN_LENGTH), laser_ranges = np.arange(0, 6, 2*pi/NUM_OF_SLICES)
    (sensors/LaserScanSensor.py:13-20); reward_min/max = min/max over the five reward values
    (collision_avoidance_env.py:766-774); game-over rule (collision_avoidance_env.py:722-736).
    """
    c = SimConfig()
    c.dt = cfg.DT
    c.reward_at_goal = cfg.REWARD_AT_GOAL
    c.reward_collision_with_agent = cfg.REWARD_COLLISION_WITH_AGENT
    c.reward_timeout = cfg.REWARD_TIMEOUT
    c.reward_infeasible = cfg.REWARD_INFEASIBLE
    c.reward_collision_with_wall = cfg.REWARD_COLLISION_WITH_WALL
    c.reward_time_step = cfg.REWARD_TIME_STEP
    c.reward_distance_to_goal = cfg.REWARD_DISTANCE_TO_GOAL
    c.reward_wiggly_behavior = cfg.REWARD_WIGGLY_BEHAVIOR
    c.wiggly_behavior_threshold = cfg.WIGGLY_BEHAVIOR_THRESHOLD
    vals = np.array([cfg.REWARD_AT_GOAL, cfg.REWARD_COLLISION_WITH_AGENT, cfg.REWARD_TIME_STEP,
                     cfg.REWARD_COLLISION_WITH_WALL, cfg.REWARD_WIGGLY_BEHAVIOR])
    c.reward_min = float(np.min(vals))
    c.reward_max = float(np.max(vals))
    c.getting_close_range = cfg.GETTING_CLOSE_RANGE
    c.near_goal_threshold = cfg.NEAR_GOAL_THRESHOLD
    c.max_time_ratio = cfg.MAX_TIME_RATIO
    c.sensing_horizon = float(cfg.SENSING_HORIZON)
    c.max_heading_change = cfg.MAX_HEADING_CHANGE
    c.map_width = cfg.MAP_WIDTH
    c.map_height = cfg.MAP_HEIGHT
    c.map_cell = cfg.SUBMAP_RESOLUTION
    if int(cfg.MAP_WIDTH / cfg.SUBMAP_RESOLUTION) != MAP_DIM or int(cfg.MAP_HEIGHT / cfg.SUBMAP_RESOLUTION) != MAP_DIM:
        raise ValueError("the static map must be %dx%d cells (Map(30, 30, 0.1))" % (MAP_DIM, MAP_DIM))
    c.laser_max_range = 6.0
    c.laser_norm_range = cfg.MAX_RANGE
    angles = np.linspace(-np.pi, np.pi, cfg.LASERSCAN_LENGTH)
    ranges = np.arange(0, 6, 2 * np.pi / cfg.NUM_OF_SLICES)
    if len(angles) > MAX_BEAMS or len(ranges) > MAX_RANGES:
        raise ValueError("at most %d beams x %d range samples" % (MAX_BEAMS, MAX_RANGES))
    for i, a in enumerate(angles):
        c.laser_angles[i] = float(a)
    for i, r in enumerate(ranges):
        c.laser_ranges[i] = float(r)
    c.laser_num_beams = len(angles)
    c.laser_num_ranges = len(ranges)
    c.ig_sens_fov = cfg.IG_DETECT_FOV * np.pi / 180
    c.ig_detect_fov = cfg.IG_DETECT_FOV
    c.ig_range = cfg.IG_DETECT_RANGE
    c.ig_cell = cfg.IG_MAP_RES * 5
    c.ig_rocc = cfg.IG_R_OCC
    c.ig_remp = cfg.IG_R_EMP
    c.ig_tolerance = cfg.IG_TOLERANCE
    c.ig_p_false_neg = cfg.IG_P_FALSE_NEG
    c.ig_p_false_pos = cfg.IG_P_FALSE_POS
    c.ig_prior = cfg.IG_PRIOR
    c.ig_dt = cfg.DT
    c.ig_map_dim = int(cfg.MAP_WIDTH / (cfg.IG_MAP_RES * 5))
    c.collision_av_w_static_agent = int(bool(cfg.COLLISION_AV_W_STATIC_AGENT))
    c.max_other_agents_observed = cfg.MAX_NUM_OTHER_AGENTS_OBSERVED
    c.max_other_agents_rows = cfg.MAX_NUM_OTHER_AGENTS_IN_ENVIRONMENT
    c.submap_w = cfg.SUBMAP_WIDTH
    c.submap_h = cfg.SUBMAP_HEIGHT
    if cfg.EVALUATE_MODE:
        c.game_over_mode = GAME_OVER_ALL if cfg.HOMOGENEOUS_TESTING else GAME_OVER_AGENT0
    elif cfg.TRAIN_SINGLE_AGENT:
        c.game_over_mode = GAME_OVER_AGENT0
    else:
        c.game_over_mode = GAME_OVER_ALL_LEARNING
    c.auto_reset = int(bool(getattr(cfg, "AUTO_RESET", False)))
    for k, v in overrides.items():
        setattr(c, k, v)
    return c
