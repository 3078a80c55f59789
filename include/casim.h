/*
 * casim -- B200-native batched simulator for the CollisionAvoidanceEnv stepping path
 * of mlodel/gym-exploration-2d.  C ABI of libcasim.so.
 *
 * This is the drop-in boundary: every entry point below is
 * what a binding of the reference's stepping path would call instead of the Python
 * implementation cited next to it (paths relative to the reference tree,
 * gym_collision_avoidance/envs/...).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no torch types.
 *   - Unless a function name ends in _host, every pointer is a DEVICE pointer owned by the
 *     caller; the library never frees caller memory.  Calls are asynchronous on the passed
 *     stream (a cudaStream_t passed as void*; NULL = legacy default stream).
 *   - Return value: 0 = ok, otherwise a cudaError_t-compatible code (or CASIM_EINVAL);
 *     casim_last_error() returns a human-readable message for the calling thread.
 *   - One handle per GPU; calls on one stream are ordered; no host threads inside.
 */
#ifndef CASIM_H
#define CASIM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CASIM_ABI_VERSION 1
#define CASIM_EINVAL 10001

/* ------------------------------------------------------------------------------------ */
/* State record: 24 four-byte words per agent, stored as 6 planes of float4 ("packs"),   */
/* plane p = state + p*E*A float4's, element (e*A + a).  Mirrors the attributes of       */
/* envs/agent.py:9-112 that the stepping path reads or writes.                           */
/* ------------------------------------------------------------------------------------ */
#define CASIM_NUM_PACKS 6
#define CASIM_STATE_WORDS 24
enum casim_state_word {
  /* pack 0 */ CASIM_PX = 0, CASIM_PY, CASIM_HEADING, CASIM_SPEED,
  /* pack 1 */ CASIM_VX, CASIM_VY, CASIM_DELTA_HEADING, CASIM_TURNING_DIR,
  /* pack 2 */ CASIM_GX, CASIM_GY, CASIM_RADIUS, CASIM_PREF_SPEED,           /* constant in an episode */
  /* pack 3 */ CASIM_T, CASIM_TIME_REMAINING, CASIM_DIST_TO_GOAL, CASIM_PAST_DIST_TO_GOAL,
  /* pack 4 */ CASIM_PAST_ACT0_V, CASIM_PAST_ACT0_W, CASIM_PAST_ACT1_V, CASIM_PAST_ACT1_W,
  /* pack 5 */ CASIM_HEADING_EGO, CASIM_STEP_NUM /*i32*/, CASIM_FLAGS /*i32*/, CASIM_TYPE /*i32*/
};

/* CASIM_FLAGS bits (envs/agent.py:67-72, collision_avoidance_env.py:720) */
enum casim_flag {
  CASIM_F_IS_AT_GOAL = 1, CASIM_F_WAS_AT_GOAL_ALREADY = 2, CASIM_F_IN_COLLISION = 4,
  CASIM_F_WAS_IN_COLLISION_ALREADY = 8, CASIM_F_RAN_OUT_OF_TIME = 16, CASIM_F_IS_DONE = 32,
  CASIM_F_IS_INFEASIBLE = 64
};

/* CASIM_TYPE = policy | dynamics<<8 | sensor_mask<<16 */
enum casim_policy {            /* envs/policies/ */
  CASIM_POLICY_STATIC = 0,         /* StaticPolicy.py:9-12            */
  CASIM_POLICY_NONCOOPERATIVE = 1, /* NonCooperativePolicy.py:10-13   */
  CASIM_POLICY_EXTERNAL = 2,       /* ExternalPolicy.py: action = caller's [speed, dheading] */
  CASIM_POLICY_LEARNING = 3,       /* LearningPolicy.py:11-16         */
  CASIM_POLICY_CARRL = 4,          /* CARRLPolicy.py:14-15 (11-entry table) */
  CASIM_POLICY_IG_GREEDY = 5,      /* ig_greedy.py:31-43 (action from casim_explore_step) */
  CASIM_POLICY_GA3C_CADRL = 6,     /* GA3CCADRLPolicy.py:34-43 (action from casim_ga3c_forward) */
  CASIM_POLICY_PEDESTRIAN = 7,     /* PedestrianDatasetPolicy.py:15-16 (zeros) */
  CASIM_POLICY_RVO = 8             /* RVOPolicy.py:53-117 + Python-RVO2/src/Agent.cpp (action from casim_rvo_step) */
};
enum casim_dynamics {          /* envs/dynamics/ */
  CASIM_DYN_UNICYCLE = 0,          /* UnicycleDynamics.py:10-31       */
  CASIM_DYN_FIRST_ORDER = 1,       /* FirstOrderDynamics.py:10-23     */
  CASIM_DYN_EXTERNAL = 2,          /* ExternalDynamics.py:8-23 + agent.py:125-145 */
  CASIM_DYN_UNICYCLE_MAX_TURN_RATE = 3, /* UnicycleDynamicsMaxTurnRate.py:11-26 */
  CASIM_DYN_UNICYCLE_MAX_ACC = 4,       /* UnicycleDynamicsMaxAcc.py:17-41: current_speed lives in the SPEED word,
                                           current_turning_rate in the TURNING_DIR word */
  CASIM_DYN_UNICYCLE_SECOND_ORDER_EULER = 5 /* UnicycleSecondOrderEulerDynamics.py:12-31:
                                           Agent.angular_speed_global_frame (agent.py:27) lives in the TURNING_DIR word */
};
enum casim_sensor {            /* envs/sensors/ */
  CASIM_SENSOR_OTHER_AGENTS = 1,   /* OtherAgentsStatesSensor.py:11-77 */
  CASIM_SENSOR_LASERSCAN = 2,      /* LaserScanSensor.py:27-58         */
  CASIM_SENSOR_OCCUPANCY_GRID = 4  /* OccupancyGridSensor.py:70-98     */
};

enum casim_game_over_mode {    /* collision_avoidance_env.py:722-736 */
  CASIM_GAME_OVER_AGENT0 = 0,      /* TRAIN_SINGLE_AGENT, or EVALUATE_MODE and not HOMOGENEOUS_TESTING */
  CASIM_GAME_OVER_ALL = 1,         /* EVALUATE_MODE and HOMOGENEOUS_TESTING */
  CASIM_GAME_OVER_ALL_LEARNING = 2 /* otherwise: all agents whose policy is_still_learning */
};

#define CASIM_MAX_BEAMS 32
#define CASIM_MAX_RANGES 32
#define CASIM_MAX_AGENTS 32      /* one warp lane per agent of an environment */
#define CASIM_MAP_DIM 300        /* Map(30 m, 30 m, 0.1 m): collision_avoidance_env.py:497 */
#define CASIM_MAP_WORDS 10       /* 32-bit words per bit-packed map row (ceil(300/32)) */
#define CASIM_NUM_STATS 8
#define CASIM_GA3C_NUM_ACTIONS 11
#define CASIM_GA3C_INPUT 75      /* GA3C_CADRL/network.py:148-151 NN_INPUT_SIZE */
#define CASIM_GA3C_NUM_PARAMS 170764

/* Snapshot of the reference's global Config (envs/config.py:3-221) plus the env members the
 * path reads (collision_avoidance_env.py:97, 755-774).  All thresholds are doubles because the
 * kernels evaluate the decision predicates in fp64, like the reference. */
typedef struct casim_config {
  double dt;                          /* Config.DT                      config.py:29 */
  double reward_at_goal;              /* config.py:30 */
  double reward_collision_with_agent; /* config.py:31 */
  double reward_timeout;              /* config.py:32 */
  double reward_infeasible;           /* config.py:33 */
  double reward_collision_with_wall;  /* config.py:34 */
  double reward_time_step;            /* config.py:37 */
  double reward_distance_to_goal;     /* config.py:38 */
  double reward_wiggly_behavior;      /* config.py:39 */
  double wiggly_behavior_threshold;   /* config.py:40 */
  double reward_min;                  /* collision_avoidance_env.py:773 */
  double reward_max;                  /* collision_avoidance_env.py:774 */
  double getting_close_range;         /* config.py:43 */
  double near_goal_threshold;         /* config.py:60 */
  double max_time_ratio;              /* config.py:61 */
  double sensing_horizon;             /* config.py:63 (inf) */
  double max_heading_change;          /* collision_avoidance_env.py:97 (LearningPolicy) */
  double map_width, map_height;       /* config.py:88-89 (metres) */
  double map_cell;                    /* config.py:85 SUBMAP_RESOLUTION */
  double laser_max_range;             /* LaserScanSensor.py:15 */
  double laser_norm_range;            /* config.py:95 MAX_RANGE */
  double laser_angles[CASIM_MAX_BEAMS];  /* np.linspace(-pi, pi, num_beams)      LaserScanSensor.py:19 */
  double laser_ranges[CASIM_MAX_RANGES]; /* np.arange(0, max_range, 2*pi/slices)  LaserScanSensor.py:20 */
  /* exploration belief (ig_greedy.py:18-29, targetMap.py:7-24, experiments/src/dmcts.py:73-78) */
  double ig_sens_fov;                 /* radians (targetMap.sensFOV) */
  double ig_detect_fov;               /* the number ig_greedy compares against an angle in radians (degrees, sic) */
  double ig_range;
  double ig_cell;                     /* target-map cell size (map_res*5 = 0.5) */
  double ig_rocc, ig_remp, ig_tolerance, ig_p_false_neg, ig_p_false_pos, ig_prior;
  double ig_dt;
  int32_t laser_num_beams;            /* config.py:55 */
  int32_t laser_num_ranges;
  int32_t collision_av_w_static_agent;/* config.py:49 */
  int32_t max_other_agents_observed;  /* config.py:72 (sort clip)      */
  int32_t max_other_agents_rows;      /* config.py:71 (rows of other_agents_states) */
  int32_t submap_w, submap_h;         /* config.py:83-84 */
  int32_t game_over_mode;             /* enum casim_game_over_mode */
  int32_t auto_reset;                 /* batched-env extension: a game_over env restarts from its
                                         initial condition at the beginning of the next step */
  int32_t ig_map_dim;                 /* cells per side of the target map (60) */
  int32_t reserved[6];
} casim_config;

/* Observation / result buffers written by casim_reset and casim_step.  NULL = not wanted
 * (a sensor buffer must be non-NULL if any agent carries that sensor). */
typedef struct casim_obs {
  float*   obs_misc;            /* [E,A,4]  rel_goal_x, rel_goal_y, num_other_agents, use_ppo
                                   (dynamics/Dynamics.py:28, OtherAgentsStatesSensor.py:75, config.py:208-213) */
  float*   other_agents_states; /* [E,A,rows,10]  OtherAgentsStatesSensor.py:35-77 */
  float*   laserscan;           /* [E,A,num_beams] LaserScanSensor.py:58 */
  uint8_t* local_grid;          /* [E,A,submap_w,submap_h] bool (rows = map rows)  OccupancyGridSensor.py:98 */
  float*   rewards;             /* [E,A]  collision_avoidance_env.py:502-567 (per-agent; agent 0's is the
                                   TRAIN_SINGLE_AGENT scalar) */
  uint8_t* done;                /* [E,A]  which_agents_done, :711-720 */
  uint8_t* game_over;           /* [E]    :722-736 */
  uint8_t* collision_flags;     /* [E,A]  bit0 collision_with_agent, bit1 collision_with_wall (:630-671); may be NULL */
  float*   nearest_dist;        /* [E,A]  dist_btwn_nearest_agent (:649), +inf if none; may be NULL */
  float*   stats;               /* [CASIM_NUM_STATS] per-step sums over the shard: n_new_collisions, n_new_at_goal,
                                   n_new_timeouts, n_game_over_envs, reward_sum, n_active_agents, n_agents, 0; may be NULL */
} casim_obs;

#define CASIM_ZEROCOPY_MAX (1u << 20)

/* Parameters of the RVO policy's simulator (RVOPolicy.py:13-28): what RVOPolicy.__init__ reads from Config. */
typedef struct casim_rvo_params {
  double neighbor_dist;      /* Config.SENSING_HORIZON (inf)                config.py:63, RVOPolicy.py:14 */
  double time_horizon;       /* Config.RVO_TIME_HORIZON (5.0)               config.py:66, RVOPolicy.py:23 */
  double radius_scale;       /* 1 + 15e-2                                   RVOPolicy.py:78 */
  double max_delta_heading;  /* pi/6                                        RVOPolicy.py:20 */
  int32_t max_neighbors;     /* Config.MAX_NUM_AGENTS_IN_ENVIRONMENT (10)   config.py:70, RVOPolicy.py:15 */
  int32_t reserved[3];
} casim_rvo_params;

typedef struct casim_handle casim_handle;

const char* casim_last_error(void);
int casim_abi_version(void);

/* Fill a config with the reference defaults (envs/config.py, LaserScanSensor.py:13-20,
 * experiments/src/dmcts.py:73-78).  Host-only, no GPU needed. */
void casim_default_config(casim_config* cfg);

/* Replaces CollisionAvoidanceEnv.__init__ (collision_avoidance_env.py:60-160) for E environments of
 * at most A agents.  Allocates only internal scratch (bit-packed maps, per-step partial sums). */
int casim_create(const casim_config* cfg, int E, int A, casim_handle** out);
int casim_destroy(casim_handle* h);
int casim_set_config(casim_handle* h, const casim_config* cfg);

/* Replaces _init_static_map + Map.__init__/get_occupancy_grid (collision_avoidance_env.py:481-500,
 * Map.py:10-36,107-123).  rects: [n_maps, max_rects, 4] = (x of corner 1, y of corner 1, x of corner 3,
 * y of corner 3) in metres; n_rects: [n_maps].  n_maps is 1 (shared by all envs) or E.  n_maps = 0 clears
 * the map ("no obstacles": the wall check of collision_avoidance_env.py:656 is skipped).
 * Also (re)builds the squared Euclidean distance field used by the exploration kernels
 * (information_models/edfMap.py:11-12).  map_out (optional): [n_maps,300,300] uint8 copy of static_map. */
int casim_set_obstacles(casim_handle* h, const float* rects, const int32_t* n_rects, int n_maps,
                        int max_rects, uint8_t* map_out, void* stream);
/* Same, from a ready-made occupancy raster [n_maps,300,300] uint8 (Map.static_map). */
int casim_set_map(casim_handle* h, const uint8_t* static_map, int n_maps, void* stream);

/* Allocation-free map build.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same
 * kernels as casim_set_obstacles, but with every intermediate (bit planes, column distances, squared distances) in one
 * file-scope `static __device__` array of the library (3 072 000 B, held by every GPU that loads the library until the
 * process exits) instead of the handle's buffers; the call itself allocates nothing and leaves the handle's own map
 * untouched (the handle supplies the Map geometry of its config only).  All pointers are DEVICE pointers.
 *   rects [n_maps,max_rects,4], n_rects [n_maps]: as casim_set_obstacles;
 *   map_out  (optional) [n_maps,300,300] uint8  = what casim_set_obstacles writes to its map_out, bit for bit;
 *   edf_out  (optional) [n_maps,300,300] double = edfMap.map, sqrt(squared distance) * cell (information_models/edfMap.py:11-12).
 * Limits: 1 <= n_maps <= CASIM_NOALLOC_MAX_MAPS and max_rects <= CASIM_NOALLOC_MAX_RECTS, anything beyond is refused with
 * CASIM_EINVAL and a casim_last_error() message.  Every call shares the one array and overwrites the previous call's
 * contents (every word read is written earlier in the same call, nothing needs clearing); calls that overlap on the GPU
 * -- different streams, or different handles on one device -- race on it: serialise them. */
#define CASIM_NOALLOC_MAX_MAPS 4
#define CASIM_NOALLOC_MAX_RECTS 256
int casim_rasterize_noalloc(casim_handle* h, const float* rects, const int32_t* n_rects, int n_maps, int max_rects,
                            uint8_t* map_out, double* edf_out, void* stream);

/* Multi-GPU statistics hook (the per-step all-reduce of the [8] episode statistics, the design's only
 * collective; the reference's counterpart is the rolling n_collisions / n_timeouts of collision_avoidance_env.py:147-148,
 * 218-221).  When an event (cudaEvent_t passed as void*) is set, every casim_step records it on its stream right after
 * the fused step kernel -- rewards, done, game_over and stats are final at that point, the sensor kernels that follow do
 * not touch them -- so that a side stream can wait on it and start the collective while the sensors still run.
 * NULL turns it off.  Works under stream capture. */
int casim_set_stats_event(casim_handle* h, void* cuda_event);

/* Optional second delivery format of the OccupancyGrid observation (OccupancyGridSensor.py:98 returns a bool array):
 * when `packed` is non-NULL every later casim_step / casim_reset / casim_observe ALSO writes each agent's grid bit-packed,
 * ceil(submap_w*submap_h/8) bytes per agent = numpy.packbits(local_grid[e,a].ravel()) (first cell in the most significant
 * bit), into packed [E,A,ceil(w*h/8)] (device pointer, 2-byte aligned, borrowed).  1/8 of the bytes of the dense uint8
 * form: what a host consumer should pull over PCIe (casim_step_host_pipelined).  Needs submap_w and submap_h to be
 * multiples of 4.  NULL turns it off.  The dense local_grid buffer stays the parity format and is always written. */
int casim_set_grid_packed_output(casim_handle* h, uint8_t* packed);

/* Trajectory tables for ExternalDynamics (dynamics/ExternalDynamics.py:11-20): traj [E,A,T,4] =
 * (pose_x, pose_y, vel_x, vel_y).  The pointer is borrowed until replaced. */
int casim_set_trajectories(casim_handle* h, const float* traj, int T);

/* Replaces Agent.__init__ for every agent + the first _get_obs (agent.py:9-112,
 * collision_avoidance_env.py:257-266).  init: [E,A,7] = start_x, start_y, goal_x, goal_y, radius,
 * pref_speed, initial_heading (NaN = None -> atan2(goal-start), agent.py:29-33); type: [E,A] CASIM_TYPE words;
 * n_agents: [E].  Writes state [6,E,A,4] and the observation buffers.  The library keeps the three
 * input pointers (borrowed) for auto_reset. */
int casim_reset(casim_handle* h, float* state, const float* init, const int32_t* type,
                const int32_t* n_agents, const casim_obs* obs, void* stream);

/* Restart SOME environments: for every environment with mask[e] != 0 (device uint8 [E], e.g. the game_over buffer the
 * last casim_step wrote) redo what casim_reset does -- Agent.__init__ from the init / type / n_agents buffers given to
 * the last casim_reset (still borrowed; the caller may have rewritten the rows of those environments with a new
 * scenario) and the first _get_obs -- and leave every other environment's state and observations untouched.  This is
 * DummyVecEnv's "if done: obs = env.reset()" (envs/wrappers.py:101-106 MultiagentDummyVecEnv, stable_baselines
 * dummy_vec_env.step_wait) for the whole batch in one launch.  rewards / done / game_over / stats are not written. */
int casim_reset_envs(casim_handle* h, float* state, const uint8_t* mask, const casim_obs* obs, void* stream);

/* Recompute sensors/observations from a caller-modified state without stepping (_get_obs, :740-753). */
int casim_observe(casim_handle* h, const float* state, const casim_obs* obs, void* stream);

/* Replaces CollisionAvoidanceEnv.step (collision_avoidance_env.py:162-232): _take_action (in-kernel policies,
 * fp32 action rounding :289, Agent.take_action agent.py:147-190, dynamics), _compute_rewards +
 * _check_for_collisions (:502-567, :630-671), _get_obs (sensors), _check_which_agents_done (:711-738).
 * ext_actions: [E,A,2] caller-supplied actions (external / learning / CARRL agents); may be NULL if none.
 * policy_actions: [E,A,2] actions chosen by casim_explore_step / casim_ga3c_forward; may be NULL if none.
 * dt: the step length (step(actions, dt=None), :181-182); dt <= 0 means Config.DT. */
int casim_step(casim_handle* h, float* state, const float* ext_actions, const float* policy_actions,
               double dt, const casim_obs* obs, void* stream);

/* n_steps consecutive CollisionAvoidanceEnv.step calls in ONE launch, for batches whose every policy is evaluated inside
 * the step kernel (Static, NonCooperative, PedestrianDataset; policy_actions, if given, is applied at every step): the
 * agents' state stays in registers between the steps, which removes the per-step launch latency that bounds a small
 * batch (collision_avoidance_env.py:162-232 called n_steps times with actions = {}).  Results: rewards [n_steps,E,A],
 * done [n_steps,E,A], game_over [n_steps,E] and stats [n_steps,8] hold every step (the caller sizes these buffers of
 * `obs` accordingly); the observations (obs_misc, other_agents_states, laserscan, local_grid, collision_flags,
 * nearest_dist) and `state` are those after the last step.  auto_reset restarts finished environments between the
 * steps of the launch exactly as between two casim_step calls.  Bit-identical to n_steps casim_step calls. */
#define CASIM_STEP_N_MAX 64
int casim_step_n(casim_handle* h, float* state, const float* policy_actions, double dt, int n_steps, const casim_obs* obs,
                 void* stream);

/* Profiling / benchmark aid: launch ONE sensor kernel -- `which` = CASIM_SENSOR_LASERSCAN (LaserScanSensor.py:27-58)
 * or CASIM_SENSOR_OCCUPANCY_GRID (OccupancyGridSensor.py:70-98) -- on the post-move poses left by the last
 * casim_step / casim_reset / casim_observe of this handle, so that the kernel can be timed alone; the OR of the two =
 * the sensor launches of a step as casim_step issues them (with a map shared by all environments both observations
 * come out of one kernel that stages the map in shared memory once). */
int casim_run_sensor(casim_handle* h, const float* state, const casim_obs* obs, int which, void* stream);

/* Host-buffer variant: ext_actions_host (pinned, may be NULL) is copied H2D, the step runs, and rewards, done and
 * game_over (and obs_misc/sensor buffers if the host pointers are non-NULL) are delivered into `host`; blocks until
 * the results are in host memory.  `dev` holds the device buffers used for the step.
 * Delivery: per-agent scalars (rewards, done, game_over, collision_flags, nearest_dist, stats, obs_misc) whose host
 * buffer is page-locked + mapped (cudaHostAlloc / cudaHostRegister, e.g. torch pin_memory) and no larger than
 * CASIM_ZEROCOPY_MAX bytes are stored by the step kernel directly into host memory over PCIe -- their `dev`
 * counterparts are then NOT updated by this call; everything else (sensor planes, pageable buffers) is copied with
 * cudaMemcpyAsync from `dev`. */
int casim_step_host(casim_handle* h, float* state, const float* ext_actions_host, float* ext_actions_dev,
                    const float* policy_actions, double dt, const casim_obs* dev, const casim_obs* host,
                    void* stream);

/* Pipelined host-buffer step for consumers that want the observations in host memory every step (the reference hands
 * them out as numpy arrays, agent.py:244-248).  Same arguments as casim_step_host plus an optional bit-packed
 * local_grid pair (see casim_set_grid_packed_output; when given, the dense local_grid stays on the device and only
 * 1/8 of its bytes crosses PCIe).  Call t ENQUEUES step t (kernels on `stream`, then the D2H copies of its large results on
 * an internal copy stream, where they overlap the kernels of step t+1) and returns as soon as step t-1 is complete in
 * the host buffers that were passed to call t-1 -- ONE STEP OF LATENCY: the caller alternates between two sets of `dev`
 * observation buffers / `host` buffers / packed buffers (set t & 1), reads set (t-1) & 1 after call t returns, and must be
 * done with a set before passing it again.  `state` is single (stepped in place).  casim_host_flush waits for the last
 * step issued.  Not to be mixed with casim_step_host on the same handle while steps are in flight. */
int casim_step_host_pipelined(casim_handle* h, float* state, const float* ext_actions_host, float* ext_actions_dev,
                              const float* policy_actions, double dt, const casim_obs* dev, const casim_obs* host,
                              uint8_t* grid_packed_dev, uint8_t* grid_packed_host, void* stream);
int casim_host_flush(casim_handle* h);

/* Replaces ig_greedy.find_next_action for every ig_greedy agent (policies/ig_greedy.py:31-93,
 * information_models/targetMap.py:43-147, edfMap.py:14-47): detector emulation from other_agents_states,
 * target-map odds update over the visible cells, one-step greedy information-gain argmax over 9 candidate
 * actions.  belief: [E,A,ig_map_dim,ig_map_dim] fp32 odds (row j = y cell, col i = x cell), RMW.
 * policy_actions [E,A,2] receives the chosen action; best_mi [E,A] (optional) the winning score;
 * cand_mi [E,A,9] (optional) every candidate's score (-1 = infeasible);
 * counters [4] (optional, uint64): cells tested, EDF samples, visible cells, clamped samples. */
int casim_explore_step(casim_handle* h, const float* state, const float* other_agents_states, float* belief,
                       float* policy_actions, float* best_mi, float* cand_mi, unsigned long long* counters,
                       void* stream);

/* The reference's seeded training-scenario generators, one environment per seed, written into the device buffers
 * that casim_reset reads: envs/test_cases.py:1192-1281 train_agents_swap_circle(number_of_agents, seed=seeds[e]) and
 * :1365-1463 train_agents_random_positions(...), :1283-1363 train_agents_pairwise_swap(...).  A seed yields the reference's own scenario (numpy's and CPython's
 * MT19937 streams are reproduced); seeds must be non-zero (`if seed:`, :1198).  2 * (max(number_of_agents, 2) / 2)
 * agents per environment (<= A); ego_policy / ego_dynamics describe agent 0 (a GA3C-CADRL ego gets
 * UnicycleDynamicsMaxAcc, :1254-1257), the other agents are RVO / NonCooperative with other_dynamics.
 * seeds [E] uint32, init [E,A,7] fp32, types [E,A] int32, n_agents [E] int32, coop [E,A] fp32 cooperation_coef
 * (optional): all device pointers. */
enum casim_scenario { CASIM_SCENARIO_SWAP_CIRCLE = 0, CASIM_SCENARIO_RANDOM_POSITIONS = 1, CASIM_SCENARIO_PAIRWISE_SWAP = 2 };
int casim_generate_scenario(casim_handle* h, int kind, int number_of_agents, const uint32_t* seeds, int ego_policy,
                            int ego_dynamics, int other_dynamics, float* init, int32_t* types, int32_t* n_agents,
                            float* coop, void* stream);

/* Replaces ig_mcts.update_belief for every ig_* agent that is not done (policies/ig_mcts.py:117-137, called in
 * agent order by collision_avoidance_env.py:341-368): the agent's target map is updated from its own pose and
 * detections (find_targets_in_obs, :139-156) and from the pose of every other explorer j of its environment with
 * j's obsvd_targets -- this step's list when j < i and j is not done, else the list of j's previous update
 * (targetMap.update with several poses, information_models/targetMap.py:86-128).
 * targets_prev / targets_new: [E,A,32,2] fp64 global-frame detections with counts n_prev / n_new [E,A]; the caller
 * swaps the two pairs between steps (all zero before the first step).  belief: [E,A,dim,dim] fp32 odds, RMW.
 * team_mask [E,A,128] (team_obsv_cells as a cell mask, layout of casim_explore_rollouts) and team_reward [E,A]
 * (:137) are optional. */
int casim_explore_team_update(casim_handle* h, const float* state, const float* other_agents_states, float* belief,
                              const double* targets_prev, const int32_t* n_prev, double* targets_new, int32_t* n_new,
                              uint32_t* team_mask, float* team_reward, void* stream);

/* Dec-MCTS simulation rollouts for every ig_* agent (policy id CASIM_POLICY_IG_GREEDY): replaces the body of
 * pydecmcts Tree._simulate (policies/pydecmcts/DecMCTS.py:233-269) with ig_mcts's callbacks -- get_next_pose
 * (policies/ig_mcts.py:173-198: first-order step, strict in-map test, EDF clearance > radius + 0.1),
 * mcts_sim_state_storer (:223-244: an infeasible step keeps the pose and observes nothing) and mcts_reward
 * (:246-253: mutual information of the cells observed along the path minus the cells the other robots'
 * paths claim; targetMap.py:43-84, 130-143).  K rollouts per agent, evaluated from the agent's current pose:
 * a tree node's state followed by its rollout is the root-to-leaf action sequence replayed here.
 * actions: [E,A,K,H] uint8, action index = 3 * vel_idx + dphi_idx into vel_list (0, 2, 4) x dphi_list
 *   (-pi/2, 0, pi/2) (:259-265); a value > 8 ends the sequence.  H <= CASIM_ROLLOUT_MAX_H; xdt = ig_mcts.xDT.
 * belief: [E,A,dim,dim] fp32 odds (read only).
 * exclude: cell masks of the other robots' observed cells, 128 uint32 per mask, bit (j * 64 + i) for cell
 *   (x cell i, y cell j); [E,A,128] shared by an agent's K rollouts, or [E,A,K,128] if exclude_per_rollout;
 *   NULL = nothing claimed.
 * rewards: [E,A,K] (0 for non-explorers).  end_pose [E,A,K,3] and obsvd [E,A,K,128] (the state's obsvd_cells
 * as a mask, BEFORE the difference -- what a robot communicates) and n_feasible [E,A,K] (int32: how many steps of
 * the sequence were applied; == the sequence length iff mcts_state_storer would accept every step, :200-221) are
 * optional. */
/* Tabulated mutual information for the rollouts.  A planner evaluates thousands of rollouts per agent on ONE belief
 * (ig_mcts.find_next_action grows its tree between two belief updates, ig_mcts.py:79-109), and mcts_reward's per-cell term
 * (targetMap.py:130-143) depends on the cell's belief alone.  casim_explore_mi_table writes that term for every cell of
 * every ig_* agent, in the 2^-44 fixed point the rewards are summed in, to table [E,A,dim,dim] int64 (device);
 * casim_explore_set_mi_table(h, table) makes casim_explore_rollouts read it instead of taking two logarithms per observed
 * cell (same integers, same rewards); NULL switches back.  The table belongs to the caller and must match the belief
 * passed to casim_explore_rollouts. */
int casim_explore_mi_table(casim_handle* h, const float* state, const float* belief, int64_t* table, void* stream);
int casim_explore_set_mi_table(casim_handle* h, const int64_t* table);
/* Restrict casim_explore_rollouts to a list of agent slots (e * A + a, int32, DEVICE memory, each a valid slot, no
 * duplicates): only those slots get CTAs and only their [slot][k] results are written -- a planner that grows the trees of
 * one explorer per environment (collision_avoidance_env.py:341-379 plans the robots one after the other) does not launch the
 * other three quarters of the batch.  NULL = every slot again.  The list belongs to the caller. */
int casim_explore_set_rollout_slots(casim_handle* h, const int32_t* slots, int n_slots);
#define CASIM_ROLLOUT_MAX_H 32
#define CASIM_ROLLOUT_PROBE 254 /* action value that ends a sequence like any value > 8 and marks it as an expansion probe: only
                                   n_feasible is wanted (DecMCTS.py:201-231 asks whether path + action is feasible, not what it
                                   sees), so the visibility / reward part is skipped: rewards = 0, obsvd = 0 */
#define CASIM_CELL_MASK_WORDS 128
int casim_explore_rollouts(casim_handle* h, const float* state, const float* belief, const uint8_t* actions, int K,
                           int H, int xdt, const uint32_t* exclude, int exclude_per_rollout, float* rewards,
                           float* end_pose, uint32_t* obsvd, int32_t* n_feasible, void* stream);

/* Dec-MCTS tree bookkeeping on the device.  Replaces, for T trees grown in lock-step,
 * the parts of pydecmcts `Tree.grow` (policies/pydecmcts/DecMCTS.py:273-360) around the simulations:
 * casim_mcts_select = _get_system_state (:182-194) + _select / _UCT (:140-153, :14-18) + the action sequences of the
 * selected node's 9 one-step extensions and `nsims` random completions (_expansion :201-231, _simulate :233-269), written
 * into the action buffer of casim_explore_rollouts (K = 9 + nsims rollouts per tree, at the tree's agent slot);
 * casim_mcts_update = _expansion (a child per feasible action, in action order) + the node's statistics (:323-339) +
 * the discounted back-propagation (:341-357), from casim_explore_rollouts' rewards / n_feasible.
 * All arrays are DEVICE memory owned by the caller; node 0 is the root (parent -1, stage 0, everything else zero, path
 * and best_seq filled with 255); nodes are numbered in creation order like the reference's graph.
 *   draws  [T,nsims,H] uint8: the random action (0..8) of every simulation step (ig_mcts.py:257; only d >= stage is used);
 *          NULL = drawn in the kernel: splitmix64(draw_seed, ((iteration * T + tree) * nsims + sim) * H + d) scaled to 0..8;
 *   comm_u [n_comms,T] double in [0,1): one draw per communicating robot (np.random.choice(n, p=q), DecMCTS.py:190);
 *   actions [E,A,K,H] uint8 / exclude [E,A,128] (or NULL: no communication): only the trees' slots are written;
 *   root_only = 1: no descent, no simulations (Tree.__init__'s _expansion of the root, :133); expand_only likewise.
 * A tree whose node array is full sets *overflow = 1 and is left unchanged. */
#define CASIM_MCTS_ACTIONS 9
typedef struct casim_mcts_trees {
  int32_t T, M, H, nsims;      /* trees, node capacity per tree, horizon (<= CASIM_ROLLOUT_MAX_H), simulations per grow */
  int32_t n_comms, comm_n;     /* communicating robots, entries per action distribution (DecMCTS.py:21-59) */
  double c_p, gamma;           /* exploration constant (:14-18), back-propagation discount (:341) */
  const int32_t* slot;         /* [T] agent slot e * A + a of each tree */
  int32_t* n_nodes;            /* [T] */
  int32_t* node;               /* [T] the node casim_mcts_select chose, consumed by casim_mcts_update */
  int32_t* parent;             /* [T,M] */
  int32_t* child;              /* [T,M,9] children in creation order, -1 = none */
  uint8_t* nchild;             /* [T,M] */
  uint8_t* stage;              /* [T,M] depth = length of the node's path */
  uint8_t* has_rollout;        /* [T,M] best_seq is set */
  double* mu;                  /* [T,M] */
  double* N;                   /* [T,M] discounted visit count */
  double* best_reward;         /* [T,M] */
  uint8_t* path;               /* [T,M,H] action sequence from the root, 255 padded */
  uint8_t* best_seq;           /* [T,M,H] best rollout through the node (requested actions, full length) */
  const double* comm_q;        /* [n_comms,T,comm_n] normalised weights */
  const uint32_t* comm_masks;  /* [n_comms,T,comm_n,128] observed-cell masks of the distributions' rollouts */
  const uint8_t* comm_valid;   /* [n_comms,T] */
  int32_t* overflow;           /* [1] */
} casim_mcts_trees;
int casim_mcts_select(casim_handle* h, const casim_mcts_trees* trees, const uint8_t* draws, uint64_t draw_seed,
                      int iteration, const double* comm_u, uint8_t* actions, uint32_t* exclude, int root_only, void* stream);
int casim_mcts_update(casim_handle* h, const casim_mcts_trees* trees, const uint8_t* actions, const float* rewards,
                      const int32_t* n_feasible, double* avg_out, int expand_only, void* stream);

/* Replaces GA3CCADRLPolicy.find_next_action for every GA3C agent (policies/GA3CCADRLPolicy.py:34-106,
 * GA3C_CADRL/network.py:45-98).  weights: fp32 parameter blob, order kernel/bias of lstm [71,256],[256];
 * layer1 [68,256],[256]; layer2 [256,256],[256]; fullyconnected1 [256,256],[256]; logits_p [256,11],[11]
 * (row-major [in,out] like TF).  Writes policy_actions [E,A,2]; optionally action_idx [E,A] (int32, -1 for
 * non-GA3C agents), probs [E,A,11] and nn_input [E,A,75].
 * use_tensor_cores: bit 0 = tcgen05 forward with fp16 operands and fp32 accumulation (else the fp32 SIMT forward); bit 1 = profiling aid, skip the
 * state-vector kernel and run the forward on the state vectors of the previous call; bit 2 (with bit 0) = "tensor cores,
 * exact decisions": every agent whose tensor-core top-2 probability margin is too small to call is re-evaluated by the fp32
 * forward in the same call, so that action_idx / policy_actions equal the fp32 forward's for every agent (the reference
 * runs this network in fp32, GA3C_CADRL/network.py:45-98). */
int casim_ga3c_set_weights(casim_handle* h, const float* weights_host, int n_params, void* stream);
int casim_ga3c_forward(casim_handle* h, const float* state, float* policy_actions, int32_t* action_idx,
                       float* probs, float* nn_input, int use_tensor_cores, void* stream);

/* RVO policy.  Replaces RVOPolicy.find_next_action (RVOPolicy.py:53-117) and the
 * ego agent's share of rvo2.PyRVOSimulator.doStep (Python-RVO2/src/RVOSimulator.cpp:163-183, Agent.cpp:43-56,
 * 286-631) for every agent whose policy id is CASIM_POLICY_RVO: writes [pref_speed, delta_heading] into
 * policy_actions [E,A,2] (other agents' entries are left untouched); pass the buffer to casim_step.
 * coop: [E,A] Agent.cooperation_coef (agent.py:10,103; RVOPolicy.py:85) or NULL = 1.0.  params: NULL = defaults.
 * casim_rvo_set_obstacles replaces sim.addObstacle (RVOPolicy.py:56-57, RVOSimulator.cpp:124-161) +
 * sim.processObstacles (RVOPolicy.py:44, KdTree.cpp:113-245, edge splitting included) for ONE obstacle set shared by
 * all environments: n_polys polygons, poly_sizes[p] vertices each, counter-clockwise, vertices [sum][2] doubles in
 * HOST memory (synchronous; at most 64 vertices after splitting; n_polys = 0 removes the set).  The obstacle ORCA
 * lines are Agent.cpp:66-284.  A handle with a static map refuses casim_rvo_step until an obstacle set is given. */
void casim_rvo_default_params(casim_rvo_params* params);
int casim_rvo_set_obstacles(casim_handle* h, const double* vertices, const int32_t* poly_sizes, int n_polys);
int casim_rvo_step(casim_handle* h, const float* state, const float* coop, const casim_rvo_params* params,
                   float* policy_actions, void* stream);

/* Flattened observation wire format.  Replaces
 * MultiagentFlattenDictWrapper.observation (envs/wrappers.py:37-45) for all environments at once: out [E,A,S] float32,
 * for each agent slot the listed keys' values ravel()ed and concatenated in the given order (S = casim_flat_row_size);
 * absent agent slots are zero (collision_avoidance_env.py:125-134).  out viewed as [E, A*S] is the VecEnv observation
 * array (experiments/src/env_utils.py:7-39); local_grid cells become 0.0 / 1.0.  keys is a HOST array. */
enum casim_obs_key {           /* names of Config.STATE_INFO_DICT, envs/config.py:104-215 */
  CASIM_KEY_DIST_TO_GOAL = 0, CASIM_KEY_REL_GOAL, CASIM_KEY_RADIUS, CASIM_KEY_HEADING_EGO_FRAME,
  CASIM_KEY_HEADING_GLOBAL_FRAME, CASIM_KEY_POS_GLOBAL_FRAME, CASIM_KEY_PREF_SPEED, CASIM_KEY_NUM_OTHER_AGENTS,
  CASIM_KEY_OTHER_AGENT_STATES, CASIM_KEY_OTHER_AGENTS_STATES, CASIM_KEY_LOCAL_GRID, CASIM_KEY_LASERSCAN,
  CASIM_KEY_USE_PPO, CASIM_KEY_COUNT
};
int casim_flat_row_size(const casim_handle* h, const int32_t* keys, int n_keys);
int casim_flatten_obs(casim_handle* h, const float* state, const casim_obs* obs, const int32_t* keys, int n_keys,
                      float* out, void* stream);

/* Introspection for benchmarks: number of kernel launches issued through this handle so far. */
unsigned long long casim_launch_count(const casim_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* CASIM_H */
