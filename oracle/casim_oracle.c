/*
 * casim_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C, double-precision, scalar restatement of the reference's stepping path
 * (mlodel/gym-exploration-2d, gym_collision_avoidance/envs/...).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and only as the checker / CPU baseline; libcasim.so never links or calls it.
 *
 * Parity status: PINNED for rows a1-a22 of SURVEY.md section 8(a) against outputs of the
 * reference itself executed in the build container (oracle/gen_golden.py ->
 * tests/golden/ (npz files); tests/test_oracle_golden.py) and against the known answers of
 * SURVEY.md Appendix B.  The GA3C-CADRL network numerics (a24): TensorFlow 1.15 is not
 * installable here and the reference holds no test vectors for it; casim_oracle_ga3c.c restates
 * GA3C_CADRL/network.py:45-98 + TF1 LSTMCell semantics and is pinned (round 2) against an independent
 * numpy float64 evaluation of the TF graph on the reference's shipped IROS18 weights
 * (oracle/gen_golden_ga3c_probs.py -> tests/golden/ga3c_probs.npz) -- not against TensorFlow itself.
 *
 * Arithmetic: the reference is float64 numpy with the chosen actions rounded to float32
 * (collision_avoidance_env.py:289).  This file follows that (numpy-1.x promotion rules; the
 * golden generator passes np.float64 headings/dt so that numpy 2 does the same).
 * Compile with -ffp-contract=off: the reference never fuses multiply-adds.
 *
 * Each function cites the reference file:line it restates (paths relative to
 * gym_collision_avoidance/envs/).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/casim.h"

#define NF CASIM_STATE_WORDS
#define MD CASIM_MAP_DIM

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ---------------------------------------------------------------------------------------- */
/* util.py:27-32                                                                             */
static double wrap(double a) {
  while (a >= M_PI) a -= 2 * M_PI;
  while (a < -M_PI) a += 2 * M_PI;
  return a;
}

static double sgn(double x) { return (x > 0) - (x < 0); }

/* ---------------------------------------------------------------------------------------- */
/* Reference defaults: config.py:29-95, LaserScanSensor.py:13-20, collision_avoidance_env.py:  */
/* 97,755-774, ig_greedy.py:18-29, experiments/src/dmcts.py:73-78.                           */
void casim_oracle_default_config(casim_config* c) {
  memset(c, 0, sizeof(*c));
  c->dt = 0.1;
  c->reward_at_goal = 3.0;
  c->reward_collision_with_agent = -10.0;
  c->reward_timeout = -10.0;
  c->reward_infeasible = 0.0;
  c->reward_collision_with_wall = -0.25;
  c->reward_time_step = -0.01;
  c->reward_distance_to_goal = 0.0;
  c->reward_wiggly_behavior = 0.0;
  c->wiggly_behavior_threshold = 0.0;
  c->reward_min = -10.0; /* min/max of {at_goal, coll_agent, time_step, coll_wall, wiggly} */
  c->reward_max = 3.0;
  c->getting_close_range = 0.2;
  c->near_goal_threshold = 0.75;
  c->max_time_ratio = 3.0;
  c->sensing_horizon = INFINITY;
  c->max_heading_change = 4.0;
  c->map_width = 30.0;
  c->map_height = 30.0;
  c->map_cell = 0.1;
  c->laser_max_range = 6.0;
  c->laser_norm_range = 6.0;
  c->laser_num_beams = 16;
  {
    /* np.linspace(-pi, pi, 16): step = (stop-start)/div; y = arange(num)*step + start; y[-1] = stop */
    int n = c->laser_num_beams;
    double start = -M_PI, stop = M_PI, step = (stop - start) / (double)(n - 1);
    for (int i = 0; i < n; i++) c->laser_angles[i] = (double)i * step + start;
    c->laser_angles[n - 1] = stop;
    /* np.arange(0, 6, 2*pi/16): len = ceil((stop-start)/step); value i = start + i*delta */
    double rstep = 2 * M_PI / 16.0;
    int nr = (int)ceil((6.0 - 0.0) / rstep);
    c->laser_num_ranges = nr;
    for (int i = 0; i < nr; i++) c->laser_ranges[i] = 0.0 + (double)i * rstep;
  }
  c->ig_sens_fov = 60.0 * M_PI / 180.0;
  c->ig_detect_fov = 60.0;
  c->ig_range = 5.0;
  c->ig_cell = 0.1 * 5;
  c->ig_rocc = 1.5;
  c->ig_remp = 0.66;
  c->ig_tolerance = 0.01;
  c->ig_p_false_neg = 0.1;
  c->ig_p_false_pos = 0.05;
  c->ig_prior = 1.0;
  c->ig_dt = 0.1;
  c->collision_av_w_static_agent = 0;
  c->max_other_agents_observed = 9;
  c->max_other_agents_rows = 9;
  c->submap_w = 60;
  c->submap_h = 60;
  c->game_over_mode = CASIM_GAME_OVER_AGENT0;
  c->auto_reset = 0;
  c->ig_map_dim = 60;
}

/* ---------------------------------------------------------------------------------------- */
/* Map.py:40-47  world_coordinates_to_map_indices                                            */
static int world_to_map(const casim_config* c, double x, double y, int* row, int* col) {
  double ox = (c->map_width / 2.) / c->map_cell, oy = (c->map_height / 2.) / c->map_cell;
  int nrow = (int)(c->map_width / c->map_cell), ncol = (int)(c->map_height / c->map_cell);
  double fr = floor(ox - y / c->map_cell), fc = floor(oy + x / c->map_cell);
  /* int(np.floor(.)) of a huge/NaN value raises in Python; clamp so that it is just "outside" */
  if (!(fr > -1e9 && fr < 1e9)) fr = -1e9;
  if (!(fc > -1e9 && fc < 1e9)) fc = -1e9;
  *row = (int)fr;
  *col = (int)fc;
  return *row >= 0 && *col >= 0 && *row < nrow && *col < ncol;
}

/* Map.py:107-123  get_occupancy_grid: rects [n,4] = corner1 (x,y), corner3 (x,y) */
void casim_oracle_rasterize(const casim_config* c, const double* rects, int n, uint8_t* map) {
  memset(map, 0, MD * MD);
  for (int k = 0; k < n; k++) {
    int r0, c0, r1, c1;
    world_to_map(c, rects[4 * k + 0], rects[4 * k + 1], &r0, &c0);
    world_to_map(c, rects[4 * k + 2], rects[4 * k + 3], &r1, &c1);
    for (int ii = r0; ii <= r1; ii++)
      for (int jj = c0; jj <= c1; jj++) {
        /* numpy would wrap negative indices / raise beyond the edge; obstacles are kept in-map */
        if (ii >= 0 && ii < MD && jj >= 0 && jj < MD) map[ii * MD + jj] = 1;
      }
  }
}

/* Map.py:67-71 get_agent_map_indices: is cell (row, col) inside the disc of an agent whose
 * centre cell is (r0, c0)?  (x - pos[1])**2 + (y - pos[0])**2 < (radius/cell)**2, fp64. */
static int in_disc(const casim_config* c, int row, int col, int r0, int c0, double radius) {
  double rc = radius / c->map_cell;
  double dx = (double)(col - c0), dy = (double)(row - r0);
  return dx * dx + dy * dy < rc * rc;
}

/* collision_avoidance_env.py:656-666: wall collision of one agent */
static int collides_with_wall(const casim_config* c, const uint8_t* map, double px, double py,
                              double radius) {
  int r0, c0;
  if (!world_to_map(c, px, py, &r0, &c0)) return 0;
  int span = (int)ceil(radius / c->map_cell) + 1;
  for (int row = r0 - span; row <= r0 + span; row++) {
    if (row < 0 || row >= MD) continue;
    for (int col = c0 - span; col <= c0 + span; col++) {
      if (col < 0 || col >= MD) continue;
      if (map[row * MD + col] && in_disc(c, row, col, r0, c0, radius)) return 1;
    }
  }
  return 0;
}

/* ---------------------------------------------------------------------------------------- */
/* agent.py:250-269 get_ref + dynamics/Dynamics.py:14-28 update_ego_frame                    */
static void get_ref_axes(const double* s, double* prll, double* orth, double* dist_out) {
  double gdx = s[CASIM_GX] - s[CASIM_PX], gdy = s[CASIM_GY] - s[CASIM_PY];
  double dist = sqrt(gdx * gdx + gdy * gdy);
  if (dist > 1e-8) {
    prll[0] = gdx / dist;
    prll[1] = gdy / dist;
  } else {
    prll[0] = gdx;
    prll[1] = gdy;
  }
  orth[0] = -prll[1];
  orth[1] = prll[0];
  if (dist_out) *dist_out = dist;
}

static void update_ego_frame(double* s) {
  double prll[2], orth[2], dist;
  get_ref_axes(s, prll, orth, &dist);
  s[CASIM_PAST_DIST_TO_GOAL] = s[CASIM_DIST_TO_GOAL]; /* agent.py:260 */
  s[CASIM_DIST_TO_GOAL] = dist;                       /* :261 */
  if (s[CASIM_T] == 0) s[CASIM_PAST_DIST_TO_GOAL] = dist; /* :262-263 */
  double ref_angle = atan2(prll[1], prll[0]);
  s[CASIM_HEADING_EGO] = wrap(s[CASIM_HEADING] - ref_angle); /* Dynamics.py:19-20 */
  /* vel_ego_frame and rel_goal are pure functions of the stored words; emitted as observations */
}

/* agent.py:9-112 Agent.__init__ for one agent.  init = sx, sy, gx, gy, radius, pref_speed, heading(NaN=None) */
static void agent_init(const casim_config* c, const double* init, int type, double* s) {
  memset(s, 0, NF * sizeof(double));
  s[CASIM_PX] = init[0];
  s[CASIM_PY] = init[1];
  s[CASIM_GX] = init[2];
  s[CASIM_GY] = init[3];
  s[CASIM_RADIUS] = init[4];
  s[CASIM_PREF_SPEED] = init[5];
  if (isnan(init[6]))
    s[CASIM_HEADING] = atan2(init[3] - init[1], init[2] - init[0]); /* :29-31 */
  else
    s[CASIM_HEADING] = init[6];
  double dx = init[0] - init[2], dy = init[1] - init[3];
  double straight = (sqrt(dx * dx + dy * dy) - c->near_goal_threshold) / init[5]; /* :59 */
  s[CASIM_TIME_REMAINING] = c->max_time_ratio * straight;                        /* :60-63 */
  s[CASIM_TYPE] = (double)type;
  update_ego_frame(s); /* :94 (t == 0 -> past_dist = dist) */
}

/* ---------------------------------------------------------------------------------------- */
/* Policies evaluated inside _take_action (collision_avoidance_env.py:298-320).              */
/* Returns the fp32-rounded action (the all_actions array is float32, :289).                 */
static const double CARRL_MAX_HEADING = M_PI / 6;

static void choose_action(const casim_config* c, const double* s, const double* ext, const double* pol,
                          float* a) {
  int type = (int)s[CASIM_TYPE];
  int policy = type & 0xff;
  int flags = (int)s[CASIM_FLAGS];
  a[0] = 0.f;
  a[1] = 0.f;
  if (flags & CASIM_F_IS_DONE) return; /* :299-300 */
  switch (policy) {
    case CASIM_POLICY_STATIC: /* StaticPolicy.py:9-12 */
      break;
    case CASIM_POLICY_NONCOOPERATIVE: /* NonCooperativePolicy.py:10-13 */
      a[0] = (float)s[CASIM_PREF_SPEED];
      a[1] = (float)(-s[CASIM_HEADING_EGO]);
      break;
    case CASIM_POLICY_EXTERNAL:
      if (ext) {
        a[0] = (float)ext[0];
        a[1] = (float)ext[1];
      }
      break;
    case CASIM_POLICY_LEARNING: /* LearningPolicy.py:11-16 */
      if (ext) {
        double heading = c->max_heading_change * (2. * ext[1] - 1.);
        double speed = s[CASIM_PREF_SPEED] * ext[0];
        a[0] = (float)speed;
        a[1] = (float)heading;
      }
      break;
    case CASIM_POLICY_CARRL: /* CARRLPolicy.py:7-15: actions[:,0]=1, actions[:,1]=linspace(-pi/6, pi/6, 11) */
      if (ext) {
        int k = (int)ext[0];
        if (k < 0) k += 11; /* numpy negative index */
        if (k < 0) k = 0;
        if (k > 10) k = 10;
        double step = (CARRL_MAX_HEADING - (-CARRL_MAX_HEADING)) / 10.0;
        double w = (k == 10) ? CARRL_MAX_HEADING : (double)k * step + (-CARRL_MAX_HEADING);
        a[0] = 1.0f;
        a[1] = (float)w;
      }
      break;
    case CASIM_POLICY_IG_GREEDY:
    case CASIM_POLICY_GA3C_CADRL:
    case CASIM_POLICY_RVO: /* RVOPolicy.py:53-117: action computed by casim_oracle_rvo_actions */
      if (pol) {
        a[0] = (float)pol[0];
        a[1] = (float)pol[1];
      }
      break;
    case CASIM_POLICY_PEDESTRIAN: /* PedestrianDatasetPolicy.py:15-16 */
    default:
      break;
  }
}

/* ---------------------------------------------------------------------------------------- */
/* Dynamics: UnicycleDynamics.py:10-31, FirstOrderDynamics.py:10-23,                          */
/* UnicycleDynamicsMaxTurnRate.py:11-26, ExternalDynamics.py:8-23 + agent.py:125-145          */
static void dynamics_step(const casim_config* c, double* s, const float* a, double dt, const double* traj,
                          int T) {
  int dyn = (((int)s[CASIM_TYPE]) >> 8) & 0xff;
  if (dyn == CASIM_DYN_EXTERNAL) {
    if (!traj || T <= 0) return;
    int step_num = (int)s[CASIM_STEP_NUM];
    int k = step_num < T - 1 ? step_num : T - 1; /* ExternalDynamics.py:11 */
    const double* row = traj + 4 * k;            /* pose_x, pose_y, vel_x, vel_y */
    /* :14-16 (x and y swapped, sic) */
    double heading = atan2(row[2], row[3]) - s[CASIM_HEADING];
    double px = row[0], py = row[1];
    /* agent.py:125-145 set_state(px, py, heading=heading) */
    if (step_num == 0) {
      s[CASIM_VX] = 0;
      s[CASIM_VY] = 0;
    } else {
      s[CASIM_VX] = (px - s[CASIM_PX]) / c->dt; /* dt_nominal */
      s[CASIM_VY] = (py - s[CASIM_PY]) / c->dt;
    }
    s[CASIM_DELTA_HEADING] = wrap(heading - s[CASIM_HEADING]);
    s[CASIM_PX] = px;
    s[CASIM_PY] = py;
    s[CASIM_SPEED] = sqrt(s[CASIM_VX] * s[CASIM_VX] + s[CASIM_VY] * s[CASIM_VY]);
    s[CASIM_HEADING] = heading; /* not wrapped, sic */
    return;
  }
  double speed = (double)a[0];
  double sel;
  if (dyn == CASIM_DYN_UNICYCLE_MAX_ACC) { /* UnicycleDynamicsMaxAcc.py:17-41 (kp 2, limits 3 rad/s, 2 m/s^2, 3 rad/s^2, 1.1 m/s) */
    double cur_speed = s[CASIM_SPEED], cur_rate = s[CASIM_TURNING_DIR];
    double rate = fmin(fmax((double)a[1] / dt, -3.0), 3.0);
    double lin_acc = fmin(fmax(2.0 * ((double)a[0] - cur_speed), -2.0), 2.0);
    double turn_acc = fmin(fmax(2.0 * (rate - cur_rate), -3.0), 3.0);
    cur_speed += lin_acc * dt;
    cur_speed = fmin(fmax(cur_speed, -1.1), 1.1);
    cur_rate += turn_acc * dt;
    s[CASIM_TURNING_DIR] = cur_rate;
    speed = cur_speed;
    sel = wrap(cur_rate * dt + s[CASIM_HEADING]);
  } else if (dyn == CASIM_DYN_UNICYCLE_SECOND_ORDER_EULER) { /* UnicycleSecondOrderEulerDynamics.py:12-31 */
    double vn = sqrt(s[CASIM_VX] * s[CASIM_VX] + s[CASIM_VY] * s[CASIM_VY]);
    speed = fmin(fmax(vn + (double)a[0] * dt, 0.0), 1.0);
    double ang = fmin(fmax(s[CASIM_TURNING_DIR] + (double)a[1] * dt, -3.0), 3.0);
    s[CASIM_TURNING_DIR] = ang;
    sel = wrap(ang * dt + s[CASIM_HEADING]);
  } else if (dyn == CASIM_DYN_FIRST_ORDER)
    sel = wrap((double)a[1] * dt + s[CASIM_HEADING]);
  else if (dyn == CASIM_DYN_UNICYCLE_MAX_TURN_RATE) {
    double rate = (double)a[1] / dt;
    if (rate < -3.0) rate = -3.0;
    if (rate > 3.0) rate = 3.0;
    sel = wrap(rate * dt + s[CASIM_HEADING]);
  } else
    sel = wrap((double)a[1] + s[CASIM_HEADING]);
  double cs = cos(sel), sn = sin(sel);
  double dx = speed * cs * dt, dy = speed * sn * dt;
  s[CASIM_PX] += dx;
  s[CASIM_PY] += dy;
  s[CASIM_VX] = speed * cs;
  s[CASIM_VY] = speed * sn;
  s[CASIM_SPEED] = speed;
  s[CASIM_DELTA_HEADING] = wrap(sel - s[CASIM_HEADING]);
  s[CASIM_HEADING] = sel;
  if (dyn == CASIM_DYN_UNICYCLE) { /* UnicycleDynamics.py:26-31 turning_dir */
    double td = s[CASIM_TURNING_DIR];
    if (fabs(td) < 1e-5)
      td = 0.11 * sgn(sel);
    else if (td * sel < 0) {
      double v = -td + sel;
      td = fmax(-M_PI, fmin(M_PI, v));
    } else
      td = sgn(td) * fmax(0.0, fabs(td) - 0.1);
    s[CASIM_TURNING_DIR] = td;
  }
}

/* agent.py:147-190 Agent.take_action */
static void take_action(const casim_config* c, double* s, const float* a, double dt, const double* traj,
                        int T) {
  int flags = (int)s[CASIM_FLAGS];
  if (flags & (CASIM_F_IS_AT_GOAL | CASIM_F_RAN_OUT_OF_TIME | CASIM_F_IN_COLLISION)) { /* :148-159 */
    if (flags & CASIM_F_IS_AT_GOAL) flags |= CASIM_F_WAS_AT_GOAL_ALREADY;
    if (flags & CASIM_F_IN_COLLISION) flags |= CASIM_F_WAS_IN_COLLISION_ALREADY;
    if (!(flags & CASIM_F_IS_AT_GOAL)) s[CASIM_T] += dt;
    s[CASIM_VX] = 0;
    s[CASIM_VY] = 0;
    s[CASIM_FLAGS] = (double)flags;
    return;
  }
  /* :162-163 */
  s[CASIM_PAST_ACT1_V] = s[CASIM_PAST_ACT0_V];
  s[CASIM_PAST_ACT1_W] = s[CASIM_PAST_ACT0_W];
  s[CASIM_PAST_ACT0_V] = (double)a[0];
  s[CASIM_PAST_ACT0_W] = (double)a[1];
  dynamics_step(c, s, a, dt, traj, T); /* :173 */
  update_ego_frame(s);                 /* :175 */
  /* :179 -> utils/end_conditions.py:3-6 */
  double ex = s[CASIM_PX] - s[CASIM_GX], ey = s[CASIM_PY] - s[CASIM_GY];
  if (ex * ex + ey * ey <= c->near_goal_threshold * c->near_goal_threshold)
    flags |= CASIM_F_IS_AT_GOAL;
  else
    flags &= ~CASIM_F_IS_AT_GOAL;
  /* :184-188 */
  s[CASIM_TIME_REMAINING] -= dt;
  s[CASIM_T] += dt;
  s[CASIM_STEP_NUM] += 1;
  if (s[CASIM_TIME_REMAINING] <= 0.0) flags |= CASIM_F_RAN_OUT_OF_TIME;
  s[CASIM_FLAGS] = (double)flags;
}

static int is_static(const double* s) { return (((int)s[CASIM_TYPE]) & 0xff) == CASIM_POLICY_STATIC; }

/* ---------------------------------------------------------------------------------------- */
/* Sensors                                                                                   */

/* OtherAgentsStatesSensor.py:11-77.  out: [rows,10]; returns num_other_agents_observed */
static int sense_other_agents(const casim_config* c, const double* env, int n, int i, double* out) {
  const double* h = env + i * NF;
  int rows = c->max_other_agents_rows;
  memset(out, 0, sizeof(double) * rows * 10);
  double key[CASIM_MAX_AGENTS];
  int idx[CASIM_MAX_AGENTS], m = 0;
  for (int j = 0; j < n; j++) {
    if (j == i) continue;
    const double* o = env + j * NF;
    double rx = o[CASIM_PX] - h[CASIM_PX], ry = o[CASIM_PY] - h[CASIM_PY];
    double dc = sqrt(rx * rx + ry * ry);
    if (dc > c->sensing_horizon) continue;
    key[m] = dc - h[CASIM_RADIUS] - o[CASIM_RADIUS];
    idx[m] = j;
    m++;
  }
  /* stable ascending insertion sort (sorted(..., key=itemgetter(1)), :28-29) */
  for (int a = 1; a < m; a++) {
    double k = key[a];
    int id = idx[a], b = a - 1;
    while (b >= 0 && key[b] > k) {
      key[b + 1] = key[b];
      idx[b + 1] = idx[b];
      b--;
    }
    key[b + 1] = k;
    idx[b + 1] = id;
  }
  /* reverse, keep the last MAX_NUM_OTHER_AGENTS_OBSERVED (:31-33) == first K of ascending, reversed */
  int K = m < c->max_other_agents_observed ? m : c->max_other_agents_observed;
  double prll[2], orth[2];
  get_ref_axes(h, prll, orth, NULL);
  int count = 0;
  for (int r = 0; r < K && r < rows; r++) {
    int j = idx[K - 1 - r];
    const double* o = env + j * NF;
    double rx = o[CASIM_PX] - h[CASIM_PX], ry = o[CASIM_PY] - h[CASIM_PY];
    double* w = out + 10 * r;
    w[0] = rx;
    w[1] = ry;
    w[2] = rx * prll[0] + ry * prll[1];
    w[3] = rx * orth[0] + ry * orth[1];
    w[4] = o[CASIM_VX] * prll[0] + o[CASIM_VY] * prll[1];
    w[5] = o[CASIM_VX] * orth[0] + o[CASIM_VY] * orth[1];
    w[6] = o[CASIM_RADIUS];
    w[7] = h[CASIM_RADIUS] + o[CASIM_RADIUS];
    w[8] = sqrt(rx * rx + ry * ry) - h[CASIM_RADIUS] - o[CASIM_RADIUS];
    w[9] = is_static(o) ? 1.0 : 2.0;
    count++;
  }
  return count;
}

/* LaserScanSensor.py:27-58 (+ Map.py:49-59, 73-79).  out: [num_beams] */
static void sense_laserscan(const casim_config* c, const uint8_t* map, const double* h, double* out) {
  int nb = c->laser_num_beams, nr = c->laser_num_ranges;
  int r0, c0;
  int ego_in_map = world_to_map(c, h[CASIM_PX], h[CASIM_PY], &r0, &c0);
  for (int b = 0; b < nb; b++) {
    double ang = c->laser_angles[b] + h[CASIM_HEADING];
    double cs = cos(ang), sn = sin(ang);
    int cum = 0;
    double range = c->laser_max_range;
    for (int k = 0; k < nr; k++) {
      double x = h[CASIM_PX] + c->laser_ranges[k] * cs;
      double y = h[CASIM_PY] + c->laser_ranges[k] * sn;
      int row, col;
      int inmap = world_to_map(c, x, y, &row, &col);
      int hit = 0;
      if (inmap && map && map[row * MD + col]) {
        int ego = ego_in_map && in_disc(c, row, col, r0, c0, h[CASIM_RADIUS]);
        hit = !ego;
      }
      cum += hit;
      if (cum == 1) range = c->laser_ranges[k]; /* last k with cumsum == 1 (:43-47) */
    }
    out[b] = 1 - range / c->laser_norm_range;
  }
}

/* Python-style banker's rounding to int for cv::saturate_cast<int>(double) == cvRound == rint */
static int cv_round(double v) { return (int)nearbyint(v); }

/* OccupancyGridSensor.py:70-98,131-143 + Map.py:81-105 + cv2.getRotationMatrix2D + cv2.warpAffine
 * (bilinear, constant-0 border), restated after OpenCV imgproc/src/imgwarp.cpp (fixed-point
 * AB_BITS=10, INTER_BITS=5).  out: [submap_h * submap_w] bool. */
static void sense_occupancy_grid(const casim_config* c, const uint8_t* map, const double* h, uint8_t* out) {
  int H = c->submap_h, W = c->submap_w; /* span_x = x_width(60), span_y = y_width(60) */
  int row0, col0;
  world_to_map(c, h[CASIM_PX], h[CASIM_PY], &row0, &col0);
  /* getSubmapByIndices(center_idx_x=row0, center_idx_y=col0, span_x=W(60), span_y=H(60)) */
  int span_x = W, span_y = H;
  int sx = row0 - span_x / 2;
  if (sx < 0) sx = 0;
  int sy = col0 - span_y / 2;
  if (sy < 0) sy = 0;
  int max_x = MD - 1, max_y = MD - 1;
  int ex = sx + span_x;
  if (ex > max_x) {
    ex = max_x;
    sx = ex - span_x;
  }
  int ey = sy + span_y;
  if (ey > max_y) {
    ey = max_y;
    sy = ey - span_y;
  }
  /* angle = -heading*180/pi (python), then cv: angle *= CV_PI/180 */
  double angle_deg = -h[CASIM_HEADING] * 180 / M_PI;
  double a = angle_deg * (M_PI / 180);
  double alpha = cos(a), beta = sin(a);
  double cx = (double)(float)col0, cy = (double)(float)row0; /* Point2f center = (agent_pos[1], agent_pos[0]) */
  double M0 = alpha, M1 = beta, M2 = (1 - alpha) * cx - beta * cy;
  double M3 = -beta, M4 = alpha, M5 = beta * cx + (1 - alpha) * cy;
  /* warpAffine without WARP_INVERSE_MAP inverts M */
  double D = M0 * M4 - M1 * M3;
  D = D != 0 ? 1. / D : 0;
  double A11 = M4 * D, A22 = M0 * D;
  M0 = A11;
  M1 *= -D;
  M3 *= -D;
  M4 = A22;
  double b1 = -M0 * M2 - M1 * M5;
  double b2 = -M3 * M2 - M4 * M5;
  M2 = b1;
  M5 = b2;
  const int AB_SCALE = 1024, ROUND_DELTA = 16;
  int nrows = ex - sx, ncols = ey - sy; /* == span unless the map is smaller */
  memset(out, 0, (size_t)H * W);
  for (int oy = 0; oy < nrows && oy < W; oy++) {
    int y = sx + oy; /* dst row */
    int X0 = cv_round((M1 * y + M2) * AB_SCALE) + ROUND_DELTA;
    int Y0 = cv_round((M4 * y + M5) * AB_SCALE) + ROUND_DELTA;
    for (int ox = 0; ox < ncols && ox < H; ox++) {
      int x = sy + ox; /* dst col */
      int adelta = cv_round(M0 * x * AB_SCALE), bdelta = cv_round(M3 * x * AB_SCALE);
      int X = (X0 + adelta) >> 5, Y = (Y0 + bdelta) >> 5;
      int sxi = X >> 5, syi = Y >> 5; /* (saturate_cast<short> never saturates for a 300x300 map) */
      int fx = X & 31, fy = Y & 31;
      int v = 0;
      if (map) {
#define SRC(r, cc) (((r) >= 0 && (r) < MD && (cc) >= 0 && (cc) < MD) ? map[(r)*MD + (cc)] : 0)
        v = SRC(syi, sxi);
        if (fx > 0) v |= SRC(syi, sxi + 1);
        if (fy > 0) v |= SRC(syi + 1, sxi);
        if (fx > 0 && fy > 0) v |= SRC(syi + 1, sxi + 1);
#undef SRC
      }
      out[oy * H + ox] = v ? 1 : 0;
    }
  }
}

/* ---------------------------------------------------------------------------------------- */
/* Output bundle (all double / uint8 so that the checker sees un-rounded values)              */
typedef struct casim_oracle_obs {
  double* obs_misc;            /* [E,A,4] */
  double* other_agents_states; /* [E,A,rows,10] */
  double* laserscan;           /* [E,A,nb] */
  uint8_t* local_grid;         /* [E,A,H,W] */
  double* rewards;             /* [E,A] */
  uint8_t* done;               /* [E,A] */
  uint8_t* game_over;          /* [E] */
  uint8_t* collision_flags;    /* [E,A] */
  double* nearest_dist;        /* [E,A] */
  double* stats;               /* [8] */
} casim_oracle_obs;

static void run_sensors(const casim_config* c, int A, int n, const double* env, const uint8_t* map,
                        int e, const casim_oracle_obs* o) {
  int rows = c->max_other_agents_rows;
  for (int i = 0; i < A; i++) {
    size_t ia = (size_t)e * A + i;
    if (i >= n) continue;
    const double* s = env + i * NF;
    int sensors = (((int)s[CASIM_TYPE]) >> 16) & 0xff;
    int policy = ((int)s[CASIM_TYPE]) & 0xff;
    int nobs = 0;
    if ((sensors & CASIM_SENSOR_OTHER_AGENTS) && o->other_agents_states)
      nobs = sense_other_agents(c, env, n, i, o->other_agents_states + ia * rows * 10);
    if ((sensors & CASIM_SENSOR_LASERSCAN) && o->laserscan)
      sense_laserscan(c, map, s, o->laserscan + ia * c->laser_num_beams);
    if ((sensors & CASIM_SENSOR_OCCUPANCY_GRID) && o->local_grid)
      sense_occupancy_grid(c, map, s, o->local_grid + ia * c->submap_h * c->submap_w);
    if (o->obs_misc) {
      double* m = o->obs_misc + ia * 4;
      m[0] = s[CASIM_GX] - s[CASIM_PX]; /* rel_goal: Dynamics.py:28 (stale == fresh for finished agents) */
      m[1] = s[CASIM_GY] - s[CASIM_PY];
      m[2] = (double)nobs;
      m[3] = policy == CASIM_POLICY_LEARNING ? 1.0 : 0.0; /* config.py:208-213 use_ppo */
    }
  }
}

/* Replaces Agent.__init__ for all agents + the first _get_obs.  state: [E,A,24] doubles. */
int casim_oracle_reset(const casim_config* c, int E, int A, const int32_t* n_agents, double* state,
                       const double* init, const int32_t* type, const uint8_t* maps, int n_maps,
                       const casim_oracle_obs* o) {
  for (int e = 0; e < E; e++) {
    int n = n_agents ? n_agents[e] : A;
    double* env = state + (size_t)e * A * NF;
    for (int i = 0; i < A; i++) {
      if (i < n)
        agent_init(c, init + ((size_t)e * A + i) * 7, type[(size_t)e * A + i], env + i * NF);
      else
        memset(env + i * NF, 0, NF * sizeof(double));
    }
    const uint8_t* map = n_maps == 0 ? NULL : maps + (size_t)(n_maps == 1 ? 0 : e) * MD * MD;
    if (o) run_sensors(c, A, n, env, map, e, o);
  }
  return 0;
}

int casim_oracle_observe(const casim_config* c, int E, int A, const int32_t* n_agents, const double* state,
                         const uint8_t* maps, int n_maps, const casim_oracle_obs* o) {
  for (int e = 0; e < E; e++) {
    int n = n_agents ? n_agents[e] : A;
    const uint8_t* map = n_maps == 0 ? NULL : maps + (size_t)(n_maps == 1 ? 0 : e) * MD * MD;
    run_sensors(c, A, n, state + (size_t)e * A * NF, map, e, o);
  }
  return 0;
}

/* One CollisionAvoidanceEnv.step for each of envs [e0, e1) (collision_avoidance_env.py:162-232). */
static void step_range(const casim_config* c, int e0, int e1, int A, const int32_t* n_agents, double* state,
                       const double* ext_actions, const double* policy_actions, double dt,
                       const uint8_t* maps, int n_maps, const double* traj, int T, const double* init,
                       const int32_t* type0, const uint8_t* prev_game_over, const casim_oracle_obs* o,
                       double* stats_out) {
  double local_stats[CASIM_NUM_STATS] = {0}; /* thread-private: no false sharing between worker threads */
  double* stats = stats_out ? local_stats : NULL;
  for (int e = e0; e < e1; e++) {
    int n = n_agents ? n_agents[e] : A;
    double* env = state + (size_t)e * A * NF;
    const uint8_t* map = n_maps == 0 ? NULL : maps + (size_t)(n_maps == 1 ? 0 : e) * MD * MD;
    /* batched-env extension: auto reset at the start of the step after game_over */
    if (c->auto_reset && prev_game_over && prev_game_over[e] && init && type0) {
      for (int i = 0; i < n; i++)
        agent_init(c, init + ((size_t)e * A + i) * 7, type0[(size_t)e * A + i], env + i * NF);
    }
    /* _take_action: all actions first (:298-331), then dynamics for everyone (:334-335) */
    float act[CASIM_MAX_AGENTS][2];
    int old_flags[CASIM_MAX_AGENTS];
    for (int i = 0; i < n; i++) {
      size_t ia = (size_t)e * A + i;
      choose_action(c, env + i * NF, ext_actions ? ext_actions + ia * 2 : NULL,
                    policy_actions ? policy_actions + ia * 2 : NULL, act[i]);
      old_flags[i] = (int)env[i * NF + CASIM_FLAGS];
    }
    for (int i = 0; i < n; i++) {
      size_t ia = (size_t)e * A + i;
      take_action(c, env + i * NF, act[i], dt, traj ? traj + ia * (size_t)T * 4 : NULL, T);
    }
    /* _check_for_collisions (:630-671) */
    int coll_agent[CASIM_MAX_AGENTS] = {0}, coll_wall[CASIM_MAX_AGENTS] = {0};
    double nearest[CASIM_MAX_AGENTS];
    for (int i = 0; i < n; i++) nearest[i] = INFINITY;
    for (int i = 0; i < n; i++)
      for (int j = i + 1; j < n; j++) {
        const double *a = env + i * NF, *b = env + j * NF;
        if (is_static(b) && !c->collision_av_w_static_agent) continue; /* :643-644 */
        double dx = a[CASIM_PX] - b[CASIM_PX], dy = a[CASIM_PY] - b[CASIM_PY];
        double d = sqrt(dx * dx + dy * dy);
        double cr = a[CASIM_RADIUS] + b[CASIM_RADIUS];
        double gap = d - cr;
        if (gap < nearest[i]) nearest[i] = gap; /* :649, i only */
        if (d <= cr) {
          coll_agent[i] = 1;
          coll_agent[j] = 1;
        }
      }
    if (map) /* :656 "if self.obstacles" */
      for (int i = 0; i < n; i++)
        coll_wall[i] = collides_with_wall(c, map, env[i * NF + CASIM_PX], env[i * NF + CASIM_PY],
                                          env[i * NF + CASIM_RADIUS]);
    /* _compute_rewards (:502-567) */
    for (int i = 0; i < n; i++) {
      size_t ia = (size_t)e * A + i;
      double* s = env + i * NF;
      int flags = (int)s[CASIM_FLAGS];
      double r = c->reward_time_step;
      if (flags & CASIM_F_IS_AT_GOAL) {
        if (!(flags & CASIM_F_WAS_AT_GOAL_ALREADY)) r = c->reward_at_goal;
      } else {
        if (!(flags & CASIM_F_WAS_IN_COLLISION_ALREADY)) {
          if (coll_agent[i]) {
            r = c->reward_collision_with_agent;
            flags |= CASIM_F_IN_COLLISION;
          } else if (coll_wall[i]) {
            r = c->reward_collision_with_wall;
            flags |= CASIM_F_IN_COLLISION;
          } else {
            if (nearest[i] <= c->getting_close_range) r += -0.1 - nearest[i] / 2.;
            double wv = s[CASIM_PAST_ACT1_V] - s[CASIM_PAST_ACT0_V];
            double ww = s[CASIM_PAST_ACT1_W] - s[CASIM_PAST_ACT0_W];
            if (sqrt(wv * wv + ww * ww) > c->wiggly_behavior_threshold) r += c->reward_wiggly_behavior;
          }
        } else if (flags & CASIM_F_RAN_OUT_OF_TIME)
          r += c->reward_timeout;
        if (flags & CASIM_F_IS_INFEASIBLE) r += c->reward_infeasible;
        r += c->reward_distance_to_goal * (s[CASIM_PAST_DIST_TO_GOAL] - s[CASIM_DIST_TO_GOAL]);
      }
      if (r < c->reward_min) r = c->reward_min;
      if (r > c->reward_max) r = c->reward_max;
      r = r / (c->reward_max - c->reward_min);
      s[CASIM_FLAGS] = (double)flags;
      if (o->rewards) o->rewards[ia] = r;
      if (o->collision_flags) o->collision_flags[ia] = (uint8_t)(coll_agent[i] | (coll_wall[i] << 1));
      if (o->nearest_dist) o->nearest_dist[ia] = nearest[i];
      if (stats) stats[4] += r;
    }
    for (int i = n; i < A; i++) {
      size_t ia = (size_t)e * A + i;
      if (o->rewards) o->rewards[ia] = 0;
      if (o->collision_flags) o->collision_flags[ia] = 0;
      if (o->nearest_dist) o->nearest_dist[ia] = 0;
      if (o->done) o->done[ia] = 0;
    }
    /* _get_obs (:740-753) */
    run_sensors(c, A, n, env, map, e, o);
    /* _check_which_agents_done (:711-738) */
    int all_done = 1, all_learning_done = 1;
    for (int i = 0; i < n; i++) {
      size_t ia = (size_t)e * A + i;
      double* s = env + i * NF;
      int flags = (int)s[CASIM_FLAGS];
      int done = (flags & (CASIM_F_IS_AT_GOAL | CASIM_F_RAN_OUT_OF_TIME | CASIM_F_IN_COLLISION)) != 0;
      if (done)
        flags |= CASIM_F_IS_DONE;
      else
        flags &= ~CASIM_F_IS_DONE;
      s[CASIM_FLAGS] = (double)flags;
      if (o->done) o->done[ia] = (uint8_t)done;
      if (!done) all_done = 0;
      if (!done && (((int)s[CASIM_TYPE]) & 0xff) == CASIM_POLICY_LEARNING) all_learning_done = 0;
      if (stats) {
        int nf = flags & ~old_flags[i];
        stats[0] += (nf & CASIM_F_IN_COLLISION) != 0;
        stats[1] += (nf & CASIM_F_IS_AT_GOAL) != 0;
        stats[2] += (nf & CASIM_F_RAN_OUT_OF_TIME) != 0;
        stats[5] += !(old_flags[i] & CASIM_F_IS_DONE);
        stats[6] += 1;
      }
    }
    int go;
    if (c->game_over_mode == CASIM_GAME_OVER_ALL)
      go = all_done;
    else if (c->game_over_mode == CASIM_GAME_OVER_ALL_LEARNING)
      go = all_learning_done;
    else
      go = n > 0 ? (((int)env[CASIM_FLAGS]) & CASIM_F_IS_DONE) != 0 : 1;
    if (o->game_over) o->game_over[e] = (uint8_t)go;
    if (stats) stats[3] += go;
  }
  if (stats_out)
    for (int k = 0; k < CASIM_NUM_STATS; k++) stats_out[k] += local_stats[k];
}

typedef struct step_job {
  const casim_config* c; int e0, e1, A; const int32_t* n_agents; double* state; const double *ext, *pol;
  double dt; const uint8_t* maps; int n_maps; const double* traj; int T; const double* init;
  const int32_t* type0; const uint8_t* prev_go; const casim_oracle_obs* o; int want_stats;
  double stats[CASIM_NUM_STATS];
} step_job;

static void* step_job_main(void* p) {
  step_job* j = (step_job*)p;
  step_range(j->c, j->e0, j->e1, j->A, j->n_agents, j->state, j->ext, j->pol, j->dt, j->maps, j->n_maps,
             j->traj, j->T, j->init, j->type0, j->prev_go, j->o, j->want_stats ? j->stats : NULL);
  return NULL;
}

int casim_oracle_step(const casim_config* c, int E, int A, const int32_t* n_agents, double* state,
                      const double* ext_actions, const double* policy_actions, double dt,
                      const uint8_t* maps, int n_maps, const double* traj, int T, const double* init,
                      const int32_t* type0, const uint8_t* prev_game_over, const casim_oracle_obs* o,
                      int n_threads) {
  if (A > CASIM_MAX_AGENTS) return CASIM_EINVAL;
  double stats[CASIM_NUM_STATS] = {0};
  if (n_threads <= 1) {
    step_range(c, 0, E, A, n_agents, state, ext_actions, policy_actions, dt, maps, n_maps, traj, T, init,
               type0, prev_game_over, o, o->stats ? stats : NULL);
  } else {
    /* environments are independent: static partition over POSIX threads (CPU baseline only) */
    if (n_threads > 64) n_threads = 64;
    step_job jobs[64];
    pthread_t th[64];
    memset(jobs, 0, sizeof(jobs));
    for (int t = 0; t < n_threads; t++) {
      step_job* j = &jobs[t];
      j->c = c; j->A = A; j->n_agents = n_agents; j->state = state; j->ext = ext_actions;
      j->pol = policy_actions; j->dt = dt; j->maps = maps; j->n_maps = n_maps; j->traj = traj; j->T = T;
      j->init = init; j->type0 = type0; j->prev_go = prev_game_over; j->o = o; j->want_stats = o->stats != NULL;
      j->e0 = (int)((long long)E * t / n_threads);
      j->e1 = (int)((long long)E * (t + 1) / n_threads);
      pthread_create(&th[t], NULL, step_job_main, j);
    }
    for (int t = 0; t < n_threads; t++) {
      pthread_join(th[t], NULL);
      for (int k = 0; k < CASIM_NUM_STATS; k++) stats[k] += jobs[t].stats[k];
    }
  }
  if (o->stats) memcpy(o->stats, stats, sizeof(stats));
  return 0;
}
