/*
 * casim_oracle_explore.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see casim_oracle.c).
 *
 * Plain-C fp64 restatement of the exploration path of mlodel/gym-exploration-2d
 * (gym_collision_avoidance/envs/...):
 *   information_models/edfMap.py:11-47      EDF, get_edf_value_from_pose, checkVisibility
 *   information_models/targetMap.py:26-147  cell <-> pose, getVisibleCells, update, MI reward
 *   policies/ig_greedy.py:31-93             detector emulation, greedy one-step action
 *
 * Parity status: PINNED against the reference classes themselves executed in the build container
 * (oracle/gen_golden_explore.py -> tests/golden/explore.npz, tests/test_oracle_golden.py), with one
 * documented deviation: ig_greedy.find_next_action passes a bare pose to targetMap.update
 * (ig_greedy.py:39), which raises once a target is detected; like ig_mcts.update_belief
 * (ig_mcts.py:117-133) this file implements update([pose], [targets]).
 * Indices that would leave the arrays (Python would wrap negative indices or raise) are clamped and
 * counted in counters[3].
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/casim.h"

#define NF CASIM_STATE_WORDS
#define MD CASIM_MAP_DIM

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* edfMap.py:11-12: scipy.ndimage.distance_transform_edt((~map).astype(int)) * cellSize.
 * Exact EDT kept as integer squared distances (cells^2): column scan, then exact row minimisation. */
void casim_oracle_edt(const uint8_t* map, int32_t* d2) {
  const int BIG = 1 << 20;
  int* g = (int*)malloc(sizeof(int) * MD * MD);
  for (int col = 0; col < MD; col++) {
    int d = BIG;
    for (int row = 0; row < MD; row++) {
      d = map[row * MD + col] ? 0 : (d + 1 < BIG ? d + 1 : BIG);
      g[row * MD + col] = d;
    }
    d = BIG;
    for (int row = MD - 1; row >= 0; row--) {
      d = map[row * MD + col] ? 0 : (d + 1 < BIG ? d + 1 : BIG);
      if (d < g[row * MD + col]) g[row * MD + col] = d;
    }
  }
  for (int row = 0; row < MD; row++)
    for (int col = 0; col < MD; col++) {
      long long best = -1;
      for (int k = 0; k < MD; k++) {
        int gv = g[row * MD + k];
        if (gv >= BIG) continue;
        long long dx = col - k, v = dx * dx + (long long)gv * gv;
        if (best < 0 || v < best) best = v;
      }
      d2[row * MD + col] = best < 0 ? (1 << 30) : (int32_t)best;
    }
  free(g);
}

typedef struct {
  const casim_config* c;
  const int32_t* d2; /* [300][300] */
  unsigned long long* cnt; /* cells tested, EDF samples, visible cells, clamped indices */
} xctx;

/* edfMap.map[yIdc, xIdc] for a world point (edfMap.py:14-19, 38-40): NO y flip (sic) */
static double edf_at(const xctx* x, double px, double py) {
  const casim_config* c = x->c;
  double fx = floor((px + c->map_width / 2) / c->map_cell), fy = floor((py + c->map_height / 2) / c->map_cell);
  int xi, yi, clamped = 0;
  if (!(fx >= 0)) { fx = 0; clamped = 1; }
  if (fx > MD - 1) { fx = MD - 1; clamped = 1; }
  if (!(fy >= 0)) { fy = 0; clamped = 1; }
  if (fy > MD - 1) { fy = MD - 1; clamped = 1; }
  xi = (int)fx; yi = (int)fy;
  if (clamped && x->cnt) x->cnt[3]++;
  if (x->cnt) x->cnt[1]++;
  return sqrt((double)x->d2[yi * MD + xi]) * c->map_cell;
}

/* edfMap.py:21-47 */
static int check_visibility(const xctx* x, double px, double py, double gx, double gy) {
  const double distIncr = 0.05, thres = 0.001;
  double dx = gx - px, dy = gy - py;
  double L = sqrt(dx * dx + dy * dy);
  double u = distIncr / L;
  while (u < 1) {
    double nx = (1 - u) * px + u * gx, ny = (1 - u) * py + u * gy;
    double minDist = edf_at(x, nx, ny);
    if (minDist < thres) return 0;
    u += minDist / L;
  }
  return 1;
}

/* targetMap.py:26-31 getCellsFromPose (one axis) */
static int cell_of(const casim_config* c, double v, double size) { return (int)floor((v + size / 2) / c->ig_cell); }
/* targetMap.py:33-36 getPoseFromCell (one axis) */
static double pose_of(const casim_config* c, int i, double size) { return i * c->ig_cell - size / 2 + c->ig_cell / 2; }
/* targetMap.py:38-41 get_pos_in_map_lims (one axis) */
static double lim(double v, double size) {
  double a = v < size / 2 ? v : size / 2;
  return a > -size / 2 ? a : -size / 2;
}

/* targetMap.py:43-84 getVisibleCells.  vis: [dim*dim] uint8 mask indexed [j*dim + i] (j = y cell, i = x cell),
 * cleared first.  Returns the number of visible cells. */
static int visible_cells(const xctx* x, const double pose[3], uint8_t* vis) {
  const casim_config* c = x->c;
  const int dim = c->ig_map_dim;
  const double W = c->map_width, H = c->map_height, R = c->ig_range, F = c->ig_sens_fov;
  memset(vis, 0, (size_t)dim * dim);
  double phi = pose[2];
  double pts[4][2];
  pts[0][0] = pose[0]; pts[0][1] = pose[1];
  pts[1][0] = lim(pose[0] + R * cos(phi), W);     pts[1][1] = lim(pose[1] + R * sin(phi), H);
  pts[2][0] = lim(pose[0] + R * cos(phi + F), W); pts[2][1] = lim(pose[1] + R * sin(phi + F), H);
  pts[3][0] = lim(pose[0] + R * cos(phi - F), W); pts[3][1] = lim(pose[1] + R * sin(phi - F), H);
  int x0 = 1 << 30, x1 = -(1 << 30), y0 = 1 << 30, y1 = -(1 << 30);
  for (int k = 0; k < 4; k++) {
    int cx = cell_of(c, pts[k][0], W), cy = cell_of(c, pts[k][1], H);
    if (cx < x0) x0 = cx;
    if (cx > x1) x1 = cx;
    if (cy < y0) y0 = cy;
    if (cy > y1) y1 = cy;
  }
  /* Python would wrap negative cell indices / raise beyond the edge: clamp and count */
  if (x0 < 0) { x0 = 0; if (x->cnt) x->cnt[3]++; }
  if (y0 < 0) { y0 = 0; if (x->cnt) x->cnt[3]++; }
  if (x1 > dim) { x1 = dim; if (x->cnt) x->cnt[3]++; }
  if (y1 > dim) { y1 = dim; if (x->cnt) x->cnt[3]++; }
  double cs = cos(phi), sn = sin(phi);
  int count = 0;
  for (int i = x0; i < x1; i++)       /* range(xIdcStart, xIdcEnd): end exclusive (:73-74) */
    for (int j = y0; j < y1; j++) {
      double cxp = pose_of(c, i, W), cyp = pose_of(c, j, H);
      double dx = cxp - pose[0], dy = cyp - pose[1];
      double r0 = cs * dx + sn * dy, r1 = -sn * dx + cs * dy; /* R = ((c, s), (-s, c)) */
      double dphi = atan2(r1, r0);
      double rn = sqrt(r0 * r0 + r1 * r1);
      if (x->cnt) x->cnt[0]++;
      if (rn < R && fabs(dphi) < F / 2) {
        if (check_visibility(x, pose[0], pose[1], cxp, cyp)) {
          vis[j * dim + i] = 1;
          count++;
        }
      }
    }
  if (x->cnt) x->cnt[2] += (unsigned long long)count;
  return count;
}

/* targetMap.py:86-128 update([pose], [targets], frame='global'); belief [dim][dim] odds, RMW */
static void belief_update(const xctx* x, const double pose[3], const double* targets, int nt, double* belief,
                          uint8_t* vis) {
  const casim_config* c = x->c;
  const int dim = c->ig_map_dim;
  double cs = cos(pose[2]), sn = sin(pose[2]);
  double det[CASIM_MAX_AGENTS][2];
  for (int t = 0; t < nt; t++) {
    double dx = targets[2 * t] - pose[0], dy = targets[2 * t + 1] - pose[1];
    det[t][0] = cs * dx + sn * dy;
    det[t][1] = -sn * dx + cs * dy;
  }
  visible_cells(x, pose, vis);
  double thr = sqrt(0.5) * c->ig_cell + c->ig_tolerance;
  for (int j = 0; j < dim; j++)
    for (int i = 0; i < dim; i++) {
      if (!vis[j * dim + i]) continue;
      double rs = c->ig_remp;
      if (nt > 0) {
        double dx = pose_of(c, i, c->map_width) - pose[0], dy = pose_of(c, j, c->map_height) - pose[1];
        double r0 = cs * dx + sn * dy, r1 = -sn * dx + cs * dy;
        for (int t = 0; t < nt; t++) {
          double ex = det[t][0] - r0, ey = det[t][1] - r1;
          if (sqrt(ex * ex + ey * ey) < thr) {
            rs = c->ig_rocc;
            break;
          }
        }
      }
      belief[j * dim + i] *= rs;
    }
}

/* targetMap.py:130-143 get_reward_from_cells, one cell */
static double cell_mi(const casim_config* c, double r) {
  double p = r / (r + 1);
  double f_p = log((r + 1) / (r + (1 / c->ig_rocc))) - log(c->ig_rocc) / (r * c->ig_rocc + 1);
  double f_n = log((r + 1) / (r + (1 / c->ig_remp))) - log(c->ig_remp) / (r * c->ig_remp + 1);
  double P_p = p * (1 - c->ig_p_false_neg) + (1 - p) * c->ig_p_false_pos;
  double P_n = p * c->ig_p_false_neg + (1 - p) * (1 - c->ig_p_false_pos);
  return P_p * f_p + P_n * f_n;
}

/* targetMap.py:145-147 get_reward_from_pose (sum in row-major cell order; the reference sums in Python-set
 * order, so agreement is to rounding) */
static double reward_from_pose(const xctx* x, const double pose[3], const double* belief, uint8_t* vis) {
  const int dim = x->c->ig_map_dim;
  visible_cells(x, pose, vis);
  double s = 0;
  for (int k = 0; k < dim * dim; k++)
    if (vis[k]) s += cell_mi(x->c, belief[k]);
  return s;
}

/* ig_greedy.py:64-93 greedy_action + get_next_pose.  action[2]; cand[9] = MI of each candidate (-1 infeasible).
 * Returns the best MI (-1 if every candidate is infeasible; the action is then (-1, -1): the integer -1 of
 * ig_greedy.py:68 broadcast into all_actions[i, :], collision_avoidance_env.py:320). */
static double greedy_action(const xctx* x, const double pose[3], double radius, const double* belief, uint8_t* vis,
                            double* action, double* cand) {
  const casim_config* c = x->c;
  const double vel_list[3] = {0.0, 2.0, 4.0}, dphi_list[3] = {-M_PI, 0, M_PI};
  double max_mi = -1;
  action[0] = -1;
  action[1] = -1;
  for (int a = 0; a < 9; a++) {
    double v = vel_list[a / 3], w = dphi_list[a % 3];
    double cs = cos(pose[2]), sn = sin(pose[2]);
    double vx = cs * v + (-sn) * 0.0, vy = sn * v + cs * 0.0; /* np.dot(R, [v, 0]) */
    double np_[3] = {pose[0] + vx * c->ig_dt, pose[1] + vy * c->ig_dt, pose[2] + w * c->ig_dt};
    cand[a] = -1;
    if (edf_at(x, np_[0], np_[1]) > radius + 0.1) {
      double mi = reward_from_pose(x, np_, belief, vis);
      cand[a] = mi;
      if (mi > max_mi) {
        max_mi = mi;
        action[0] = v;
        action[1] = w;
      }
    }
  }
  return max_mi;
}

/* ig_greedy.py:45-62 find_targets_in_obs: rows [rows][10] of other_agents_states; pose = (px, py, heading).
 * detect_fov is compared in DEGREES against an angle in radians (sic, :55-56). */
static int find_targets(const casim_config* c, const double* rows, int nrows, const double pose[3], double* targets) {
  int nt = 0;
  for (int k = 0; k < nrows; k++) {
    const double* r = rows + 10 * k;
    if (r[9] == 1.0) {
      double dphi = atan2(r[1], r[0]) - pose[2];
      int in_fov = fabs(dphi) <= c->ig_detect_fov / 2.0;
      double rn = sqrt(r[0] * r[0] + r[1] * r[1]);
      if (in_fov && rn <= c->ig_range && nt < CASIM_MAX_AGENTS) {
        targets[2 * nt] = r[0] + pose[0];
        targets[2 * nt + 1] = r[1] + pose[1];
        nt++;
      }
    }
  }
  return nt;
}

/* ig_greedy.find_next_action (ig_greedy.py:31-43) for every ig_greedy agent that is not done
 * (collision_avoidance_env.py:298-300).  All arrays are doubles except maps / vis_out (uint8) and
 * counters (uint64[4]).  With auto_reset and prev_game_over[e] set, the agent acts from its initial
 * condition with a fresh prior belief (batched-env extension, see include/casim.h). */
typedef struct {
  const casim_config* c; int e0, e1, A; const int32_t* n_agents; const double* state; const double* oas;
  double* belief; const uint8_t* maps; int n_maps; const int32_t* d2_shared; double* policy_actions; double* best_mi;
  double* cand_mi; uint8_t* vis_out; const double* init; const uint8_t* prev_game_over;
  unsigned long long cnt[4];
} xjob;

static void* explore_range(void* pj) {
  xjob* J = (xjob*)pj;
  const casim_config* c = J->c;
  const int A = J->A, dim = c->ig_map_dim, rows = c->max_other_agents_rows;
  const double *state = J->state, *oas = J->oas, *init = J->init;
  double *belief = J->belief, *policy_actions = J->policy_actions, *best_mi = J->best_mi, *cand_mi = J->cand_mi;
  uint8_t* vis_out = J->vis_out;
  int32_t* d2_own = J->d2_shared ? NULL : (int32_t*)malloc(sizeof(int32_t) * MD * MD);
  uint8_t* vis = (uint8_t*)malloc((size_t)dim * dim);
  uint8_t* vis2 = (uint8_t*)malloc((size_t)dim * dim);
  memset(J->cnt, 0, sizeof(J->cnt));
  for (int e = J->e0; e < J->e1; e++) {
    if (d2_own) casim_oracle_edt(J->maps + (size_t)e * MD * MD, d2_own);
    xctx x = {c, d2_own ? d2_own : J->d2_shared, J->cnt};
    int n = J->n_agents ? J->n_agents[e] : A;
    int restart = c->auto_reset && J->prev_game_over && init && J->prev_game_over[e];
    for (int i = 0; i < n; i++) {
      size_t g = (size_t)e * A + i;
      const double* s = state + g * NF;
      int type = (int)s[CASIM_TYPE];
      if ((type & 0xff) != CASIM_POLICY_IG_GREEDY) continue;
      double* b = belief + g * dim * dim;
      double* act = policy_actions + 2 * g;
      double cand[9];
      double pose[3], radius, targets[2 * CASIM_MAX_AGENTS];
      int nt = 0;
      if (restart) {
        const double* in = init + g * 7;
        pose[0] = in[0]; pose[1] = in[1];
        pose[2] = isnan(in[6]) ? atan2(in[3] - in[1], in[2] - in[0]) : in[6];
        radius = in[4];
        for (int k = 0; k < dim * dim; k++) b[k] = c->ig_prior;
        for (int j = 0; j < n; j++) {
          if (j == i) continue;
          if (((int)state[((size_t)e * A + j) * NF + CASIM_TYPE] & 0xff) != CASIM_POLICY_STATIC) continue;
          const double* oin = init + ((size_t)e * A + j) * 7;
          double row[10] = {oin[0] - in[0], oin[1] - in[1], 0, 0, 0, 0, 0, 0, 0, 1.0};
          nt += find_targets(c, row, 1, pose, targets + 2 * nt);
        }
      } else {
        if ((int)s[CASIM_FLAGS] & CASIM_F_IS_DONE) {
          act[0] = 0; act[1] = 0;
          if (best_mi) best_mi[g] = -1;
          if (cand_mi) for (int k = 0; k < 9; k++) cand_mi[9 * g + k] = -1;
          if (vis_out) memset(vis_out + g * dim * dim, 0, (size_t)dim * dim);
          continue;
        }
        pose[0] = s[CASIM_PX]; pose[1] = s[CASIM_PY]; pose[2] = s[CASIM_HEADING];
        radius = s[CASIM_RADIUS];
        nt = find_targets(c, oas + g * rows * 10, rows, pose, targets);
      }
      belief_update(&x, pose, targets, nt, b, vis);
      if (vis_out) memcpy(vis_out + g * dim * dim, vis, (size_t)dim * dim);
      double m = greedy_action(&x, pose, radius, b, vis2, act, cand);
      if (best_mi) best_mi[g] = m;
      if (cand_mi) memcpy(cand_mi + 9 * g, cand, sizeof(cand));
    }
  }
  free(d2_own); free(vis); free(vis2);
  return NULL;
}

int casim_oracle_explore_step(const casim_config* c, int E, int A, const int32_t* n_agents, const double* state,
                              const double* oas, double* belief, const uint8_t* maps, int n_maps,
                              double* policy_actions, double* best_mi, double* cand_mi,
                              unsigned long long* counters, uint8_t* vis_out, const double* init,
                              const uint8_t* prev_game_over, int n_threads) {
  if (n_maps <= 0 || !maps) return CASIM_EINVAL;
  /* the EDF is built once per map (edfMap.__init__), not per step: keep the last shared map's transform */
  static uint8_t* cache_map = NULL;
  static int32_t* cache_d2 = NULL;
  const int32_t* d2_shared = NULL;
  if (n_maps == 1) {
    if (!cache_map) {
      cache_map = (uint8_t*)malloc(MD * MD);
      cache_d2 = (int32_t*)malloc(sizeof(int32_t) * MD * MD);
      memset(cache_map, 0xff, MD * MD);
    }
    if (memcmp(cache_map, maps, MD * MD) != 0) {
      casim_oracle_edt(maps, cache_d2);
      memcpy(cache_map, maps, MD * MD);
    }
    d2_shared = cache_d2;
  }
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  xjob jobs[64];
  pthread_t th[64];
  for (int t = 0; t < n_threads; t++) {
    xjob* j = &jobs[t];
    j->c = c; j->A = A; j->n_agents = n_agents; j->state = state; j->oas = oas; j->belief = belief; j->maps = maps;
    j->n_maps = n_maps; j->d2_shared = d2_shared; j->policy_actions = policy_actions; j->best_mi = best_mi;
    j->cand_mi = cand_mi; j->vis_out = vis_out; j->init = init; j->prev_game_over = prev_game_over;
    j->e0 = (int)((long long)E * t / n_threads);
    j->e1 = (int)((long long)E * (t + 1) / n_threads);
    if (n_threads == 1) explore_range(j); else pthread_create(&th[t], NULL, explore_range, j);
  }
  if (n_threads > 1) for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
  if (counters) {
    memset(counters, 0, 4 * sizeof(unsigned long long));
    for (int t = 0; t < n_threads; t++) for (int k = 0; k < 4; k++) counters[k] += jobs[t].cnt[k];
  }
  return 0;
}

/* Standalone primitives for the golden tests (targetMap.getVisibleCells / get_reward_from_pose,
 * edfMap.checkVisibility) on one map. */
int casim_oracle_visible_cells(const casim_config* c, const uint8_t* map, const double* pose, uint8_t* vis) {
  int32_t* d2 = (int32_t*)malloc(sizeof(int32_t) * MD * MD);
  casim_oracle_edt(map, d2);
  xctx x = {c, d2, NULL};
  int n = visible_cells(&x, pose, vis);
  free(d2);
  return n;
}

double casim_oracle_reward_from_pose(const casim_config* c, const uint8_t* map, const double* pose,
                                     const double* belief) {
  int32_t* d2 = (int32_t*)malloc(sizeof(int32_t) * MD * MD);
  uint8_t* vis = (uint8_t*)malloc((size_t)c->ig_map_dim * c->ig_map_dim);
  casim_oracle_edt(map, d2);
  xctx x = {c, d2, NULL};
  double r = reward_from_pose(&x, pose, belief, vis);
  free(d2); free(vis);
  return r;
}

/* Dec-MCTS simulation rollouts: policies/ig_mcts.py:173-198 get_next_pose, :223-244 mcts_sim_state_storer,
 * :246-253 mcts_reward, driven the way pydecmcts/DecMCTS.py:233-269 _simulate drives them.  One agent:
 * K action sequences of H action indices (index = 3 * vel_idx + dphi_idx into vel_list (0, 2, 4) x dphi_list
 * (-pi/2, 0, pi/2), ig_mcts.py:259-265; 255 = the sequence has ended).  A tree node's state followed by a
 * rollout equals the whole root-to-leaf sequence replayed here, because the tree only keeps feasible steps
 * (mcts_state_storer, :200-221) and a feasible step does the same thing in both storers.
 * An infeasible step leaves the pose where it was and observes nothing (:237-243).  The cells seen from the
 * ROOT pose are not part of obsvd_cells (:226-228).  exclude: [dim*dim] cells claimed by the other robots'
 * paths (mcts_reward's other_agents_obsvd_cells) or NULL.  rewards [K]; end_pose [K][3] and obsvd
 * [K][dim*dim] (the state's obsvd_cells, before the difference) are optional.  The sum runs in row-major cell order (the reference: Python-set order). */
int casim_oracle_explore_rollouts(const casim_config* c, const uint8_t* map, const double* pose, double radius,
                                  const double* belief, const uint8_t* actions, int K, int H, int xdt,
                                  const uint8_t* exclude, double* rewards, double* end_pose, uint8_t* obsvd,
                                  int32_t* n_feasible) {
  const int dim = c->ig_map_dim;
  int32_t* d2 = (int32_t*)malloc(sizeof(int32_t) * MD * MD);
  uint8_t* vis = (uint8_t*)malloc((size_t)dim * dim);
  uint8_t* seen = (uint8_t*)malloc((size_t)dim * dim);
  casim_oracle_edt(map, d2);
  xctx x = {c, d2, NULL};
  const double vel_list[3] = {0.0, 2.0, 4.0}, dphi_list[3] = {-0.5 * M_PI, 0, 0.5 * M_PI};
  for (int k = 0; k < K; k++) {
    double cur[3] = {pose[0], pose[1], pose[2]};
    int applied = 0;
    memset(seen, 0, (size_t)dim * dim);
    for (int t = 0; t < H; t++) {
      int a = actions[(size_t)k * H + t];
      if (a > 8) break;
      double v = vel_list[a / 3], w = dphi_list[a % 3];
      double nx[3] = {cur[0], cur[1], cur[2]};
      int ok = 1;
      for (int i = 0; i < xdt && ok; i++) {
        double cs = cos(nx[2]), sn = sin(nx[2]);
        double vx = cs * v + (-sn) * 0.0, vy = sn * v + cs * 0.0;
        nx[0] = nx[0] + vx * c->ig_dt;
        nx[1] = nx[1] + vy * c->ig_dt;
        nx[2] = nx[2] + w * c->ig_dt;
        if (v == 0.0) continue;
        int in_map = c->map_width / 2 > nx[0] && c->map_height / 2 > nx[1] && nx[0] > -c->map_width / 2 &&
                     nx[1] > -c->map_height / 2;
        if (!in_map || !(edf_at(&x, nx[0], nx[1]) > radius + 0.1)) ok = 0;
      }
      if (!ok) continue; /* pose kept, nothing observed */
      cur[0] = nx[0]; cur[1] = nx[1]; cur[2] = nx[2];
      applied++;
      visible_cells(&x, cur, vis);
      for (int q = 0; q < dim * dim; q++) seen[q] |= vis[q];
    }
    double s = 0;
    for (int q = 0; q < dim * dim; q++) {
      if (seen[q] && !(exclude && exclude[q])) s += cell_mi(c, belief[q]);
    }
    rewards[k] = s;
    if (n_feasible) n_feasible[k] = applied;
    if (end_pose) { end_pose[3 * k] = cur[0]; end_pose[3 * k + 1] = cur[1]; end_pose[3 * k + 2] = cur[2]; }
    if (obsvd) memcpy(obsvd + (size_t)k * dim * dim, seen, (size_t)dim * dim);
  }
  free(d2); free(vis); free(seen);
  return 0;
}

/* ig_mcts.update_belief (policies/ig_mcts.py:117-137) for every explorer agent of a batch, in the order
 * collision_avoidance_env.py:341-368 calls it (ascending agent index, done agents skipped, :298-299):
 *   poses   = [own pose] + [pose of every OTHER explorer j (ascending j), done or not]
 *   targets = [own detections (find_targets_in_obs, :139-156)] + [agents[j].policy.obsvd_targets]
 * where obsvd_targets of j is what j detected at ITS latest update: this step's if j < i and j is not done, else
 * the previous call's (targets_prev / n_prev; none before j's first update, :127).  targetMap.update multiplies
 * every visible cell once per pose that sees it (targetMap.py:86-128); team_obsv_cells is the union and team_reward
 * the mutual information of the union on the UPDATED map (:136-137).
 * state [E][A][NF], oas [E][A][rows][10], belief [E][A][dim][dim] RMW, one shared map.
 * targets_prev/new [E][A][CASIM_MAX_AGENTS][2] global-frame doubles with counts n_prev/new [E][A] (a done agent
 * keeps its previous list).  team_mask [E][A][dim*dim] uint8 and team_reward [E][A] are optional. */
int casim_oracle_team_update(const casim_config* c, int E, int A, const int32_t* n_agents, const double* state,
                             const double* oas, double* belief, const uint8_t* map, const double* targets_prev,
                             const int32_t* n_prev, double* targets_new, int32_t* n_new, uint8_t* team_mask,
                             double* team_reward) {
  const int dim = c->ig_map_dim, rows = c->max_other_agents_rows;
  const int TM = CASIM_MAX_AGENTS;
  int32_t* d2 = (int32_t*)malloc(sizeof(int32_t) * MD * MD);
  uint8_t* vis = (uint8_t*)malloc((size_t)dim * dim);
  uint8_t* uni = (uint8_t*)malloc((size_t)dim * dim);
  casim_oracle_edt(map, d2);
  xctx x = {c, d2, NULL};
  memcpy(targets_new, targets_prev, sizeof(double) * (size_t)E * A * TM * 2);
  memcpy(n_new, n_prev, sizeof(int32_t) * (size_t)E * A);
  for (int e = 0; e < E; e++) {
    int n = n_agents ? n_agents[e] : A;
    for (int i = 0; i < n; i++) {
      size_t g = (size_t)e * A + i;
      const double* s = state + g * NF;
      if (((int)s[CASIM_TYPE] & 0xff) != CASIM_POLICY_IG_GREEDY) continue;
      if (team_reward) team_reward[g] = 0;
      if (team_mask) memset(team_mask + g * dim * dim, 0, (size_t)dim * dim);
      if ((int)s[CASIM_FLAGS] & CASIM_F_IS_DONE) continue;
      double* b = belief + g * dim * dim;
      double pose[3] = {s[CASIM_PX], s[CASIM_PY], s[CASIM_HEADING]};
      /* own detections become this agent's obsvd_targets: later agents of this step read them from targets_new */
      n_new[g] = find_targets(c, oas + g * rows * 10, rows, pose, targets_new + g * TM * 2);
      memset(uni, 0, (size_t)dim * dim);
      belief_update(&x, pose, targets_new + g * TM * 2, n_new[g], b, vis);
      for (int q = 0; q < dim * dim; q++) uni[q] |= vis[q];
      for (int j = 0; j < n; j++) {
        size_t gj = (size_t)e * A + j;
        const double* sj = state + gj * NF;
        if (j == i || ((int)sj[CASIM_TYPE] & 0xff) != CASIM_POLICY_IG_GREEDY) continue;
        double pj[3] = {sj[CASIM_PX], sj[CASIM_PY], sj[CASIM_HEADING]};
        /* j < i and not done: already replaced by this step's list above; otherwise still the previous one */
        belief_update(&x, pj, targets_new + gj * TM * 2, n_new[gj], b, vis);
        for (int q = 0; q < dim * dim; q++) uni[q] |= vis[q];
      }
      if (team_mask) memcpy(team_mask + g * dim * dim, uni, (size_t)dim * dim);
      if (team_reward) {
        double r = 0;
        for (int q = 0; q < dim * dim; q++)
          if (uni[q]) r += cell_mi(c, b[q]);
        team_reward[g] = r;
      }
    }
  }
  free(d2); free(vis); free(uni);
  return 0;
}
