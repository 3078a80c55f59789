/*
 * TEST INFRASTRUCTURE -- CPU restatement of the reference's seeded training-scenario generators
 *   gym_collision_avoidance/envs/test_cases.py:1192-1281  train_agents_swap_circle(seed=s)
 *   gym_collision_avoidance/envs/test_cases.py:1365-1463  train_agents_random_positions(seed=s)
 *   gym_collision_avoidance/envs/test_cases.py:1283-1363  train_agents_pairwise_swap(seed=s)
 *   gym_collision_avoidance/envs/test_cases.py:129-133    is_pose_valid
 * for a batch of environments, one seed per environment.  The reference draws from numpy's legacy global generator
 * (np.random.seed(s); np.random.uniform) and from Python's `random` (random.seed(s); random.choice): both are the
 * published MT19937 of Matsumoto & Nishimura (init_genrand / init_by_array / genrand_int32, 53-bit doubles from two
 * outputs), restated here, so that a seed yields the reference's scenario itself, not merely one like it.
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use this file.  Pinned by
 * tests/golden/scenarios.npz (oracle/gen_golden_scenarios.py runs the reference's functions).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "../include/casim.h"

typedef struct { uint32_t mt[624]; int idx; } mt_t;

static void mt_init_genrand(mt_t* m, uint32_t s) { /* numpy legacy seed(int): _legacy_seeding -> mt19937_seed */
  m->mt[0] = s;
  for (int i = 1; i < 624; i++) m->mt[i] = 1812433253u * (m->mt[i - 1] ^ (m->mt[i - 1] >> 30)) + (uint32_t)i;
  m->idx = 624;
}

static void mt_init_by_array(mt_t* m, const uint32_t* key, int len) { /* CPython random.seed(int): key = |int| words */
  mt_init_genrand(m, 19650218u);
  int i = 1, j = 0;
  for (int k = (624 > len ? 624 : len); k; k--) {
    m->mt[i] = (m->mt[i] ^ ((m->mt[i - 1] ^ (m->mt[i - 1] >> 30)) * 1664525u)) + key[j] + (uint32_t)j;
    i++; j++;
    if (i >= 624) { m->mt[0] = m->mt[623]; i = 1; }
    if (j >= len) j = 0;
  }
  for (int k = 623; k; k--) {
    m->mt[i] = (m->mt[i] ^ ((m->mt[i - 1] ^ (m->mt[i - 1] >> 30)) * 1566083941u)) - (uint32_t)i;
    i++;
    if (i >= 624) { m->mt[0] = m->mt[623]; i = 1; }
  }
  m->mt[0] = 0x80000000u;
  m->idx = 624;
}

static uint32_t mt_next(mt_t* m) {
  if (m->idx >= 624) {
    for (int k = 0; k < 624; k++) {
      uint32_t y = (m->mt[k] & 0x80000000u) | (m->mt[(k + 1) % 624] & 0x7fffffffu);
      m->mt[k] = m->mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    m->idx = 0;
  }
  uint32_t y = m->mt[m->idx++];
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

/* numpy legacy random_sample / uniform(low, high) = low + (high - low) * sample */
static double mt_uniform(mt_t* m, double low, double high) {
  uint32_t a = mt_next(m) >> 5, b = mt_next(m) >> 6;
  double u = (a * 67108864.0 + b) / 9007199254740992.0;
  return low + (high - low) * u;
}

/* CPython random.choice over a 2-element list: _randbelow_with_getrandbits(2): getrandbits(2) until < 2 */
static int py_choice2(mt_t* m) {
  uint32_t r;
  do r = mt_next(m) >> 30; while (r >= 2);
  return (int)r;
}

/* CPython Random._randbelow_with_getrandbits(n): getrandbits(n.bit_length()) until < n */
static int py_randbelow(mt_t* m, int n) {
  int k = 0;
  while ((n >> k) != 0) k++;
  uint32_t r;
  do r = mt_next(m) >> (32 - k); while (r >= (uint32_t)n);
  return (int)r;
}

/* test_cases.py:129-133 */
static int pose_valid(double x, double y, const double (*list)[2], int n, double distance) {
  for (int k = 0; k < n; k++) {
    double dx = x - list[k][0], dy = y - list[k][1];
    if (sqrt(dx * dx + dy * dy) < distance) return 0;
  }
  return 1;
}

#define SC_MAX CASIM_MAX_AGENTS

static int32_t type_word(int policy, int dyn, int sensors) { return policy | (dyn << 8) | (sensors << 16); }

static void put_agent(double* init, int32_t* types, double* coop, int slot, const double* start, const double* goal,
                      int policy, int dyn, int sensors, double c) {
  double* r = init + 7 * slot;
  r[0] = start[0]; r[1] = start[1]; r[2] = goal[0]; r[3] = goal[1];
  r[4] = 0.5;      /* radius, test_cases.py:1195 / :1368 */
  r[5] = 1.0;      /* pref_speed */
  r[6] = NAN;      /* initial_heading None: towards the goal (agent.py:21-33) */
  types[slot] = type_word(policy, dyn, sensors);
  if (coop) coop[slot] = c;
}

/* kind: CASIM_SCENARIO_SWAP_CIRCLE (0), CASIM_SCENARIO_RANDOM_POSITIONS (1) or CASIM_SCENARIO_PAIRWISE_SWAP (2).  seeds [E] (non-zero: `if seed:`).
 * init [E][A][7] doubles, types [E][A], n_agents [E], coop [E][A] (optional; cooperation_coef, agent.py).  Slots past
 * the generated agents are zeroed.  Returns 0, or CASIM_EINVAL if 2 * (max(number_of_agents, 2) / 2) > A. */
int casim_oracle_scenario(int kind, int E, int A, int number_of_agents, const uint32_t* seeds, int ego_policy,
                          int ego_dynamics, int other_dynamics, double* init, int32_t* types, int32_t* n_agents,
                          double* coop) {
  const int n = number_of_agents > 2 ? number_of_agents : 2;
  const int P = n / 2;
  if (2 * P > A || n > SC_MAX || (kind < 0 || kind > 2)) return CASIM_EINVAL;
  const int ga3c = ego_policy == CASIM_POLICY_GA3C_CADRL;
  const int ego_dyn = ga3c ? CASIM_DYN_UNICYCLE_MAX_ACC : ego_dynamics; /* :1254-1257, :1435-1438 */
  const int OAS = CASIM_SENSOR_OTHER_AGENTS, GRID = CASIM_SENSOR_OCCUPANCY_GRID;
  for (int e = 0; e < E; e++) {
    double* in = init + (size_t)e * A * 7;
    int32_t* ty = types + (size_t)e * A;
    double* co = coop ? coop + (size_t)e * A : NULL;
    memset(in, 0, sizeof(double) * A * 7);
    memset(ty, 0, sizeof(int32_t) * A);
    if (co) memset(co, 0, sizeof(double) * A);
    n_agents[e] = 2 * P;
    mt_t m;
    mt_init_genrand(&m, seeds[e]); /* np.random.seed(seed) */
    if (kind == 0) {
      double pos[2 * SC_MAX][2];
      int np_ = 0;
      double d = mt_uniform(&m, 4.0, 8.0), a = mt_uniform(&m, -M_PI, M_PI);
      double x = d * cos(a), y = d * sin(a);
      pos[np_][0] = -x; pos[np_][1] = -y; np_++;
      pos[np_][0] = x; pos[np_][1] = y; np_++;
      for (int ag = 0; ag < P - 1; ag++) {
        int ok = 0;
        while (!ok) {
          d = mt_uniform(&m, 4.0, 8.0); a = mt_uniform(&m, -M_PI, M_PI);
          x = d * cos(a); y = d * sin(a);
          ok = pose_valid(-x, -y, pos, np_, 1.5) && pose_valid(x, y, pos, np_, 1.5);
        }
        pos[np_][0] = -x; pos[np_][1] = -y; np_++;
        pos[np_][0] = x; pos[np_][1] = y; np_++;
      }
      for (int ag = 0; ag < P; ag++) {
        int pa = mt_uniform(&m, 0, 1) > 0.8 ? CASIM_POLICY_NONCOOPERATIVE : CASIM_POLICY_RVO;
        if (ag == 0) put_agent(in, ty, co, 0, pos[0], pos[1], ego_policy, ego_dyn, OAS | GRID, 1.0);
        else put_agent(in, ty, co, 2 * ag, pos[2 * ag], pos[2 * ag + 1], pa, other_dynamics, OAS | GRID, 0.5);
        int pb = mt_uniform(&m, 0, 1) > 0.8 ? CASIM_POLICY_NONCOOPERATIVE : CASIM_POLICY_RVO;
        put_agent(in, ty, co, 2 * ag + 1, pos[2 * ag + 1], pos[2 * ag], pb, other_dynamics, OAS, 0.5);
      }
    } else if (kind == 2) { /* train_agents_pairwise_swap, test_cases.py:1283-1363 */
      double ip[SC_MAX][2];
      int pol[SC_MAX];
      int k = 0;
      ip[0][0] = mt_uniform(&m, -7.5, 7.5); ip[0][1] = mt_uniform(&m, -7.5, 7.5); k = 1;
      for (int ag = 0; ag < n - 1; ag++) {
        double x, y;
        do { x = mt_uniform(&m, -7.5, 7.5); y = mt_uniform(&m, -7.5, 7.5); } while (!pose_valid(x, y, ip, k, 2.0));
        ip[k][0] = x; ip[k][1] = y; k++;
      }
      /* the policy draws come from numpy's stream, the shuffle (:1324) from Python's: independent, so numpy first */
      for (int q = 0; q < 2 * P; q++) pol[q] = mt_uniform(&m, 0, 1) > 0.8 ? CASIM_POLICY_NONCOOPERATIVE : CASIM_POLICY_RVO;
      mt_init_by_array(&m, &seeds[e], 1);
      for (int i = n - 1; i >= 1; i--) { /* random.shuffle */
        int j = py_randbelow(&m, i + 1);
        double tx = ip[i][0], ty = ip[i][1];
        ip[i][0] = ip[j][0]; ip[i][1] = ip[j][1];
        ip[j][0] = tx; ip[j][1] = ty;
      }
      for (int ag = 0; ag < P; ag++) {
        if (ag == 0) put_agent(in, ty, co, 0, ip[0], ip[1], ego_policy, ego_dyn, OAS, 1.0);
        else put_agent(in, ty, co, 2 * ag, ip[2 * ag], ip[2 * ag + 1], pol[2 * ag], other_dynamics, OAS, 0.5);
        put_agent(in, ty, co, 2 * ag + 1, ip[2 * ag + 1], ip[2 * ag], pol[2 * ag + 1], other_dynamics, OAS, 0.5);
      }
    } else {
      double ip[SC_MAX][2], gp[SC_MAX][2];
      int k = 0, ok = 0;
      double x, y, gx, gy;
      while (!ok) {
        x = mt_uniform(&m, -7.5, 7.5); y = mt_uniform(&m, -7.5, 7.5);
        gx = mt_uniform(&m, -7.5, 7.5); gy = mt_uniform(&m, -7.5, 7.5);
        double g1[1][2] = {{gx, gy}};
        ok = pose_valid(x, y, g1, 1, 4.0);
      }
      ip[k][0] = x; ip[k][1] = y; gp[k][0] = gx; gp[k][1] = gy; k++;
      for (int ag = 0; ag < n - 1; ag++) {
        ok = 0;
        while (!ok) {
          x = mt_uniform(&m, -7.5, 7.5); y = mt_uniform(&m, -7.5, 7.5);
          int v1 = pose_valid(x, y, ip, k, 1.5);
          gx = mt_uniform(&m, -7.5, 7.5); gy = mt_uniform(&m, -7.5, 7.5);
          int v2 = pose_valid(gx, gy, gp, k, 1.5);
          double i1[1][2] = {{x, y}};
          int v3 = pose_valid(gx, gy, i1, 1, 4.0);
          ok = v1 && v2 && v3;
        }
        ip[k][0] = x; ip[k][1] = y; gp[k][0] = gx; gp[k][1] = gy; k++;
      }
      mt_init_by_array(&m, &seeds[e], 1); /* random.seed(seed): the policy choices come from Python's generator */
      const int opts[2] = {CASIM_POLICY_RVO, CASIM_POLICY_NONCOOPERATIVE}; /* :1377 */
      for (int ag = 0; ag < P; ag++) {
        int pa = opts[py_choice2(&m)];
        if (ag == 0) put_agent(in, ty, co, 0, ip[0], ga3c ? gp[1] : gp[0], ego_policy, ego_dyn, OAS, 1.0); /* sic :1437 */
        else put_agent(in, ty, co, 2 * ag, ip[2 * ag], gp[2 * ag], pa, other_dynamics, OAS, 0.5);
        int pb = opts[py_choice2(&m)];
        put_agent(in, ty, co, 2 * ag + 1, ip[2 * ag + 1], gp[2 * ag + 1], pb, other_dynamics, OAS, 0.5);
      }
    }
  }
  return 0;
}
