/*
 * casim_scenario.cu -- the reference's seeded training-scenario generators, one environment per thread, written
 * straight into the device buffers casim_reset reads (no host loop, no H2D of the initial conditions):
 *   gym_collision_avoidance/envs/test_cases.py:1192-1281  train_agents_swap_circle(seed=s)
 *   gym_collision_avoidance/envs/test_cases.py:1365-1463  train_agents_random_positions(seed=s)
 *   gym_collision_avoidance/envs/test_cases.py:1283-1363  train_agents_pairwise_swap(seed=s)
 *   gym_collision_avoidance/envs/test_cases.py:129-133    is_pose_valid (rejection sampling on pairwise distance)
 * A seed gives the reference's own scenario: the generators draw from numpy's legacy global generator
 * (np.random.seed(s), np.random.uniform) and from CPython's `random` (random.seed(s), random.choice), both the
 * published MT19937 (init_genrand / init_by_array / genrand_int32; 53-bit doubles from two outputs), which each thread
 * runs in its own 624-word state (local memory: the generator is sequential by definition, the batch is the
 * parallelism).  Positions are computed in fp64 and stored as the fp32 initial-condition rows of casim_reset.
 */
#include "casim_device.cuh"

namespace {

struct Mt {
  uint32_t mt[624];
  int idx;
};

__device__ void mt_init_genrand(Mt& m, uint32_t s) { /* np.random.seed(int) */
  m.mt[0] = s;
  for (int i = 1; i < 624; i++) m.mt[i] = 1812433253u * (m.mt[i - 1] ^ (m.mt[i - 1] >> 30)) + (uint32_t)i;
  m.idx = 624;
}

__device__ void mt_init_by_key(Mt& m, uint32_t key) { /* random.seed(int < 2^32): init_by_array with a one-word key */
  mt_init_genrand(m, 19650218u);
  int i = 1;
  for (int k = 624; k; k--) {
    m.mt[i] = (m.mt[i] ^ ((m.mt[i - 1] ^ (m.mt[i - 1] >> 30)) * 1664525u)) + key; /* + key[j] + j with j == 0 */
    i++;
    if (i >= 624) { m.mt[0] = m.mt[623]; i = 1; }
  }
  for (int k = 623; k; k--) {
    m.mt[i] = (m.mt[i] ^ ((m.mt[i - 1] ^ (m.mt[i - 1] >> 30)) * 1566083941u)) - (uint32_t)i;
    i++;
    if (i >= 624) { m.mt[0] = m.mt[623]; i = 1; }
  }
  m.mt[0] = 0x80000000u;
  m.idx = 624;
}

__device__ uint32_t mt_next(Mt& m) {
  if (m.idx >= 624) {
    for (int k = 0; k < 624; k++) {
      const int k1 = k + 1 < 624 ? k + 1 : 0, k397 = k + 397 < 624 ? k + 397 : k + 397 - 624;
      const uint32_t y = (m.mt[k] & 0x80000000u) | (m.mt[k1] & 0x7fffffffu);
      m.mt[k] = m.mt[k397] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    m.idx = 0;
  }
  uint32_t y = m.mt[m.idx++];
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

__device__ double mt_uniform(Mt& m, double low, double high) { /* legacy uniform: low + (high - low) * random_sample() */
  const uint32_t a = mt_next(m) >> 5, b = mt_next(m) >> 6;
  const double u = ((double)a * 67108864.0 + (double)b) / 9007199254740992.0;
  return low + (high - low) * u;
}

__device__ int py_choice2(Mt& m) { /* random.choice of a 2-list: getrandbits(2) until < 2 */
  uint32_t r;
  do r = mt_next(m) >> 30; while (r >= 2u);
  return (int)r;
}

__device__ int py_randbelow(Mt& m, int n) { /* Random._randbelow_with_getrandbits: getrandbits(n.bit_length()) until < n */
  const int k = 32 - __clz(n);
  uint32_t r;
  do r = mt_next(m) >> (32 - k); while (r >= (uint32_t)n);
  return (int)r;
}

__device__ bool pose_valid(double x, double y, const double (*list)[2], int n, double distance) {
  for (int k = 0; k < n; k++) {
    const double dx = x - list[k][0], dy = y - list[k][1];
    if (sqrt(dx * dx + dy * dy) < distance) return false;
  }
  return true;
}

struct ScenarioArgs {
  int kind, E, A, n, ego_policy, ego_dyn, other_dyn, ga3c;
  const uint32_t* seeds;
  float* init;       /* [E][A][7] */
  int32_t* types;    /* [E][A] */
  int32_t* n_agents; /* [E] */
  float* coop;       /* [E][A] or NULL */
};

__device__ void put_agent(const ScenarioArgs& p, int e, int slot, const double* start, const double* goal, int policy,
                          int dyn, int sensors, float c) {
  float* r = p.init + ((size_t)e * p.A + slot) * 7;
  r[0] = (float)start[0]; r[1] = (float)start[1]; r[2] = (float)goal[0]; r[3] = (float)goal[1];
  r[4] = 0.5f;                    /* radius, test_cases.py:1195 / :1368 */
  r[5] = 1.0f;                    /* pref_speed */
  r[6] = __int_as_float(0x7fc00000); /* initial_heading None: towards the goal */
  p.types[(size_t)e * p.A + slot] = policy | (dyn << 8) | (sensors << 16);
  if (p.coop) p.coop[(size_t)e * p.A + slot] = c;
}

__global__ void __launch_bounds__(64) k_scenario(const ScenarioArgs p) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.E) return;
  const int n = p.n, P = n / 2;
  const int OAS = CASIM_SENSOR_OTHER_AGENTS, GRID = CASIM_SENSOR_OCCUPANCY_GRID;
  for (int s = 2 * P; s < p.A; s++) { /* unused slots */
    float* r = p.init + ((size_t)e * p.A + s) * 7;
    for (int k = 0; k < 7; k++) r[k] = 0.f;
    p.types[(size_t)e * p.A + s] = 0;
    if (p.coop) p.coop[(size_t)e * p.A + s] = 0.f;
  }
  p.n_agents[e] = 2 * P;
  const uint32_t seed = p.seeds[e];
  Mt m;
  mt_init_genrand(m, seed);
  if (p.kind == CASIM_SCENARIO_SWAP_CIRCLE) {
    double pos[2 * CASIM_MAX_AGENTS][2];
    int np_ = 0;
    double d = mt_uniform(m, 4.0, 8.0), a = mt_uniform(m, -CASIM_PI, CASIM_PI);
    double x = d * cos(a), y = d * sin(a);
    pos[np_][0] = -x; pos[np_][1] = -y; np_++;
    pos[np_][0] = x; pos[np_][1] = y; np_++;
    for (int ag = 0; ag < P - 1; ag++) {
      bool ok = false;
      while (!ok) {
        d = mt_uniform(m, 4.0, 8.0);
        a = mt_uniform(m, -CASIM_PI, CASIM_PI);
        x = d * cos(a); y = d * sin(a);
        ok = pose_valid(-x, -y, pos, np_, 1.5) && pose_valid(x, y, pos, np_, 1.5);
      }
      pos[np_][0] = -x; pos[np_][1] = -y; np_++;
      pos[np_][0] = x; pos[np_][1] = y; np_++;
    }
    for (int ag = 0; ag < P; ag++) {
      const int pa = mt_uniform(m, 0.0, 1.0) > 0.8 ? CASIM_POLICY_NONCOOPERATIVE : CASIM_POLICY_RVO;
      if (ag == 0) put_agent(p, e, 0, pos[0], pos[1], p.ego_policy, p.ego_dyn, OAS | GRID, 1.0f);
      else put_agent(p, e, 2 * ag, pos[2 * ag], pos[2 * ag + 1], pa, p.other_dyn, OAS | GRID, 0.5f);
      const int pb = mt_uniform(m, 0.0, 1.0) > 0.8 ? CASIM_POLICY_NONCOOPERATIVE : CASIM_POLICY_RVO;
      put_agent(p, e, 2 * ag + 1, pos[2 * ag + 1], pos[2 * ag], pb, p.other_dyn, OAS, 0.5f);
    }
  } else if (p.kind == CASIM_SCENARIO_PAIRWISE_SWAP) { /* test_cases.py:1283-1363 */
    double ip[CASIM_MAX_AGENTS][2];
    int pol[CASIM_MAX_AGENTS];
    int k = 1;
    ip[0][0] = mt_uniform(m, -7.5, 7.5);
    ip[0][1] = mt_uniform(m, -7.5, 7.5);
    for (int ag = 0; ag < n - 1; ag++) {
      double x, y;
      do {
        x = mt_uniform(m, -7.5, 7.5);
        y = mt_uniform(m, -7.5, 7.5);
      } while (!pose_valid(x, y, ip, k, 2.0));
      ip[k][0] = x; ip[k][1] = y; k++;
    }
    /* policy draws: numpy's stream; the shuffle (:1324): Python's.  Independent streams, so numpy's draws first. */
    for (int q = 0; q < 2 * P; q++) pol[q] = mt_uniform(m, 0.0, 1.0) > 0.8 ? CASIM_POLICY_NONCOOPERATIVE : CASIM_POLICY_RVO;
    mt_init_by_key(m, seed);
    for (int i = n - 1; i >= 1; i--) { /* random.shuffle */
      const int j = py_randbelow(m, i + 1);
      const double tx = ip[i][0], ty = ip[i][1];
      ip[i][0] = ip[j][0]; ip[i][1] = ip[j][1];
      ip[j][0] = tx; ip[j][1] = ty;
    }
    for (int ag = 0; ag < P; ag++) {
      if (ag == 0) put_agent(p, e, 0, ip[0], ip[1], p.ego_policy, p.ego_dyn, OAS, 1.0f);
      else put_agent(p, e, 2 * ag, ip[2 * ag], ip[2 * ag + 1], pol[2 * ag], p.other_dyn, OAS, 0.5f);
      put_agent(p, e, 2 * ag + 1, ip[2 * ag + 1], ip[2 * ag], pol[2 * ag + 1], p.other_dyn, OAS, 0.5f);
    }
  } else {
    double ip[CASIM_MAX_AGENTS][2], gp[CASIM_MAX_AGENTS][2];
    int k = 0;
    bool ok = false;
    double x = 0, y = 0, gx = 0, gy = 0;
    while (!ok) {
      x = mt_uniform(m, -7.5, 7.5); y = mt_uniform(m, -7.5, 7.5);
      gx = mt_uniform(m, -7.5, 7.5); gy = mt_uniform(m, -7.5, 7.5);
      const double g1[1][2] = {{gx, gy}};
      ok = pose_valid(x, y, g1, 1, 4.0);
    }
    ip[k][0] = x; ip[k][1] = y; gp[k][0] = gx; gp[k][1] = gy; k++;
    for (int ag = 0; ag < n - 1; ag++) {
      ok = false;
      while (!ok) {
        x = mt_uniform(m, -7.5, 7.5); y = mt_uniform(m, -7.5, 7.5);
        const bool v1 = pose_valid(x, y, ip, k, 1.5);
        gx = mt_uniform(m, -7.5, 7.5); gy = mt_uniform(m, -7.5, 7.5);
        const bool v2 = pose_valid(gx, gy, gp, k, 1.5);
        const double i1[1][2] = {{x, y}};
        const bool v3 = pose_valid(gx, gy, i1, 1, 4.0);
        ok = v1 && v2 && v3;
      }
      ip[k][0] = x; ip[k][1] = y; gp[k][0] = gx; gp[k][1] = gy; k++;
    }
    mt_init_by_key(m, seed); /* random.seed(seed): the policy choices come from Python's generator (:1426, :1451) */
    for (int ag = 0; ag < P; ag++) {
      const int pa = py_choice2(m) ? CASIM_POLICY_NONCOOPERATIVE : CASIM_POLICY_RVO; /* [RVOPolicy, NonCooperativePolicy], :1377 */
      if (ag == 0) put_agent(p, e, 0, ip[0], p.ga3c ? gp[1] : gp[0], p.ego_policy, p.ego_dyn, OAS, 1.0f); /* sic :1437 */
      else put_agent(p, e, 2 * ag, ip[2 * ag], gp[2 * ag], pa, p.other_dyn, OAS, 0.5f);
      const int pb = py_choice2(m) ? CASIM_POLICY_NONCOOPERATIVE : CASIM_POLICY_RVO;
      put_agent(p, e, 2 * ag + 1, ip[2 * ag + 1], gp[2 * ag + 1], pb, p.other_dyn, OAS, 0.5f);
    }
  }
}

}  // namespace

cudaError_t casim_launch_scenario(casim_handle* h, int kind, int number_of_agents, const uint32_t* seeds, int ego_policy,
                                  int ego_dynamics, int other_dynamics, float* init, int32_t* types, int32_t* n_agents,
                                  float* coop, cudaStream_t s) {
  ScenarioArgs a;
  a.kind = kind;
  a.E = h->E;
  a.A = h->A;
  a.n = number_of_agents > 2 ? number_of_agents : 2;
  a.ga3c = ego_policy == CASIM_POLICY_GA3C_CADRL;
  a.ego_policy = ego_policy;
  a.ego_dyn = a.ga3c ? CASIM_DYN_UNICYCLE_MAX_ACC : ego_dynamics; /* :1254-1257, :1435-1438 */
  a.other_dyn = other_dynamics;
  a.seeds = seeds;
  a.init = init;
  a.types = types;
  a.n_agents = n_agents;
  a.coop = coop;
  h->launches++;
  k_scenario<<<(h->E + 63) / 64, 64, 0, s>>>(a);
  return cudaGetLastError();
}
