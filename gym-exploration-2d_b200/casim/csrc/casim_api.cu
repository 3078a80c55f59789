/*
 * casim_api.cu -- the C ABI of libcasim.so (include/casim.h): handle management, configuration
 * snapshot, and the entry points that launch the kernels.  No torch types, no host threads, no
 * host synchronisation except in the *_host variant and in create/destroy.
 */
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "casim_internal.h"

static thread_local char g_err[512] = "";

void casim_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

#define CK(call)                                                                      \
  do {                                                                                \
    cudaError_t _e = (call);                                                          \
    if (_e != cudaSuccess) {                                                          \
      casim_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return (int)_e;                                                                 \
    }                                                                                 \
  } while (0)

#define REQUIRE(cond, msg)                     \
  do {                                         \
    if (!(cond)) {                             \
      casim_set_error("casim: %s", msg);       \
      return CASIM_EINVAL;                     \
    }                                          \
  } while (0)

extern "C" {

const char* casim_last_error(void) { return g_err; }
int casim_abi_version(void) { return CASIM_ABI_VERSION; }

/* Reference defaults: config.py:29-95, LaserScanSensor.py:13-20, collision_avoidance_env.py:97,755-774,
 * ig_greedy.py:18-29, experiments/src/dmcts.py:73-78. */
void casim_default_config(casim_config* c) {
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
  c->reward_min = -10.0;
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
    /* np.linspace(-pi, pi, 16) and np.arange(0, 6, 2*pi/16) */
    const double PI = 3.14159265358979323846;
    int n = c->laser_num_beams;
    double start = -PI, stop = PI, step = (stop - start) / (double)(n - 1);
    for (int i = 0; i < n; i++) c->laser_angles[i] = (double)i * step + start;
    c->laser_angles[n - 1] = stop;
    double rstep = 2 * PI / 16.0;
    int nr = (int)ceil((6.0 - 0.0) / rstep);
    c->laser_num_ranges = nr;
    for (int i = 0; i < nr; i++) c->laser_ranges[i] = 0.0 + (double)i * rstep;
    c->ig_sens_fov = 60.0 * PI / 180.0;
  }
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

static int derive(const casim_config* c, DevCfg* d) {
  memset(d, 0, sizeof(*d));
  REQUIRE(c->laser_num_beams >= 0 && c->laser_num_beams <= CASIM_MAX_BEAMS, "laser_num_beams out of range");
  REQUIRE(c->laser_num_ranges >= 0 && c->laser_num_ranges <= CASIM_MAX_RANGES, "laser_num_ranges out of range");
  REQUIRE(c->map_cell > 0, "map_cell must be positive");
  REQUIRE((int)(c->map_width / c->map_cell) == CASIM_MAP_DIM && (int)(c->map_height / c->map_cell) == CASIM_MAP_DIM,
          "the static map must be 300x300 cells");
  REQUIRE(c->max_other_agents_rows >= 1 && c->max_other_agents_rows <= CASIM_MAX_AGENTS, "max_other_agents_rows");
  REQUIRE(c->submap_w >= 1 && c->submap_w <= 64 && c->submap_h >= 1 && c->submap_h <= 64, "submap must be <= 64x64");
  d->dt = c->dt;
  d->dt_nominal = c->dt;
  d->r_at_goal = c->reward_at_goal;
  d->r_coll_agent = c->reward_collision_with_agent;
  d->r_timeout = c->reward_timeout;
  d->r_infeasible = c->reward_infeasible;
  d->r_coll_wall = c->reward_collision_with_wall;
  d->r_time_step = c->reward_time_step;
  d->r_dist_goal = c->reward_distance_to_goal;
  d->r_wiggly = c->reward_wiggly_behavior;
  d->wiggly_thr = c->wiggly_behavior_threshold;
  d->r_min = c->reward_min;
  d->r_max = c->reward_max;
  d->r_inv_span = 1.0 / (c->reward_max - c->reward_min);
  d->close_range = c->getting_close_range;
  d->near_goal = c->near_goal_threshold;
  d->max_time_ratio = c->max_time_ratio;
  d->sensing_horizon = c->sensing_horizon;
  d->max_heading_change = c->max_heading_change;
  d->map_ox = (c->map_width / 2.) / c->map_cell;
  d->map_oy = (c->map_height / 2.) / c->map_cell;
  d->cell = c->map_cell;
  d->inv_cell = 1.0 / c->map_cell;
  d->laser_max_range = c->laser_max_range;
  d->laser_norm = c->laser_norm_range;
  d->nb = c->laser_num_beams;
  d->nr = c->laser_num_ranges;
  for (int i = 0; i < d->nb; i++) {
    d->laser_ang[i] = c->laser_angles[i];
    d->laser_cos[i] = cos(c->laser_angles[i]);
    d->laser_sin[i] = sin(c->laser_angles[i]);
  }
  for (int i = 0; i < d->nr; i++) d->laser_r[i] = c->laser_ranges[i];
  d->laser_sorted = 1;
  for (int i = 1; i < d->nr; i++)
    if (!(c->laser_ranges[i] > c->laser_ranges[i - 1])) d->laser_sorted = 0;
  /* LaserScanSensor.py:58 `1 - ranges / max_range`, fp64 then stored fp32: one IEEE division per table entry, here instead of per CTA */
  for (int i = 0; i < CASIM_MAX_RANGES; i++) d->laser_out[i] = i < d->nr ? (float)(1.0 - d->laser_r[i] / d->laser_norm) : 0.f;
  d->laser_out[CASIM_MAX_RANGES] = (float)(1.0 - d->laser_max_range / d->laser_norm);
  d->ig_fov_half = c->ig_sens_fov / 2;
  d->ig_detect_fov_half = c->ig_detect_fov / 2.0;
  d->ig_range = c->ig_range;
  d->ig_cell = c->ig_cell;
  d->ig_rocc = c->ig_rocc;
  d->ig_remp = c->ig_remp;
  d->ig_tol = c->ig_tolerance;
  d->ig_pfn = c->ig_p_false_neg;
  d->ig_pfp = c->ig_p_false_pos;
  d->ig_dt = c->ig_dt;
  d->ig_det_thresh = sqrt(0.5) * c->ig_cell + c->ig_tolerance;
  d->ig_fov = c->ig_sens_fov;
  d->ig_tan_half = (c->ig_sens_fov / 2 < 1.5) ? tan(c->ig_sens_fov / 2) : 0.0;
  d->ig_prior = c->ig_prior;
  d->map_hw = c->map_width / 2;
  d->map_hh = c->map_height / 2;
  REQUIRE(c->ig_map_dim >= 1 && c->ig_map_dim <= 64, "ig_map_dim must be in 1..64");
  d->coll_w_static = c->collision_av_w_static_agent;
  d->max_obs = c->max_other_agents_observed;
  d->rows = c->max_other_agents_rows;
  d->submap_w = c->submap_w;
  d->submap_h = c->submap_h;
  d->game_over_mode = c->game_over_mode;
  d->auto_reset = c->auto_reset;
  d->ig_dim = c->ig_map_dim;
  return 0;
}

int casim_create(const casim_config* cfg, int E, int A, casim_handle** out) {
  REQUIRE(cfg && out, "null argument");
  REQUIRE(E >= 1, "E must be >= 1");
  REQUIRE(A >= 1 && A <= CASIM_MAX_AGENTS, "A must be in 1..32");
  casim_handle* h = (casim_handle*)calloc(1, sizeof(casim_handle));
  REQUIRE(h, "out of host memory");
  h->cfg = *cfg;
  int rc = derive(cfg, &h->dc);
  if (rc) {
    free(h);
    return rc;
  }
  h->E = E;
  h->A = A;
  cudaError_t e = cudaGetDevice(&h->device);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, h->device);
  int epw = 32 / A;
  (void)epw;
  h->max_ctas = E; /* upper bound of the core grid: one warp per environment, one warp per CTA */
  size_t N = (size_t)E * A;
  if (e == cudaSuccess) e = cudaMalloc((void**)&h->pose64, N * 4 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc((void**)&h->prev_game_over, (size_t)E);
  if (e == cudaSuccess) e = cudaMalloc((void**)&h->state0, N * 24 * sizeof(float));
  if (e == cudaSuccess) e = cudaMemset(h->state0, 0, N * 24 * sizeof(float));
  /* statistics accumulators: 8 words 64 B apart per step of a launch (casim_step_n: up to CASIM_STEP_N_MAX steps) */
  const size_t partial_words = (size_t)(h->max_ctas > 32 ? h->max_ctas : 32) * CASIM_NUM_STATS > (size_t)CASIM_STEP_N_MAX * 64
                                   ? (size_t)(h->max_ctas > 32 ? h->max_ctas : 32) * CASIM_NUM_STATS
                                   : (size_t)CASIM_STEP_N_MAX * 64;
  if (e == cudaSuccess) e = cudaMalloc((void**)&h->partials, partial_words * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc((void**)&h->ticket, sizeof(unsigned int));
  if (e == cudaSuccess) e = cudaMemset(h->prev_game_over, 0, (size_t)E);
  if (e == cudaSuccess) e = cudaMemset(h->ticket, 0, sizeof(unsigned int));
  if (e == cudaSuccess) e = cudaMemset(h->partials, 0, partial_words * sizeof(double));
  if (e == cudaSuccess) e = cudaMemset(h->pose64, 0, N * 4 * sizeof(double));
  if (e != cudaSuccess) {
    casim_set_error("casim_create: %s", cudaGetErrorString(e));
    casim_destroy(h);
    return (int)e;
  }
  *out = h;
  return 0;
}

int casim_destroy(casim_handle* h) {
  if (!h) return 0;
  cudaFree(h->pose64);
  cudaFree(h->prev_game_over);
  cudaFree(h->state0);
  casim_rvo_obstacles_free(h);
  cudaFree(h->partials);
  cudaFree(h->ticket);
  cudaFree(h->map_planes);
  cudaFree(h->edf_d2);
  cudaFree(h->edf);
  cudaFree(h->edf_scratch);
  cudaFree(h->grid_maps);
  cudaFree(h->grid2_map);
  cudaFree(h->grid_work);
  cudaFree(h->ga3c_w);
  cudaFree(h->ga3c_wb);
  cudaFree(h->ga3c_x);
  cudaFree(h->ga3c_xb);
  cudaFree(h->ga3c_fix);
  if (h->pipe_stream) {
    for (int k = 0; k < 2; k++) {
      cudaEventDestroy(h->pipe_ev_kernels[k]);
      cudaEventDestroy(h->pipe_ev_copied[k]);
    }
    cudaStreamDestroy(h->pipe_stream);
  }
  free(h);
  return 0;
}

int casim_set_config(casim_handle* h, const casim_config* cfg) {
  REQUIRE(h && cfg, "null argument");
  DevCfg d;
  int rc = derive(cfg, &d);
  if (rc) return rc;
  h->cfg = *cfg;
  h->dc = d;
  return 0;
}

static int ensure_maps(casim_handle* h, int n_maps) {
  if (n_maps <= h->n_maps_alloc) return 0;
  cudaFree(h->map_planes);
  cudaFree(h->edf_d2);
  cudaFree(h->edf);
  cudaFree(h->edf_scratch);
  cudaFree(h->grid_maps);
  h->grid_maps = nullptr;
  h->map_planes = nullptr;
  h->edf_d2 = h->edf_scratch = nullptr;
  h->edf = nullptr;
  h->n_maps_alloc = 0;
  CK(cudaMalloc((void**)&h->map_planes, (size_t)n_maps * CASIM_NUM_PLANES * CASIM_PLANE_WORDS * sizeof(uint32_t)));
  CK(cudaMalloc((void**)&h->edf_d2, (size_t)n_maps * CASIM_MAP_DIM * CASIM_MAP_DIM * sizeof(int32_t)));
  CK(cudaMalloc((void**)&h->edf, (size_t)n_maps * CASIM_MAP_DIM * CASIM_MAP_DIM * sizeof(double)));
  CK(cudaMalloc((void**)&h->edf_scratch, (size_t)n_maps * CASIM_MAP_DIM * CASIM_MAP_DIM * sizeof(int32_t)));
  CK(cudaMalloc((void**)&h->grid_maps, (size_t)n_maps * CASIM_GMAP_WORDS * sizeof(uint32_t)));
  if (!h->grid2_map) {
    CK(cudaMalloc((void**)&h->grid2_map, CASIM_G2_BYTES));
    CK(cudaMalloc((void**)&h->grid_work, 2 * sizeof(unsigned int)));
    CK(cudaMemset(h->grid_work, 0, 2 * sizeof(unsigned int)));
    h->grid2_ready = 0;
  }
  h->n_maps_alloc = n_maps;
  return 0;
}

int casim_set_obstacles(casim_handle* h, const float* rects, const int32_t* n_rects, int n_maps, int max_rects,
                        uint8_t* map_out, void* stream) {
  REQUIRE(h, "null handle");
  REQUIRE(n_maps == 0 || n_maps == 1 || n_maps == h->E, "n_maps must be 0, 1 or E");
  if (n_maps == 0) {
    h->n_maps = 0;
    return 0;
  }
  REQUIRE(rects && n_rects && max_rects >= 0, "null rectangles");
  int rc = ensure_maps(h, n_maps);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  CK(casim_launch_raster(h, rects, n_rects, n_maps, max_rects, s));
  CK(casim_launch_map_derive(h, n_maps, map_out, s));
  h->n_maps = n_maps;
  return 0;
}

int casim_set_map(casim_handle* h, const uint8_t* static_map, int n_maps, void* stream) {
  REQUIRE(h, "null handle");
  REQUIRE(n_maps == 0 || n_maps == 1 || n_maps == h->E, "n_maps must be 0, 1 or E");
  if (n_maps == 0) {
    h->n_maps = 0;
    return 0;
  }
  REQUIRE(static_map, "null map");
  int rc = ensure_maps(h, n_maps);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  CK(casim_launch_pack_map(h, static_map, n_maps, s));
  CK(casim_launch_map_derive(h, n_maps, nullptr, s));
  h->n_maps = n_maps;
  return 0;
}

int casim_set_stats_event(casim_handle* h, void* cuda_event) {
  REQUIRE(h, "null handle");
  h->stats_event = (cudaEvent_t)cuda_event;
  return 0;
}

int casim_set_grid_packed_output(casim_handle* h, uint8_t* packed) {
  REQUIRE(h, "null handle");
  REQUIRE((((uintptr_t)packed) & 1) == 0, "packed local_grid buffer must be 2-byte aligned");
  h->grid_packed = packed;
  return 0;
}

int casim_set_trajectories(casim_handle* h, const float* traj, int T) {
  REQUIRE(h, "null handle");
  REQUIRE((traj == nullptr) == (T == 0), "traj and T must both be set or both be empty");
  h->traj = traj;
  h->T = T;
  return 0;
}

static int check_obs(const casim_handle* h, const casim_obs* obs) {
  REQUIRE(obs, "null casim_obs");
  (void)h;
  return 0;
}

static int run_sensors(casim_handle* h, const float* state, const casim_obs* obs, cudaStream_t s) {
  /* both map sensors on a shared map: ONE persistent kernel stages the map once and produces both observations */
  if (obs->laserscan && obs->local_grid && casim_grid_can_fuse_laser(h, obs->local_grid)) {
    CK(casim_launch_grid(h, state, obs->local_grid, s, obs->laserscan));
    return 0;
  }
  if (obs->laserscan) CK(casim_launch_laser(h, state, obs->laserscan, s));
  if (obs->local_grid) CK(casim_launch_grid(h, state, obs->local_grid, s));
  return 0;
}

int casim_reset(casim_handle* h, float* state, const float* init, const int32_t* type, const int32_t* n_agents,
                const casim_obs* obs, void* stream) {
  REQUIRE(h && state && init && type, "null argument");
  int rc = check_obs(h, obs);
  if (rc) return rc;
  h->init = init;
  h->type0 = type;
  h->n_agents = n_agents;
  cudaStream_t s = (cudaStream_t)stream;
  CK(casim_launch_core(h, CASIM_MODE_RESET, state, nullptr, nullptr, h->cfg.dt, obs, s));
  return run_sensors(h, state, obs, s);
}

int casim_reset_envs(casim_handle* h, float* state, const uint8_t* mask, const casim_obs* obs, void* stream) {
  REQUIRE(h && state && mask, "null argument");
  REQUIRE(h->init && h->type0, "casim_reset_envs needs a previous casim_reset (it re-reads its init / type / n_agents buffers)");
  int rc = check_obs(h, obs);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  CK(casim_launch_core(h, CASIM_MODE_RESET, state, nullptr, nullptr, h->cfg.dt, obs, s, mask));
  /* the sensor kernels recompute every agent from the fp64 pose scratch; untouched environments get identical bytes */
  return run_sensors(h, state, obs, s);
}

int casim_observe(casim_handle* h, const float* state, const casim_obs* obs, void* stream) {
  REQUIRE(h && state, "null argument");
  int rc = check_obs(h, obs);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  CK(casim_launch_core(h, CASIM_MODE_OBSERVE, const_cast<float*>(state), nullptr, nullptr, h->cfg.dt, obs, s));
  return run_sensors(h, state, obs, s);
}

int casim_step(casim_handle* h, float* state, const float* ext_actions, const float* policy_actions, double dt,
               const casim_obs* obs, void* stream) {
  REQUIRE(h && state, "null argument");
  REQUIRE((((uintptr_t)ext_actions) & 7) == 0, "ext_actions must be 8-byte aligned");
  int rc = check_obs(h, obs);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  double dtd = dt > 0 ? dt : h->cfg.dt;
  CK(casim_launch_core(h, CASIM_MODE_STEP, state, ext_actions, policy_actions, dtd, obs, s));
  if (h->stats_event) CK(cudaEventRecord(h->stats_event, s)); /* rewards / done / game_over / stats are final from here */
  return run_sensors(h, state, obs, s);
}

int casim_step_n(casim_handle* h, float* state, const float* policy_actions, double dt, int n_steps, const casim_obs* obs,
                 void* stream) {
  REQUIRE(h && state, "null argument");
  REQUIRE(n_steps >= 1 && n_steps <= CASIM_STEP_N_MAX, "casim_step_n: n_steps must be in 1..CASIM_STEP_N_MAX");
  int rc = check_obs(h, obs);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  double dtd = dt > 0 ? dt : h->cfg.dt;
  CK(casim_launch_core(h, CASIM_MODE_STEP, state, nullptr, policy_actions, dtd, obs, s, nullptr, n_steps));
  if (h->stats_event) CK(cudaEventRecord(h->stats_event, s));
  return run_sensors(h, state, obs, s);
}

int casim_run_sensor(casim_handle* h, const float* state, const casim_obs* obs, int which, void* stream) {
  REQUIRE(h && state && obs, "null argument");
  cudaStream_t s = (cudaStream_t)stream;
  if (which == CASIM_SENSOR_LASERSCAN) {
    REQUIRE(obs->laserscan, "laserscan buffer missing");
    CK(casim_launch_laser(h, state, obs->laserscan, s));
  } else if (which == CASIM_SENSOR_OCCUPANCY_GRID) {
    REQUIRE(obs->local_grid, "local_grid buffer missing");
    CK(casim_launch_grid(h, state, obs->local_grid, s));
  } else if (which == (CASIM_SENSOR_LASERSCAN | CASIM_SENSOR_OCCUPANCY_GRID)) {
    REQUIRE(obs->local_grid && obs->laserscan, "sensor buffer missing");
    return run_sensors(h, state, obs, s); /* what casim_step launches: fused when the map is shared */
  } else {
    REQUIRE(false, "which must be CASIM_SENSOR_LASERSCAN, CASIM_SENSOR_OCCUPANCY_GRID or both");
  }
  return 0;
}

/* Device-visible alias of a pinned (page-locked, mapped) host buffer, or NULL when `p` is pageable / not mapped. */
static void* mapped_alias(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return at.type == cudaMemoryTypeHost ? at.devicePointer : nullptr;
}

/* Shared body of the host-buffer steps: actions in (zero-copy read or H2D copy on `s`), the step with the small results
 * stored straight into mapped host memory, and the list of (dst, src, bytes) D2H copies that remain to be issued. */
struct HostPending { void* dst; const void* src; size_t bytes; };

static int host_step_launch(casim_handle* h, float* state, const float* ext_actions_host, float* ext_actions_dev,
                            const float* policy_actions, double dt, const casim_obs* dev, const casim_obs* host,
                            cudaStream_t s, HostPending* pend, int* n_pend) {
  size_t N = (size_t)h->E * h->A;
  static const int zc_on = casim_tuning_int("CASIM_ZEROCOPY", 1);
  /* Actions.  Pinned + mapped and small: the step kernel reads them from host memory itself (every lane, one
   * coalesced 8-byte read) and refreshes ext_actions_dev on the way -- the copy-engine hop in front of a ~10 us
   * kernel costs more than the kernel.  Otherwise: cudaMemcpyAsync H2D. */
  const float* ext_src = nullptr;
  if (ext_actions_host) {
    REQUIRE(ext_actions_dev, "ext_actions_dev required with ext_actions_host");
    void* al = (zc_on && N * 2 * sizeof(float) <= CASIM_ZEROCOPY_MAX) ? mapped_alias(ext_actions_host) : nullptr;
    if (al) {
      ext_src = (const float*)al;
      h->ext_copy_dst = ext_actions_dev;
    } else {
      CK(cudaMemcpyAsync(ext_actions_dev, ext_actions_host, N * 2 * sizeof(float), cudaMemcpyHostToDevice, s));
      ext_src = ext_actions_dev;
    }
  }
  /* Result delivery.  A result whose host buffer is pinned + mapped and at most CASIM_ZEROCOPY_MAX bytes is
   * written by the kernels STRAIGHT into host memory over PCIe (posted, coalesced stores; the field's device
   * buffer is not touched by this call): no copy-engine hop and no extra stream operation per field, which is
   * what bounds a small step (C2: four D2H copies cost more than the step kernel).  Larger results (sensor
   * planes) and pageable buffers go through cudaMemcpyAsync after the kernels. */
  casim_obs run = *dev;
  int np = 0;
#define OUT(field, nbytes, zc_ok)                                                                \
  if (host->field) {                                                                             \
    void* al = ((zc_ok) && zc_on && (size_t)(nbytes) <= CASIM_ZEROCOPY_MAX) ? mapped_alias(host->field) : nullptr; \
    if (al)                                                                                      \
      run.field = reinterpret_cast<decltype(run.field)>(al);                                     \
    else {                                                                                       \
      REQUIRE(dev->field, "host buffer given without its device buffer: " #field);               \
      pend[np].dst = host->field; pend[np].src = dev->field; pend[np].bytes = (nbytes); np++;    \
    }                                                                                            \
  }
  OUT(rewards, N * sizeof(float), 1)
  OUT(done, N, 1)
  OUT(game_over, (size_t)h->E, 1)
  OUT(collision_flags, N, 1)
  OUT(nearest_dist, N * sizeof(float), 1)
  OUT(stats, CASIM_NUM_STATS * sizeof(float), 1)
  OUT(obs_misc, N * 4 * sizeof(float), 1)
  /* sensor planes are inputs of casim_explore_step / casim_ga3c_forward / the next observe: device copy stays current */
  OUT(other_agents_states, N * (size_t)h->dc.rows * 10 * sizeof(float), 0)
  OUT(laserscan, N * (size_t)h->dc.nb * sizeof(float), 0)
  OUT(local_grid, N * (size_t)h->dc.submap_w * h->dc.submap_h, 0)
#undef OUT
  int rc = casim_step(h, state, ext_src, policy_actions, dt, &run, s);
  h->ext_copy_dst = nullptr;
  *n_pend = np;
  return rc;
}

int casim_step_host(casim_handle* h, float* state, const float* ext_actions_host, float* ext_actions_dev,
                    const float* policy_actions, double dt, const casim_obs* dev, const casim_obs* host,
                    void* stream) {
  REQUIRE(h && state && dev && host, "null argument");
  cudaStream_t s = (cudaStream_t)stream;
  HostPending pend[12];
  int np = 0;
  int rc = host_step_launch(h, state, ext_actions_host, ext_actions_dev, policy_actions, dt, dev, host, s, pend, &np);
  if (rc) return rc;
  for (int i = 0; i < np; i++) CK(cudaMemcpyAsync(pend[i].dst, pend[i].src, pend[i].bytes, cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  return 0;
}

int casim_step_host_pipelined(casim_handle* h, float* state, const float* ext_actions_host, float* ext_actions_dev,
                              const float* policy_actions, double dt, const casim_obs* dev, const casim_obs* host,
                              uint8_t* grid_packed_dev, uint8_t* grid_packed_host, void* stream) {
  REQUIRE(h && state && dev && host, "null argument");
  REQUIRE((grid_packed_dev == nullptr) == (grid_packed_host == nullptr), "packed grid: give both buffers or neither");
  cudaStream_t s = (cudaStream_t)stream;
  if (!h->pipe_stream) {
    CK(cudaStreamCreateWithFlags(&h->pipe_stream, cudaStreamNonBlocking));
    for (int k = 0; k < 2; k++) {
      CK(cudaEventCreateWithFlags(&h->pipe_ev_kernels[k], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&h->pipe_ev_copied[k], cudaEventDisableTiming));
    }
  }
  const int slot = (int)(h->pipe_steps & 1ull);
  /* step t: kernels on the caller's stream; with a packed grid the dense local_grid stays on the device */
  casim_obs host_run = *host;
  uint8_t* keep_packed = h->grid_packed;
  if (grid_packed_dev) {
    REQUIRE(dev->local_grid, "packed grid delivery needs the device local_grid buffer");
    h->grid_packed = grid_packed_dev;
    host_run.local_grid = nullptr;
  }
  HostPending pend[12];
  int np = 0;
  int rc = host_step_launch(h, state, ext_actions_host, ext_actions_dev, policy_actions, dt, dev, &host_run, s, pend, &np);
  h->grid_packed = keep_packed;
  if (rc) return rc;
  if (grid_packed_dev) {
    pend[np].dst = grid_packed_host;
    pend[np].src = grid_packed_dev;
    pend[np].bytes = (size_t)h->E * h->A * (size_t)((h->dc.submap_w * h->dc.submap_h + 7) / 8);
    np++;
  }
  CK(cudaEventRecord(h->pipe_ev_kernels[slot], s));
  /* its D2H copies on the copy stream: they overlap the kernels of step t+1 */
  CK(cudaStreamWaitEvent(h->pipe_stream, h->pipe_ev_kernels[slot], 0));
  for (int i = 0; i < np; i++)
    CK(cudaMemcpyAsync(pend[i].dst, pend[i].src, pend[i].bytes, cudaMemcpyDeviceToHost, h->pipe_stream));
  CK(cudaEventRecord(h->pipe_ev_copied[slot], h->pipe_stream));
  /* what the caller gets back now: step t-1, complete in ITS host buffers */
  if (h->pipe_steps > 0) CK(cudaEventSynchronize(h->pipe_ev_copied[slot ^ 1]));
  h->pipe_steps++;
  return 0;
}

int casim_host_flush(casim_handle* h) {
  REQUIRE(h, "null handle");
  if (h->pipe_stream && h->pipe_steps > 0) CK(cudaEventSynchronize(h->pipe_ev_copied[(int)((h->pipe_steps - 1) & 1ull)]));
  return 0;
}

int casim_explore_step(casim_handle* h, const float* state, const float* other_agents_states, float* belief,
                       float* policy_actions, float* best_mi, float* cand_mi, unsigned long long* counters,
                       void* stream) {
  REQUIRE(h && state && other_agents_states && belief && policy_actions, "null argument");
  REQUIRE(h->n_maps > 0, "casim_explore_step needs a static map (casim_set_obstacles / casim_set_map)");
  CK(casim_launch_explore(h, state, other_agents_states, belief, policy_actions, best_mi, cand_mi, counters,
                          (cudaStream_t)stream));
  return 0;
}

int casim_generate_scenario(casim_handle* h, int kind, int number_of_agents, const uint32_t* seeds, int ego_policy,
                            int ego_dynamics, int other_dynamics, float* init, int32_t* types, int32_t* n_agents,
                            float* coop, void* stream) {
  REQUIRE(h && seeds && init && types && n_agents, "null argument");
  REQUIRE(kind >= CASIM_SCENARIO_SWAP_CIRCLE && kind <= CASIM_SCENARIO_PAIRWISE_SWAP, "unknown scenario kind");
  const int n = number_of_agents > 2 ? number_of_agents : 2;
  REQUIRE(n <= CASIM_MAX_AGENTS && 2 * (n / 2) <= h->A, "scenario: more agents than slots per environment");
  CK(casim_launch_scenario(h, kind, number_of_agents, seeds, ego_policy, ego_dynamics, other_dynamics, init, types, n_agents,
                           coop, (cudaStream_t)stream));
  return 0;
}

int casim_explore_team_update(casim_handle* h, const float* state, const float* other_agents_states, float* belief,
                              const double* targets_prev, const int32_t* n_prev, double* targets_new, int32_t* n_new,
                              uint32_t* team_mask, float* team_reward, void* stream) {
  REQUIRE(h && state && other_agents_states && belief && targets_prev && n_prev && targets_new && n_new, "null argument");
  REQUIRE(targets_prev != targets_new && n_prev != n_new, "team update: previous and new target lists must not alias");
  REQUIRE(h->n_maps > 0, "casim_explore_team_update needs a static map (casim_set_obstacles / casim_set_map)");
  CK(casim_launch_team_update(h, state, other_agents_states, belief, targets_prev, n_prev, targets_new, n_new, team_mask,
                              team_reward, (cudaStream_t)stream));
  return 0;
}

int casim_explore_mi_table(casim_handle* h, const float* state, const float* belief, int64_t* table, void* stream) {
  REQUIRE(h && state && belief && table, "null argument");
  CK(casim_launch_mi_table(h, state, belief, reinterpret_cast<long long*>(table), (cudaStream_t)stream));
  return 0;
}

int casim_explore_set_mi_table(casim_handle* h, const int64_t* table) {
  REQUIRE(h, "null handle");
  h->mi_table = reinterpret_cast<const long long*>(table);
  return 0;
}

int casim_explore_set_rollout_slots(casim_handle* h, const int32_t* slots, int n_slots) {
  REQUIRE(h, "null handle");
  REQUIRE(slots == nullptr || (n_slots >= 1 && (size_t)n_slots <= (size_t)h->E * h->A), "rollout slots: 1 <= n_slots <= E * A");
  h->rollout_slots = slots;
  h->n_rollout_slots = slots ? n_slots : 0;
  return 0;
}

int casim_explore_rollouts(casim_handle* h, const float* state, const float* belief, const uint8_t* actions, int K,
                           int H, int xdt, const uint32_t* exclude, int exclude_per_rollout, float* rewards,
                           float* end_pose, uint32_t* obsvd, int32_t* n_feasible, void* stream) {
  REQUIRE(h && state && belief && actions && rewards, "null argument");
  REQUIRE(h->n_maps > 0, "casim_explore_rollouts needs a static map (casim_set_obstacles / casim_set_map)");
  REQUIRE(K >= 1 && H >= 1 && H <= CASIM_ROLLOUT_MAX_H, "rollouts: K >= 1 and 1 <= H <= CASIM_ROLLOUT_MAX_H");
  REQUIRE(xdt >= 1, "rollouts: xdt >= 1");
  REQUIRE((size_t)h->E * h->A * (size_t)K <= 0x7fffffffull, "rollouts: E * A * K exceeds the grid limit");
  CK(casim_launch_rollouts(h, state, belief, actions, K, H, xdt, exclude, exclude_per_rollout, rewards, end_pose, obsvd,
                           n_feasible, (cudaStream_t)stream));
  return 0;
}

static int mcts_check(const casim_handle* h, const casim_mcts_trees* t) {
  REQUIRE(h && t, "null argument");
  REQUIRE(t->T >= 1 && t->M >= 1 && t->H >= 1 && t->H <= CASIM_ROLLOUT_MAX_H && t->nsims >= 0, "mcts: bad tree dimensions");
  REQUIRE(t->slot && t->n_nodes && t->node && t->parent && t->child && t->nchild && t->stage && t->has_rollout && t->mu &&
              t->N && t->best_reward && t->path && t->best_seq && t->overflow, "mcts: null tree array");
  REQUIRE(t->n_comms == 0 || (t->comm_q && t->comm_masks && t->comm_valid && t->comm_n >= 1), "mcts: null communication array");
  return 0;
}

int casim_mcts_select(casim_handle* h, const casim_mcts_trees* trees, const uint8_t* draws, uint64_t draw_seed,
                      int iteration, const double* comm_u, uint8_t* actions, uint32_t* exclude, int root_only, void* stream) {
  int rc = mcts_check(h, trees);
  if (rc) return rc;
  REQUIRE(actions, "mcts: null action buffer");
  REQUIRE(iteration >= 0, "mcts: negative iteration");
  REQUIRE(!exclude || trees->n_comms == 0 || comm_u, "mcts: communication needs comm_u");
  CK(casim_launch_mcts_select(h, trees, draws, (unsigned long long)draw_seed, iteration, comm_u, actions, exclude, root_only,
                              (cudaStream_t)stream));
  return 0;
}

int casim_mcts_update(casim_handle* h, const casim_mcts_trees* trees, const uint8_t* actions, const float* rewards,
                      const int32_t* n_feasible, double* avg_out, int expand_only, void* stream) {
  int rc = mcts_check(h, trees);
  if (rc) return rc;
  REQUIRE(actions && rewards && n_feasible, "mcts: null rollout result");
  REQUIRE(expand_only || trees->nsims >= 1, "mcts: statistics need at least one simulation");
  CK(casim_launch_mcts_update(h, trees, actions, rewards, n_feasible, avg_out, expand_only, trees->overflow,
                              (cudaStream_t)stream));
  return 0;
}

int casim_ga3c_set_weights(casim_handle* h, const float* weights_host, int n_params, void* stream) {
  REQUIRE(h && weights_host, "null argument");
  REQUIRE(n_params == CASIM_GA3C_NUM_PARAMS, "GA3C-CADRL parameter blob must hold 170764 floats");
  CK(casim_ga3c_upload(h, weights_host, n_params, (cudaStream_t)stream));
  return 0;
}

int casim_ga3c_forward(casim_handle* h, const float* state, float* policy_actions, int32_t* action_idx, float* probs,
                       float* nn_input, int use_tensor_cores, void* stream) {
  REQUIRE(h && state && policy_actions, "null argument");
  REQUIRE(h->ga3c_w, "casim_ga3c_set_weights has not been called");
  CK(casim_launch_ga3c(h, state, policy_actions, action_idx, probs, nn_input, use_tensor_cores,
                       (cudaStream_t)stream));
  return 0;
}

int casim_flat_row_size(const casim_handle* h, const int32_t* keys, int n_keys) {
  if (!h || !keys || n_keys < 0) return -1;
  return casim_flat_row_size_impl(h, keys, n_keys);
}

int casim_flatten_obs(casim_handle* h, const float* state, const casim_obs* obs, const int32_t* keys, int n_keys,
                      float* out, void* stream) {
  REQUIRE(h && state && obs && keys && out, "null argument");
  REQUIRE(n_keys > 0 && n_keys <= 16, "between 1 and 16 observation keys");
  REQUIRE(casim_flat_row_size_impl(h, keys, n_keys) > 0, "unknown observation key");
  for (int i = 0; i < n_keys; i++) {
    const int k = keys[i];
    if (k == CASIM_KEY_REL_GOAL || k == CASIM_KEY_NUM_OTHER_AGENTS || k == CASIM_KEY_USE_PPO)
      REQUIRE(obs->obs_misc, "flatten: obs_misc buffer missing");
    if (k == CASIM_KEY_OTHER_AGENT_STATES || k == CASIM_KEY_OTHER_AGENTS_STATES)
      REQUIRE(obs->other_agents_states, "flatten: other_agents_states buffer missing");
    if (k == CASIM_KEY_LASERSCAN) REQUIRE(obs->laserscan, "flatten: laserscan buffer missing");
    if (k == CASIM_KEY_LOCAL_GRID) REQUIRE(obs->local_grid, "flatten: local_grid buffer missing");
  }
  CK(casim_launch_flatten(h, state, obs, keys, n_keys, out, (cudaStream_t)stream));
  return 0;
}

void casim_rvo_default_params(casim_rvo_params* p) {
  if (!p) return;
  memset(p, 0, sizeof(*p));
  p->neighbor_dist = INFINITY;      /* Config.SENSING_HORIZON, config.py:63 */
  p->time_horizon = 5.0;            /* Config.RVO_TIME_HORIZON, config.py:66 */
  p->radius_scale = 1 + 15e-2;      /* RVOPolicy.py:78 */
  p->max_delta_heading = M_PI / 6;  /* RVOPolicy.py:20 */
  p->max_neighbors = 10;            /* Config.MAX_NUM_AGENTS_IN_ENVIRONMENT, config.py:70 */
}

int casim_rvo_set_obstacles(casim_handle* h, const double* vertices, const int32_t* poly_sizes, int n_polys) {
  REQUIRE(h, "null argument");
  casim_rvo_obstacles_free(h);
  if (n_polys <= 0) return 0;
  REQUIRE(vertices && poly_sizes, "null argument");
  CK(cudaDeviceSynchronize()); /* the previous set may still be read by a kernel in flight */
  int rc = casim_rvo_obstacles_build(h, vertices, poly_sizes, n_polys);
  REQUIRE(rc != -1, "casim_rvo_set_obstacles: a polygon needs at least 2 vertices");
  REQUIRE(rc != -2, "casim_rvo_set_obstacles: more than 64 obstacle vertices after edge splitting");
  REQUIRE(rc == 0, "casim_rvo_set_obstacles: CUDA allocation / upload failed");
  return 0;
}

int casim_rvo_step(casim_handle* h, const float* state, const float* coop, const casim_rvo_params* params,
                   float* policy_actions, void* stream) {
  REQUIRE(h && state && policy_actions, "null argument");
  REQUIRE(h->n_maps == 0 || h->rvo_n > 0,
          "casim_rvo_step: the handle has a static map but no RVO obstacle polygons; call casim_rvo_set_obstacles");
  casim_rvo_params def;
  if (!params) {
    casim_rvo_default_params(&def);
    params = &def;
  }
  REQUIRE(params->max_neighbors >= 0 && params->time_horizon > 0, "bad casim_rvo_params");
  CK(casim_launch_rvo(h, state, coop, params, policy_actions, (cudaStream_t)stream));
  return 0;
}

}  /* extern "C" */

int casim_pdl_enabled(void) {
  static const int on = casim_tuning_int("CASIM_PDL", 1) != 0;
  return on;
}

extern "C" {

unsigned long long casim_launch_count(const casim_handle* h) { return h ? h->launches : 0ull; }

} /* extern "C" */
