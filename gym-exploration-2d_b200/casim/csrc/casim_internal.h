/*
 * casim_internal.h -- shared between the translation units of libcasim.so (not part of the ABI).
 *
 * Every kernel receives the derived configuration `DevCfg` BY VALUE (kernel parameter space =
 * constant bank), so no __constant__ symbol has to be shared across translation units and no
 * relocatable device code is needed.
 */
#ifndef CASIM_INTERNAL_H
#define CASIM_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../../include/casim.h"

#define CASIM_NUM_SMS_FALLBACK 148
#define CASIM_PLANE_WORDS (CASIM_MAP_DIM * CASIM_MAP_WORDS) /* 3000 words = 12 000 B per bit plane */
#define CASIM_NUM_PLANES 4 /* S, S|right, S|down, S|right|down|diag  (OccupancyGrid bilinear footprint) */

/* OccupancyGrid fast path (casim_sensors.cu k_grid): the four footprint planes again, zero-padded by
 * CASIM_GPAD cells on every side (96 = 3 words, so padded word w is map word w-3), plus a 2-bit classification per
 * anchor cell: bit 0 = no occupied cell in the 6x6 box anchored there, bit 1 = all 36 cells occupied and in-map. */
#define CASIM_GPAD 96
#define CASIM_GROWS (CASIM_MAP_DIM + 2 * CASIM_GPAD)  /* 492 padded rows */
#define CASIM_GWORDS 16                                /* 512 padded columns */
#define CASIM_GCLS_WORDS 32                            /* 2 bits x 512 anchors per row */
#define CASIM_GBOX 6
#define CASIM_GPLANE_WORDS (CASIM_GROWS * CASIM_GWORDS)
#define CASIM_GMAP_WORDS (CASIM_NUM_PLANES * CASIM_GPLANE_WORDS + CASIM_GROWS * CASIM_GCLS_WORDS)

/* OccupancyGrid shared-map path (casim_sensors.cu k_grid_tma): ONE map padded by CASIM_G2_PAD cells on every side, laid
 * out so that a persistent CTA stages all of it into shared memory with TMA tile loads:
 *   4 footprint bit planes [384 rows][12 words]  (73 728 B)
 *   1 classification byte per EVEN anchor [192][192] (36 864 B): what the 8x8 box of plane 0 anchored there holds --
 *     bits 0-2: 0 other, 1 all free, 2 all occupied, 3 columns >= p occupied, 4 columns < p occupied, 5 rows >= p occupied,
 *     6 rows < p occupied (every row / column of the box identical); bits 3-5: p (1..7, relative to the anchor).
 * Seen as a 2-D tensor of 48-byte rows (12 uint32) it has 2304 rows; the kernel loads it as nine 256-row boxes. */
#define CASIM_G2_PAD 42
#define CASIM_G2_DIM (CASIM_MAP_DIM + 2 * CASIM_G2_PAD) /* 384 */
#define CASIM_G2_WORDS 12
#define CASIM_G2_PLANE_WORDS (CASIM_G2_DIM * CASIM_G2_WORDS)
#define CASIM_G2_CLS_DIM (CASIM_G2_DIM / 2)
#define CASIM_G2_CLS_BYTES (CASIM_G2_CLS_DIM * CASIM_G2_CLS_DIM)
#define CASIM_G2_BYTES (CASIM_NUM_PLANES * CASIM_G2_PLANE_WORDS * 4 + CASIM_G2_CLS_BYTES) /* 110 592 */
#define CASIM_G2_ROW_BYTES (CASIM_G2_WORDS * 4)
#define CASIM_G2_ROWS (CASIM_G2_BYTES / CASIM_G2_ROW_BYTES) /* 2304 */
#define CASIM_G2_BOX_ROWS 256
#define CASIM_G2_BOXES (CASIM_G2_ROWS / CASIM_G2_BOX_ROWS) /* 9 */

/* Derived, device-side view of casim_config (thresholds stay doubles: predicates are fp64). */
struct DevCfg {
  double dt, dt_nominal;
  double r_at_goal, r_coll_agent, r_timeout, r_infeasible, r_coll_wall, r_time_step, r_dist_goal;
  double r_wiggly, wiggly_thr, r_min, r_max, r_inv_span;
  double close_range, near_goal, max_time_ratio, sensing_horizon, max_heading_change;
  double map_ox, map_oy; /* (width/2)/cell, (height/2)/cell : Map.py:31-32 */
  double cell, inv_cell;
  double laser_max_range, laser_norm;
  double laser_cos[CASIM_MAX_BEAMS], laser_sin[CASIM_MAX_BEAMS], laser_ang[CASIM_MAX_BEAMS];
  double laser_r[CASIM_MAX_RANGES];
  float laser_out[CASIM_MAX_RANGES + 1]; /* value reported for "range = sample k" (k == nr slot kept at index CASIM_MAX_RANGES: no hit) */
  /* exploration */
  double ig_fov_half, ig_detect_fov_half, ig_range, ig_cell, ig_rocc, ig_remp, ig_tol, ig_pfn, ig_pfp, ig_dt;
  double ig_det_thresh; /* sqrt(0.5)*cell + tolerance */
  double ig_fov, ig_tan_half, ig_prior; /* full FOV (radians), tan(FOV/2) (0 when FOV/2 >= pi/2), prior odds */
  double map_hw, map_hh; /* map_width/2, map_height/2 (metres) */
  int nb, nr;
  int coll_w_static, max_obs, rows, submap_w, submap_h, game_over_mode, auto_reset, ig_dim;
  int laser_sorted; /* laser_r strictly increasing (np.arange): the ego-disc test is confined to a prefix of the samples */
};

struct casim_handle {
  casim_config cfg;
  DevCfg dc;
  int E, A;
  int device;
  int num_sms;
  /* maps */
  int n_maps;              /* 0 = no obstacles */
  uint32_t* map_planes;    /* [n_maps_alloc][4][300][10] bit planes (plane 0 = static_map) */
  double* edf;             /* [n_maps_alloc][300][300] edfMap.map itself: sqrt(edf_d2) * cell in fp64 */
  int32_t* edf_d2;         /* [n_maps_alloc][300][300] squared distance (cells^2) to the nearest occupied cell;
                              edfMap.map[r][c] == sqrt(edf_d2) * cell exactly (information_models/edfMap.py:11-12) */
  int32_t* edf_scratch;    /* [n_maps_alloc][300][300] column pass */
  uint32_t* grid_maps;     /* [n_maps_alloc][CASIM_GMAP_WORDS]: 4 padded planes then the classification plane */
  int n_maps_alloc;
  uint32_t* grid2_map;     /* [CASIM_G2_BYTES] shared-map layout of map 0 for k_grid_tma (built when n_maps == 1) */
  alignas(64) unsigned char grid2_tmap[128]; /* CUtensorMap over grid2_map (driver-encoded, passed as a kernel parameter) */
  int grid2_ready;         /* tensor map encoded for the current grid2_map */
  unsigned int* grid_work; /* [2] self-resetting work counter + finished-warp counter of the persistent k_grid_tma */
  uint8_t* grid_packed;    /* caller's optional bit-packed local_grid buffer (casim_set_grid_packed_output) or NULL */
  const long long* mi_table; /* caller's optional tabulated cell_mi of the belief (casim_explore_set_mi_table) or NULL */
  const int32_t* rollout_slots; /* caller's optional list of agent slots for casim_explore_rollouts (device) or NULL */
  int n_rollout_slots;
  /* borrowed from the caller (casim_reset / casim_set_trajectories) */
  const float* init;
  const int32_t* type0;
  const int32_t* n_agents;
  const float* traj;
  int T;
  /* scratch */
  double* pose64;          /* [E*A][4] post-move px, py, heading, radius in fp64 for the sensor kernels */
  uint8_t* prev_game_over; /* [E] */
  float* state0;           /* [6][E*A] float4: the episode-start state written by casim_reset; auto_reset re-loads it
                              instead of re-running Agent.__init__ (fp64 atan2/sqrt/div) inside the step kernel */
  double* partials;        /* [max_ctas][8] */
  unsigned int* ticket;    /* last-block counter */
  int max_ctas;
  /* GA3C */
  float* ga3c_w;           /* fp32 parameter blob on the device */
  void* ga3c_wb;           /* bf16 tiles for the tensor-core path */
  float* ga3c_x;           /* [E*A][76] network input scratch */
  int32_t* ga3c_fix;       /* [1 + E*A] counter + list of the agents the fp32 kernel re-evaluates after the bf16 forward */
  void* ga3c_xb;           /* [tiles][10][128] x 16 B: bf16 LSTM input chunks for the tensor-core path */
  /* RVO obstacle set (casim_rvo_set_obstacles): processed vertices + obstacle tree, see casim_rvo.cu */
  float2 *rvo_point, *rvo_unitdir;
  int *rvo_next, *rvo_prev, *rvo_convex;
  int4* rvo_node;
  int rvo_n, rvo_root;
  float* ext_copy_dst;     /* set by casim_step_host around its casim_step call (see CoreArgs::ext_copy_dst) */
  /* casim_step_host_pipelined: copy stream + per-slot events (kernels done / copies done), steps issued so far */
  cudaStream_t pipe_stream;
  cudaEvent_t pipe_ev_kernels[2], pipe_ev_copied[2];
  unsigned long long pipe_steps;
  cudaEvent_t stats_event; /* casim_set_stats_event: recorded right after the step kernel (the statistics are final then) */
  unsigned long long launches;
};

/* ---- launchers (host functions defined in the .cu files) ---- */
enum { CASIM_MODE_STEP = 0, CASIM_MODE_RESET = 1, CASIM_MODE_OBSERVE = 2, CASIM_MODE_STEP_N = 3 };

cudaError_t casim_launch_core(casim_handle* h, int mode, float* state, const float* ext_actions,
                              const float* policy_actions, double dt, const casim_obs* obs, cudaStream_t s,
                              const uint8_t* reset_mask = nullptr, int n_steps = 1);
cudaError_t casim_launch_laser(casim_handle* h, const float* state, float* laserscan, cudaStream_t s);
cudaError_t casim_launch_grid(casim_handle* h, const float* state, uint8_t* local_grid, cudaStream_t s,
                              float* fused_laser = nullptr);
bool casim_grid_can_fuse_laser(const casim_handle* h, const uint8_t* local_grid);
cudaError_t casim_launch_raster(casim_handle* h, const float* rects, const int32_t* n_rects, int n_maps,
                                int max_rects, cudaStream_t s);
cudaError_t casim_launch_pack_map(casim_handle* h, const uint8_t* static_map, int n_maps, cudaStream_t s);
cudaError_t casim_launch_map_derive(casim_handle* h, int n_maps, uint8_t* map_out, cudaStream_t s);
cudaError_t casim_grid2_encode_tmap(casim_handle* h);
cudaError_t casim_launch_explore(casim_handle* h, const float* state, const float* oas, float* belief,
                                 float* policy_actions, float* best_mi, float* cand_mi,
                                 unsigned long long* counters, cudaStream_t s);
cudaError_t casim_launch_scenario(casim_handle* h, int kind, int number_of_agents, const uint32_t* seeds, int ego_policy,
                                  int ego_dynamics, int other_dynamics, float* init, int32_t* types, int32_t* n_agents,
                                  float* coop, cudaStream_t s);
cudaError_t casim_launch_team_update(casim_handle* h, const float* state, const float* oas, float* belief,
                                     const double* targets_prev, const int32_t* n_prev, double* targets_new,
                                     int32_t* n_new, uint32_t* team_mask, float* team_reward, cudaStream_t s);
cudaError_t casim_launch_mi_table(casim_handle* h, const float* state, const float* belief, long long* table, cudaStream_t s);
cudaError_t casim_launch_rollouts(casim_handle* h, const float* state, const float* belief, const uint8_t* actions, int K,
                                  int H, int xdt, const uint32_t* exclude, int exclude_per_rollout, float* rewards,
                                  float* end_pose, uint32_t* obsvd, int32_t* n_feasible, cudaStream_t s);
cudaError_t casim_launch_ga3c(casim_handle* h, const float* state, float* policy_actions, int32_t* action_idx,
                              float* probs, float* nn_input, int use_tc, cudaStream_t s);
cudaError_t casim_launch_rvo(casim_handle* h, const float* state, const float* coop, const casim_rvo_params* prm,
                             float* policy_actions, cudaStream_t s);
int casim_rvo_obstacles_build(casim_handle* h, const double* verts, const int32_t* poly_sizes, int n_polys);
void casim_rvo_obstacles_free(casim_handle* h);
cudaError_t casim_launch_flatten(casim_handle* h, const float* state, const casim_obs* obs, const int32_t* keys,
                                 int n_keys, float* out, cudaStream_t s);
int casim_flat_row_size_impl(const casim_handle* h, const int32_t* keys, int n_keys);
cudaError_t casim_ga3c_upload(casim_handle* h, const float* weights_host, int n_params, cudaStream_t s);

cudaError_t casim_launch_mcts_select(casim_handle* h, const casim_mcts_trees* t, const uint8_t* draws,
                                     unsigned long long draw_seed, int iteration, const double* comm_u, uint8_t* actions,
                                     uint32_t* exclude, int root_only, cudaStream_t s);
cudaError_t casim_launch_mcts_update(casim_handle* h, const casim_mcts_trees* t, const uint8_t* actions, const float* rewards,
                                     const int32_t* n_feasible, double* avg_out, int expand_only, int* overflow,
                                     cudaStream_t s);
void casim_set_error(const char* fmt, ...);

/* Tuning knobs (launch shapes, zero-copy, PDL, trace dumps) are read from the environment ONLY in a -DCASIM_TUNING
 * build (make TUNING=1, used by the sweeps under profiles/); the release library has them compiled out, so that no
 * environment variable can silently change what the product path launches. */
#ifdef CASIM_TUNING
#include <stdlib.h>
static inline int casim_tuning_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}
static inline const char* casim_tuning_str(const char* name) { return getenv(name); }
#else
static inline int casim_tuning_int(const char*, int dflt) { return dflt; }
static inline const char* casim_tuning_str(const char*) { return nullptr; }
#endif

/* Programmatic dependent launch (PDL): the per-step kernels run back to back on one stream and each is short
 * (a single wave at the named batch sizes), so the launch latency between dependent kernels is a large share of a
 * step.  Launched with the programmatic-stream-serialization attribute, the next kernel is scheduled while the
 * previous one drains; every such kernel executes casim_pdl_enter() before its first global-memory access
 * (griddepcontrol.wait: all memory of the preceding kernel is visible afterwards), so the data dependence is kept.
 * Works under stream capture (programmatic edges in the CUDA graph).  CASIM_PDL=0 turns it off. */
int casim_pdl_enabled(void);

#ifdef __CUDACC__
__device__ __forceinline__ void casim_pdl_enter() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

template <typename Arg>
static inline cudaError_t casim_launch_pdl(void (*kern)(const Arg), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                           const Arg& arg) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = casim_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, arg);
}
#endif

#endif
