/*
 * casim_explore.cu -- the exploration policy step: ig_greedy.find_next_action for every ig_greedy agent
 * (citations relative to gym_collision_avoidance/envs/ of mlodel/gym-exploration-2d):
 *   policies/ig_greedy.py:45-62     find_targets_in_obs  (detector emulation from other_agents_states rows)
 *   information_models/targetMap.py:43-84   getVisibleCells (bounding box, range / FOV wedge, line of sight)
 *   information_models/edfMap.py:21-47      checkVisibility (sphere tracing on the Euclidean distance field)
 *   information_models/targetMap.py:86-128  update          (odds *= rOcc / rEmp on the visible cells)
 *   information_models/targetMap.py:130-147 get_reward_from_cells / get_reward_from_pose (mutual information)
 *   policies/ig_greedy.py:64-93     greedy_action / get_next_pose (9 candidates, EDF feasibility, strict-> argmax)
 *
 * One CTA per agent slot (slots that are not ig_greedy agents exit at once).  Ten poses are evaluated: pose 0
 * = the agent's pose (belief update), poses 1..9 = the candidate next poses (scoring on the UPDATED belief).
 * Two phases per batch of poses, both atomic-free on the data path (each belief cell is owned by one thread):
 *   phase 1  one thread per cell of the pose's bounding box: ego-frame rotation, range and FOV tests; cells
 *            inside the wedge are appended to a shared-memory queue (keeps phase 2 balanced: ~13 % of the box
 *            survives);
 *   phase 2  one thread per queued cell: sphere tracing on the exact squared EDT, then the odds multiply
 *            (pose 0) or the cell's mutual information (poses 1..9), accumulated per candidate as a 64-bit
 *            fixed-point integer so that the sum -- hence the argmax and its ties -- does not depend on the order
 *            in which threads arrive.
 *
 * Arithmetic: fp64 like the reference for every decision (cell indices, range / FOV / visibility, feasibility,
 * detection radius); the two wedge predicates are decided from squared / cross-multiplied forms with a
 * relative guard band of 1e-12 and fall back to the reference's sqrt / atan2 inside the band.  The belief is
 * stored as fp32 odds (clamped to FLT_MAX); MI is evaluated in fp64.
 */
#include <float.h>

#include "casim_device.cuh"

namespace {

#define XQ_CAP 4096        /* queue entries: >= ig_dim^2 (ig_dim <= 64), so one pose always fits */
#ifndef X_THREADS
#define X_THREADS 128 /* C4 sweep of the three-phase kernel (profiles/r2e_sweep_explore_threads.txt): 96: 83.9 us, 128: 69.8,
                        * 192: 83.1 (spills at 64 registers), 256: 90.0.  k_team_update shares the value. */
#endif
#define X_NPOSE 10
#ifndef X_MIN_CTAS
#define X_MIN_CTAS (1024 / X_THREADS) /* 64 registers per thread */
#endif
#define X_MI_SCALE 17592186044416.0 /* 2^44 */

struct ExploreArgs {
  DevCfg c;
  const float4* state;
  const float* oas;          /* [E*A][rows][10] */
  float* belief;             /* [E*A][dim][dim] */
  float* policy_actions;     /* [E*A][2] */
  float* best_mi;            /* [E*A] or NULL */
  float* cand_mi;            /* [E*A][9] or NULL */
  unsigned long long* counters; /* [4] or NULL */
  const int32_t* n_agents;
  const double* edf;
  const float* init;
  const uint8_t* prev_game_over;
  int E, A, n_maps;
};

struct PoseBox {
  double x, y, cs, sn;
  int x0, y0, w, h;       /* box origin and extent in target-map cells (w*h cells) */
  int feasible;
  int inv_h;              /* ceil(2^18 / h): (off * inv_h) >> 18 == off / h for off < 4096, h <= 64 */
};

/* edfMap.map[yIdc, xIdc] (edfMap.py:14-19,38-40; no y flip, sic); indices clamped into the array */
__device__ __forceinline__ double edf_value(const DevCfg& c, const double* __restrict__ d2, double px, double py,
                                            unsigned& clamped) {
  /* floor((p + half) / cell): the product with 1/cell differs from the quotient by at most an ulp, so its floor is the
   * quotient's unless it lands within 1e-9 of an integer; only then is the division itself evaluated */
  double tx = (px + c.map_hw) * c.inv_cell, ty = (py + c.map_hh) * c.inv_cell;
  double fx = floor(tx), fy = floor(ty);
  const double rx = tx - fx, ry = ty - fy; /* exact (tx and its floor are within a factor of two, or the result is < 1) */
  if (!(rx > 1e-9 && rx < 1.0 - 1e-9)) fx = floor((px + c.map_hw) / c.cell);
  if (!(ry > 1e-9 && ry < 1.0 - 1e-9)) fy = floor((py + c.map_hh) / c.cell);
  /* double -> int saturates and maps NaN to 0: the clamp of edfMap.py:16-19 in two integer min/max per axis */
  const int xr = __double2int_rz(fx), yr = __double2int_rz(fy);
  const int xi = min(max(xr, 0), CASIM_MAP_DIM - 1), yi = min(max(yr, 0), CASIM_MAP_DIM - 1);
  clamped += (unsigned)(xi != xr) + (unsigned)(yi != yr);
  return __ldg(d2 + yi * CASIM_MAP_DIM + xi);
}

/* edfMap.py:21-47 checkVisibility */
__device__ __forceinline__ bool line_of_sight(const DevCfg& c, const double* __restrict__ d2, double px, double py,
                                              double gx, double gy, unsigned& samples, unsigned& clamped) {
  double dx = gx - px, dy = gy - py;
  double L = sqrt(dx * dx + dy * dy);
  double u = 0.05 / L;
  while (u < 1.0) {
    double nx = (1.0 - u) * px + u * gx, ny = (1.0 - u) * py + u * gy;
    double md = edf_value(c, d2, nx, ny, clamped);
    samples++;
    if (md < 0.001) return false;
    u += md / L;
  }
  return true;
}

/* targetMap.py:33-36 getPoseFromCell (one axis) */
__device__ __forceinline__ double cell_centre(const DevCfg& c, int i, double half) {
  return (double)i * c.ig_cell - half + c.ig_cell / 2;
}
__device__ __forceinline__ int cell_index(const DevCfg& c, double v, double half) {
  double f = floor((v + half) / c.ig_cell);
  f = fmin(fmax(f, -1.0e6), 1.0e6);
  return (int)f;
}
__device__ __forceinline__ double lim_axis(double v, double half) { return fmax(fmin(v, half), -half); }

/* range and FOV wedge test of targetMap.py:75-79: sqrt(r0^2+r1^2) < R and |atan2(r1, r0)| < F/2 */
__device__ __forceinline__ bool in_wedge(const DevCfg& c, double r0, double r1) {
  const double R = c.ig_range;
  double s = r0 * r0 + r1 * r1, R2 = R * R;
  bool in_range;
  if (s < R2 * (1.0 - 1e-12)) in_range = true;
  else if (s > R2 * (1.0 + 1e-12)) in_range = false;
  else in_range = sqrt(s) < R;
  if (!in_range) return false;
  if (c.ig_tan_half > 0.0) {
    if (r0 <= 0.0) return false; /* |dphi| >= pi/2 > F/2 */
    double a = fabs(r1), b = c.ig_tan_half * r0;
    double t = a - b;
    if (fabs(t) > 1e-12 * (a + b)) return t < 0.0;
  }
  return fabs(atan2(r1, r0)) < c.ig_fov_half;
}

/* targetMap.py:130-143, one cell */
__device__ __forceinline__ double cell_mi(const DevCfg& c, double r, double log_rocc, double log_remp) {
  double p = r / (r + 1.0);
  double f_p = log((r + 1.0) / (r + (1.0 / c.ig_rocc))) - log_rocc / (r * c.ig_rocc + 1.0);
  double f_n = log((r + 1.0) / (r + (1.0 / c.ig_remp))) - log_remp / (r * c.ig_remp + 1.0);
  double P_p = p * (1.0 - c.ig_pfn) + (1.0 - p) * c.ig_pfp;
  double P_n = p * c.ig_pfn + (1.0 - p) * (1.0 - c.ig_pfp);
  return P_p * f_p + P_n * f_n;
}

/* bounding box of targetMap.py:46-66 for one pose: cells of {pose, centre, left, right}, end-exclusive */
__device__ __forceinline__ void pose_box(const DevCfg& c, double x, double y, double phi, PoseBox& b, unsigned& clamped) {
  const double R = c.ig_range, F = c.ig_fov;
  double s0, c0, s1, c1, s2, c2;
  sincos(phi, &s0, &c0);
  sincos(phi + F, &s1, &c1);
  sincos(phi - F, &s2, &c2);
  double xs[4] = {x, lim_axis(x + R * c0, c.map_hw), lim_axis(x + R * c1, c.map_hw), lim_axis(x + R * c2, c.map_hw)};
  double ys[4] = {y, lim_axis(y + R * s0, c.map_hh), lim_axis(y + R * s1, c.map_hh), lim_axis(y + R * s2, c.map_hh)};
  int x0 = INT_MAX, x1 = INT_MIN, y0 = INT_MAX, y1 = INT_MIN;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    int cx = cell_index(c, xs[k], c.map_hw), cy = cell_index(c, ys[k], c.map_hh);
    x0 = min(x0, cx); x1 = max(x1, cx);
    y0 = min(y0, cy); y1 = max(y1, cy);
  }
  const int dim = c.ig_dim;
  if (x0 < 0) { x0 = 0; clamped++; }
  if (y0 < 0) { y0 = 0; clamped++; }
  if (x1 > dim) { x1 = dim; clamped++; }
  if (y1 > dim) { y1 = dim; clamped++; }
  b.x = x; b.y = y; b.cs = c0; b.sn = s0;
  b.x0 = x0; b.y0 = y0;
  b.w = max(x1 - x0, 0);
  b.h = max(y1 - y0, 0);
  b.inv_h = b.h > 0 ? ((1 << 18) + b.h - 1) / b.h : 0;
}

/* One CTA per explorer.  COUNT: instrumentation counters (cells tested / EDF samples / visible cells / clamped indices)
 * compiled in only when the caller passes a counters buffer. */
template <bool COUNT>
__global__ void __launch_bounds__(X_THREADS, X_MIN_CTAS) k_explore(const ExploreArgs p) {
  const DevCfg& c = p.c;
  const int g = blockIdx.x;
  const int e = g / p.A, ai = g - e * p.A;
  const int n = p.n_agents ? min(p.n_agents[e], p.A) : p.A;
  if (ai >= n) return;
  const size_t N = (size_t)p.E * p.A;
  const float4 p5 = p.state[5 * N + g];
  const int type = __float_as_int(p5.w), flags = __float_as_int(p5.z);
  if ((type & 0xff) != CASIM_POLICY_IG_GREEDY) return;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const bool restart = c.auto_reset && p.prev_game_over && p.init && p.prev_game_over[e];
  if (!restart && (flags & CASIM_F_IS_DONE)) { /* collision_avoidance_env.py:298-300: no policy query */
    if (tid == 0) {
      p.policy_actions[2 * g] = 0.f;
      p.policy_actions[2 * g + 1] = 0.f;
      if (p.best_mi) p.best_mi[g] = -1.f;
    }
    if (p.cand_mi && tid < 9) p.cand_mi[9 * (size_t)g + tid] = -1.f;
    return;
  }

  __shared__ PoseBox s_pose[X_NPOSE];
  __shared__ double s_det[CASIM_MAX_AGENTS][2];
  __shared__ double s_log[2];                        /* log(ig_rocc), log(ig_remp) */
  __shared__ double s_cc[2][64];                     /* getPoseFromCell per axis (targetMap.py:33-36): x of column i, y of row j */
  __shared__ int s_ndet;
  __shared__ int s_off[X_NPOSE + 1];                 /* prefix of the box sizes: cells of pose k are [s_off[k], s_off[k+1]); pose 2: none */
  __shared__ unsigned int s_wq[X_THREADS / 32][64];  /* per-warp queue: wedge survivors (pose << 12) | (j << 6) | i, then seen cells */
  __shared__ long long s_miw[X_THREADS / 32][X_NPOSE]; /* per-warp fixed-point mutual-information sums */
  __shared__ unsigned int s_vis[64 * 64];            /* per target-map cell: bit 16+k = inside pose k's wedge, bit k = visible from candidate pose k */
  __shared__ unsigned int s_cnt[4];

  const int dim = c.ig_dim;
  float* belief = p.belief + (size_t)g * dim * dim;
  const double* d2 = p.edf + (size_t)(p.n_maps == 1 ? 0 : e) * CASIM_MAP_DIM * CASIM_MAP_DIM;
  unsigned cnt_tested = 0, cnt_samples = 0, cnt_visible = 0, cnt_clamped = 0;

  /* the explorer's pose: every thread reads it itself (uniform loads), nobody waits for a broadcast through shared memory */
  double px, py, ph, prad;
  if (restart) { /* agent.py:21-33 with the initial condition kept by casim_reset */
    const float* in = p.init + (size_t)g * 7;
    px = in[0]; py = in[1]; prad = in[4];
    ph = isnan(in[6]) ? atan2((double)in[3] - (double)in[1], (double)in[2] - (double)in[0]) : (double)in[6];
  } else {
    const float4 p0 = p.state[g], p2 = p.state[2 * N + g];
    px = p0.x; py = p0.y; ph = p0.z; prad = p2.z;
  }
  for (int k = tid; k < 64 * 64; k += X_THREADS) s_vis[k] = 0u;
  for (int k = tid; k < 128; k += X_THREADS) s_cc[k >> 6][k & 63] = cell_centre(c, k & 63, (k >> 6) ? c.map_hh : c.map_hw);
  if (COUNT && tid < 4) s_cnt[tid] = 0u;
  if (restart)
    for (int k = tid; k < dim * dim; k += X_THREADS) belief[k] = (float)c.ig_prior;

  if (wid == 0) {
    /* ---- detector emulation (ig_greedy.py:45-62) + ego-frame detections (targetMap.py:91-103) ---- */
    double rx = 0, ry = 0;
    bool hit = false;
    if (restart) {
      if (tid < n && tid != ai && (__float_as_int(p.state[5 * N + (size_t)e * p.A + tid].w) & 0xff) == CASIM_POLICY_STATIC) {
        const float* oin = p.init + ((size_t)e * p.A + tid) * 7;
        rx = (double)oin[0] - px;
        ry = (double)oin[1] - py;
        hit = true;
      }
    } else if (tid < c.rows) {
      const float* row = p.oas + ((size_t)g * c.rows + tid) * 10;
      if (row[9] == 1.0f) {
        rx = row[0];
        ry = row[1];
        hit = true;
      }
    }
    if (hit) {
      double dphi = atan2(ry, rx) - ph;
      bool in_fov = fabs(dphi) <= c.ig_detect_fov_half; /* degrees vs radians, sic */
      double rn = sqrt(rx * rx + ry * ry);
      hit = in_fov && rn <= c.ig_range;
    }
    unsigned m = __ballot_sync(FULL_MASK, hit);
    if (hit) {
      int k = __popc(m & ((1u << tid) - 1u));
      /* target = r + pose; detection = R_minus (target - pose) */
      double tx = rx + px, ty = ry + py;
      double dx = tx - px, dy = ty - py;
      double cs, sn;
      sincos(ph, &sn, &cs);
      s_det[k][0] = cs * dx + sn * dy;
      s_det[k][1] = -sn * dx + cs * dy;
    }
    if (tid == 0) s_ndet = __popc(m);
  } else if (wid == 1) {
    /* ---- the ten poses: 0 = current, 1..9 = get_next_pose(action a) (ig_greedy.py:81-93), lanes 0..9 ---- */
    const int k = lane;
    int cells = 0;
    if (k < X_NPOSE) {
      double x = px, y = py, h = ph;
      int feasible = 1;
      if (k > 0) {
        int a = k - 1;
        double v = 2.0 * (double)(a / 3), w = (a % 3 == 0) ? -CASIM_PI : ((a % 3 == 2) ? CASIM_PI : 0.0);
        double sn, cs;
        sincos(ph, &sn, &cs);
        double vx = cs * v + (-sn) * 0.0, vy = sn * v + cs * 0.0;
        x = px + vx * c.ig_dt;
        y = py + vy * c.ig_dt;
        h = ph + w * c.ig_dt;
        unsigned cl = 0;
        feasible = edf_value(c, d2, x, y, cl) > prad + 0.1;
        cnt_samples++;
        cnt_clamped += cl;
      }
      PoseBox b;
      pose_box(c, x, y, h, b, cnt_clamped);
      b.feasible = feasible;
      if (!feasible) b.w = b.h = 0;
      s_pose[k] = b;
      cells = k == 2 ? 0 : b.w * b.h; /* pose 2 rides on pose 0 (phase 1) */
    }
    int incl = cells; /* inclusive prefix over the warp */
#pragma unroll
    for (int d = 1; d < 16; d <<= 1) {
      const int up = __shfl_up_sync(FULL_MASK, incl, d);
      if (k >= d) incl += up;
    }
    if (k < X_NPOSE) s_off[k + 1] = incl;
    if (k == 0) s_off[0] = 0;
  }
  if (tid == X_THREADS - 1) s_log[0] = log(c.ig_rocc);
  if (tid == X_THREADS - 2) s_log[1] = log(c.ig_remp);
  __syncthreads();

  const int ndet = s_ndet;
  unsigned int* wq = s_wq[wid];

  /* ---- phase 1: range / FOV wedge test (targetMap.py:73-79) of every cell of every pose's box, one dense sweep over the
   * concatenated boxes; a survivor sets bit 16+k of its cell's word.  Pose 2 (action v = 0, w = 0) IS pose 0 -- same
   * position, heading and box, bit for bit -- so its tests are not repeated: its bit is set with pose 0's. ---- */
  {
    const int total = s_off[X_NPOSE];
    const unsigned twin = s_pose[2].feasible ? (1u << 18) : 0u;
    int k = 0;
    for (int idx = tid; idx < total; idx += X_THREADS) {
      while (idx >= s_off[k + 1]) k++; /* idx only grows: the pose cursor never rewinds */
      const int off = idx - s_off[k];
      const PoseBox& b = s_pose[k];
      const int ii = (off * b.inv_h) >> 18, jj = off - ii * b.h;
      const int i = b.x0 + ii, j = b.y0 + jj;
      const double dx = s_cc[0][i] - b.x, dy = s_cc[1][j] - b.y;
      const double r0 = b.cs * dx + b.sn * dy, r1 = -b.sn * dx + b.cs * dy; /* R = ((c, s), (-s, c)) */
      cnt_tested += (k == 0 && twin) ? 2u : 1u;
      if (in_wedge(c, r0, r1)) atomicOr(&s_vis[j * 64 + i], (1u << (16 + k)) | (k == 0 ? twin : 0u));
    }
  }
  __syncthreads();

  /* ---- phase 2: lines of sight.  The ten poses have only THREE distinct positions (speed 0 / 2 / 4 along the
   * heading; the turn rate changes the heading alone: ig_greedy.py:81-93) and edfMap.checkVisibility depends on the two
   * end points only, so one ray per (position, cell) serves every pose of that position whose wedge holds the cell:
   * groups {0,1,2,3}, {4,5,6}, {7,8,9}.  Cells are compacted into the warp's queue and traced 32 at a time. ---- */
  int ux0 = INT_MAX, uy0 = INT_MAX, ux1 = INT_MIN, uy1 = INT_MIN; /* union box of all poses */
  for (int kk = 0; kk < X_NPOSE; kk++) {
    const PoseBox& b = s_pose[kk];
    if (b.w <= 0 || b.h <= 0) continue;
    ux0 = min(ux0, b.x0); uy0 = min(uy0, b.y0);
    ux1 = max(ux1, b.x0 + b.w); uy1 = max(uy1, b.y0 + b.h);
  }
  const int uw = ux1 > ux0 ? ux1 - ux0 : 0, uh = uy1 > uy0 ? uy1 - uy0 : 0; /* (all boxes empty: nothing to do) */
  const int ncell = uw * uh;

  auto trace = [&](unsigned en) {
    const int grp = (int)(en >> 12), j = (int)((en >> 6) & 63u), i = (int)(en & 63u);
    const unsigned gmask = grp == 0 ? 0x00fu : (grp == 1 ? 0x070u : 0x380u);
    const PoseBox& b = s_pose[grp == 0 ? 0 : (grp == 1 ? 4 : 7)];
    const unsigned bits = (s_vis[j * 64 + i] >> 16) & gmask; /* the upper half is final since the barrier above */
    const double gx = s_cc[0][i], gy = s_cc[1][j];
    unsigned ns = 0;
    const bool vis = line_of_sight(c, d2, b.x, b.y, gx, gy, ns, cnt_clamped);
    if (COUNT) cnt_samples += ns * __popc(bits); /* counters report what the reference does: one ray per (pose, cell) */
    if (!vis) return;
    if (COUNT) cnt_visible += __popc(bits);
    if (bits & 1u) { /* targetMap.py:104-126; pose 0's cell is traced exactly once: no race on the belief */
      float* cellp = belief + j * dim + i;
      double rs = c.ig_remp;
      if (ndet > 0) {
        double dx = gx - b.x, dy = gy - b.y;
        double r0 = b.cs * dx + b.sn * dy, r1 = -b.sn * dx + b.cs * dy;
        for (int t = 0; t < ndet; t++) {
          double ex = s_det[t][0] - r0, ey = s_det[t][1] - r1;
          if (sqrt(ex * ex + ey * ey) < c.ig_det_thresh) {
            rs = c.ig_rocc;
            break;
          }
        }
      }
      *cellp = fminf((float)((double)*cellp * rs), FLT_MAX);
    }
    /* The candidate poses' wedges overlap heavily and a cell's mutual information depends on the cell alone:
     * record who sees it, evaluate the two fp64 logs once per cell below (not once per pose and cell). */
    if (bits & ~1u) atomicOr(&s_vis[j * 64 + i], bits & ~1u);
  };
  {
    int wqn = 0;
    for (int base = wid * 32; base < ncell; base += X_THREADS) { /* warp-uniform trip count */
      const int idx = base + lane;
      unsigned w = 0u, cell = 0u;
      if (idx < ncell) {
        const int ii = idx / uh, jj = idx - ii * uh;
        const int i = ux0 + ii, j = uy0 + jj;
        w = s_vis[j * 64 + i] >> 16;
        cell = ((unsigned)j << 6) | (unsigned)i;
      }
      if (__ballot_sync(FULL_MASK, w != 0u) == 0u) continue;
#pragma unroll
      for (int grp = 0; grp < 3; grp++) {
        const bool has = (w & (grp == 0 ? 0x00fu : (grp == 1 ? 0x070u : 0x380u))) != 0u;
        const unsigned m = __ballot_sync(FULL_MASK, has);
        if (m == 0u) continue;
        if (has) wq[wqn + __popc(m & ((1u << lane) - 1u))] = ((unsigned)grp << 12) | cell;
        wqn += __popc(m);
        __syncwarp();
        if (wqn >= 32) {
          wqn -= 32;
          const unsigned e2 = wq[wqn + lane];
          __syncwarp();
          trace(e2);
        }
      }
    }
    if (lane < wqn) trace(wq[lane]);
  }
  __syncthreads(); /* pose 0's odds update and every visibility bit are in place */

  /* ---- phase 3: targetMap.py:130-143 over the union box, on the UPDATED belief.  Seen cells are
   * compacted into the warp's queue first so that the two fp64 logs per cell run on full warps; the per-pose sums are
   * integer (fixed point, 2^-44): independent of the order of arrival, accumulated with warp reductions on three
   * 22-bit pieces (lane k of every warp carries pose k's sum), no shared-memory atomics. ---- */
  {
    const double log_rocc = s_log[0], log_remp = s_log[1];
    long long acc = 0;
    auto score = [&](bool valid, unsigned cell) {
      unsigned m = 0u;
      long long v = 0;
      if (valid) {
        const int j = (int)(cell >> 6), i = (int)(cell & 63u);
        m = s_vis[j * 64 + i] & 0x3feu;
        v = __double2ll_rn(cell_mi(c, (double)belief[j * dim + i], log_rocc, log_remp) * X_MI_SCALE);
      }
      const unsigned lo = (unsigned)(v & 0x3fffffll), mid = (unsigned)((v >> 22) & 0x3fffffll);
      const int hi = (int)(v >> 44); /* arithmetic shift: v == (hi << 44) + (mid << 22) + lo for negative v too */
      unsigned um = __reduce_or_sync(FULL_MASK, m);
      while (um) {
        const int kk = __ffs(um) - 1;
        um &= um - 1;
        const bool on = (m >> kk) & 1u;
        const unsigned a = __reduce_add_sync(FULL_MASK, on ? lo : 0u), b2 = __reduce_add_sync(FULL_MASK, on ? mid : 0u);
        const int c2 = __reduce_add_sync(FULL_MASK, on ? hi : 0);
        if (lane == kk)
          acc = (long long)((unsigned long long)acc + (unsigned long long)a + ((unsigned long long)b2 << 22) +
                            ((unsigned long long)(long long)c2 << 44));
      }
    };
    int qn = 0;
    for (int base = wid * 32; base < ncell; base += X_THREADS) { /* warp-uniform trip count */
      const int idx = base + lane;
      bool seen = false;
      unsigned cell = 0u;
      if (idx < ncell) {
        const int ii = idx / uh, jj = idx - ii * uh;
        const int i = ux0 + ii, j = uy0 + jj;
        seen = (s_vis[j * 64 + i] & 0x3feu) != 0u;
        cell = ((unsigned)j << 6) | (unsigned)i;
      }
      const unsigned m = __ballot_sync(FULL_MASK, seen);
      if (seen) wq[qn + __popc(m & ((1u << lane) - 1u))] = cell;
      qn += __popc(m);
      __syncwarp();
      if (qn >= 32) {
        qn -= 32;
        const unsigned c2 = wq[qn + lane];
        __syncwarp();
        score(true, c2);
      }
    }
    if (qn > 0) score(lane < qn, lane < qn ? wq[lane] : 0u);
    if (lane < X_NPOSE) s_miw[wid][lane] = acc;
  }
  __syncthreads();

  /* ---- greedy argmax, strict '>' in action order, starting from -1 (ig_greedy.py:64-78) ---- */
  if (tid == 0) {
    double max_mi = -1.0;
    float a0 = -1.f, a1 = -1.f; /* the integer -1 broadcast into all_actions[i, :] */
    for (int a = 0; a < 9; a++) {
      double mi = -1.0;
      if (s_pose[a + 1].feasible) {
        unsigned long long sum = 0ull;
        for (int w = 0; w < X_THREADS / 32; w++) sum += (unsigned long long)s_miw[w][a + 1];
        mi = (double)(long long)sum / X_MI_SCALE;
        if (mi > max_mi) {
          max_mi = mi;
          a0 = 2.0f * (float)(a / 3);
          a1 = (a % 3 == 0) ? (float)-CASIM_PI : ((a % 3 == 2) ? (float)CASIM_PI : 0.f);
        }
      }
      if (p.cand_mi) p.cand_mi[9 * (size_t)g + a] = (float)mi;
    }
    p.policy_actions[2 * g] = a0;
    p.policy_actions[2 * g + 1] = a1;
    if (p.best_mi) p.best_mi[g] = (float)max_mi;
  }
  if (COUNT) { /* instrumentation only: per-CTA totals, one global atomic per counter */
    unsigned v0 = __reduce_add_sync(FULL_MASK, cnt_tested), v1 = __reduce_add_sync(FULL_MASK, cnt_samples);
    unsigned v2 = __reduce_add_sync(FULL_MASK, cnt_visible), v3 = __reduce_add_sync(FULL_MASK, cnt_clamped);
    if ((tid & 31) == 0) {
      atomicAdd(&s_cnt[0], v0);
      atomicAdd(&s_cnt[1], v1);
      atomicAdd(&s_cnt[2], v2);
      atomicAdd(&s_cnt[3], v3);
    }
    __syncthreads();
    if (tid < 4) atomicAdd(p.counters + tid, (unsigned long long)s_cnt[tid]);
  }
}

/* ---------------------------------------------------------------------------------------------------------
 * Dec-MCTS simulation rollouts (policies/ig_mcts.py:173-198 get_next_pose, :223-244 mcts_sim_state_storer,
 * :246-253 mcts_reward; driven like pydecmcts/DecMCTS.py:233-269 _simulate).  One CTA per (agent, rollout):
 *   1. thread 0 replays the action sequence (the only serial part: each pose depends on whether the previous
 *      step was feasible) and writes one bounding box per feasible step;
 *   2. the boxes are processed like k_explore's candidate poses -- wedge test per box cell, survivors queued,
 *      sphere tracing per queued cell -- and every visible cell sets its bit in a 4096-bit shared-memory mask
 *      (obsvd_cells is a set: a cell seen at several steps counts once, and is traced once if a previous batch
 *      already saw it);
 *   3. the mask minus the other robots' cells is scored: fp64 mutual information per cell, 64-bit fixed-point
 *      sum (order-independent).
 * --------------------------------------------------------------------------------------------------------- */
#define XR_QCAP 2048         /* box cells per batch (queue entries are 16 bit: 4-bit pose, 6+6-bit cell) */
#define XR_MASK_WORDS (64 * 64 / 32)

struct RolloutArgs {
  DevCfg c;
  const float4* state;
  const float* belief;        /* [E*A][dim][dim] */
  const uint8_t* actions;     /* [E*A][K][H] */
  const uint32_t* exclude;    /* [E*A][128] or [E*A][K][128] or NULL; bit (j*64 + i) */
  float* rewards;             /* [E*A][K] */
  float* end_pose;            /* [E*A][K][3] or NULL */
  uint32_t* obsvd;            /* [E*A][K][128] or NULL */
  int32_t* n_feasible;        /* [E*A][K] or NULL: steps of the sequence that were applied */
  const int32_t* n_agents;
  const double* edf;
  const long long* mi_table;  /* [E*A][dim][dim] fixed-point cell_mi of the belief (casim_explore_mi_table) or NULL */
  const int32_t* slots;       /* [n_slots] agent slots to evaluate (casim_explore_set_rollout_slots) or NULL = every slot */
  int E, A, n_maps, K, H, xdt, exclude_per_rollout, qcap;
};

template <int XR_THREADS>
__global__ void __launch_bounds__(XR_THREADS) k_rollouts(const RolloutArgs p) {
  const DevCfg& c = p.c;
  /* one CTA per (agent slot, rollout); with a slot list only the listed slots get CTAs, results stay at [slot][k] */
  const int si = blockIdx.x / p.K;
  const int g = p.slots ? p.slots[si] : si;
  const int gk = g * p.K + (blockIdx.x - si * p.K);
  const int e = g / p.A, ai = g - e * p.A;
  const int n = p.n_agents ? min(p.n_agents[e], p.A) : p.A;
  const size_t N = (size_t)p.E * p.A;
  const int tid = threadIdx.x;
  const float4 p5 = p.state[5 * N + g];
  const bool explorer = ai < n && (__float_as_int(p5.w) & 0xff) == CASIM_POLICY_IG_GREEDY;
  if (!explorer) {
    if (tid == 0) {
      p.rewards[gk] = 0.f;
      if (p.n_feasible) p.n_feasible[gk] = 0;
      if (p.end_pose) {
        const float4 p0 = p.state[g];
        p.end_pose[3 * (size_t)gk] = p0.x; p.end_pose[3 * (size_t)gk + 1] = p0.y; p.end_pose[3 * (size_t)gk + 2] = p0.z;
      }
    }
    if (p.obsvd)
      for (int w = tid; w < XR_MASK_WORDS; w += XR_THREADS) p.obsvd[(size_t)gk * XR_MASK_WORDS + w] = 0u;
    return;
  }

  __shared__ PoseBox s_pose[CASIM_ROLLOUT_MAX_H];
  __shared__ unsigned short s_queue[XR_QCAP];
  __shared__ unsigned int s_seen[XR_MASK_WORDS];
  __shared__ unsigned long long s_sum;
  __shared__ int s_qn, s_nposes, s_probe;

  const int dim = c.ig_dim;
  const double* d2 = p.edf + (size_t)(p.n_maps == 1 ? 0 : e) * CASIM_MAP_DIM * CASIM_MAP_DIM;
  unsigned cnt_samples = 0, cnt_clamped = 0;
  for (int w = tid; w < XR_MASK_WORDS; w += XR_THREADS) s_seen[w] = 0u;
  if (p.xdt == 1 && tid < 32) {
    /* ---- the replay on one warp, lane t = step t.  The only serial dependence is feasibility (an infeasible step keeps
     * pose and heading), so the warp SPECULATES that every remaining step is feasible: headings and positions are the same
     * running sums the serial loop forms (each lane accumulates its own prefix in step order: identical operations,
     * identical bits), the fp64 sincos and the clearance look-ups of all steps run side by side, and the first infeasible
     * step ends the pass -- the steps before it are final, the walk resumes behind it from the last good pose.  One pass
     * per infeasible step instead of one sincos + one dependent EDF load per step. ---- */
    const int lane = tid;
    if (lane == 0) {
      s_qn = 0;
      s_sum = 0ull;
    }
    const float4 p0 = p.state[g], p2 = p.state[2 * N + g];
    double x = p0.x, y = p0.y, h = p0.z;
    const double rad = p2.z;
    const int a = lane < p.H ? (int)p.actions[(size_t)gk * p.H + lane] : 255;
    const unsigned endm = __ballot_sync(FULL_MASK, a > 8);
    const int len = endm ? __ffs(endm) - 1 : 32;
    const int a_end = __shfl_sync(FULL_MASK, a, min(len, 31));
    const double v = a <= 8 ? 2.0 * (double)(a / 3) : 0.0;
    const double w = a <= 8 ? ((a % 3 == 0) ? -0.5 * CASIM_PI : ((a % 3 == 2) ? 0.5 * CASIM_PI : 0.0)) : 0.0;
    int t0 = 0, np_ = 0;
    while (t0 < len) {
      double hh = h, xx = x, yy = y; /* the state before step `lane` if steps t0 .. lane-1 are all feasible */
      for (int k = t0; k < len; k++) {
        const double wk = __shfl_sync(FULL_MASK, w, k);
        if (k < lane) hh = hh + wk * c.ig_dt;
      }
      double sn, cs;
      sincos(hh, &sn, &cs);
      const double vx = cs * v + (-sn) * 0.0, vy = sn * v + cs * 0.0;
      for (int k = t0; k < len; k++) {
        const double vxk = __shfl_sync(FULL_MASK, vx, k), vyk = __shfl_sync(FULL_MASK, vy, k);
        if (k < lane) {
          xx = xx + vxk * c.ig_dt;
          yy = yy + vyk * c.ig_dt;
        }
      }
      const double nx = xx + vx * c.ig_dt, ny = yy + vy * c.ig_dt, nh = hh + w * c.ig_dt;
      const bool mine = lane >= t0 && lane < len;
      bool ok = true;
      if (mine && v != 0.0) {
        const bool in_map = c.map_hw > nx && c.map_hh > ny && nx > -c.map_hw && ny > -c.map_hh;
        ok = in_map && edf_value(c, d2, nx, ny, cnt_clamped) > rad + 0.1;
      }
      const unsigned bad = __ballot_sync(FULL_MASK, mine && !ok);
      const int f = bad ? __ffs(bad) - 1 : len; /* ig_mcts.py:237-243: at step f the pose stays, nothing is observed */
      if (lane >= t0 && lane < f) {
        PoseBox& q = s_pose[np_ + lane - t0];
        q.x = nx; q.y = ny; q.cs = nh; /* heading parked in .cs until the boxes are built */
      }
      if (f > t0) {
        x = __shfl_sync(FULL_MASK, nx, f - 1);
        y = __shfl_sync(FULL_MASK, ny, f - 1);
        h = __shfl_sync(FULL_MASK, nh, f - 1);
      }
      np_ += f - t0;
      t0 = f + 1;
    }
    if (lane == 0) {
      s_probe = len < p.H && a_end == CASIM_ROLLOUT_PROBE;
      s_nposes = np_;
      if (p.n_feasible) p.n_feasible[gk] = np_;
      if (p.end_pose) {
        p.end_pose[3 * (size_t)gk] = (float)x; p.end_pose[3 * (size_t)gk + 1] = (float)y;
        p.end_pose[3 * (size_t)gk + 2] = (float)h;
      }
    }
  } else if (p.xdt != 1 && tid == 0) { /* sub-stepped actions (xDT > 1): the plain serial replay */
    s_qn = 0;
    s_sum = 0ull;
    const float4 p0 = p.state[g], p2 = p.state[2 * N + g];
    double x = p0.x, y = p0.y, h = p0.z;
    const double rad = p2.z;
    const uint8_t* seq = p.actions + (size_t)gk * p.H;
    int np_ = 0;
    s_probe = 0;
    for (int t = 0; t < p.H; t++) {
      const int a = seq[t];
      if (a > 8) { /* end of the sequence; CASIM_ROLLOUT_PROBE: the caller only wants n_feasible (an expansion probe) */
        s_probe = a == CASIM_ROLLOUT_PROBE;
        break;
      }
      const double v = 2.0 * (double)(a / 3);
      const double w = (a % 3 == 0) ? -0.5 * CASIM_PI : ((a % 3 == 2) ? 0.5 * CASIM_PI : 0.0);
      double nx = x, ny = y, nh = h;
      bool ok = true;
      for (int i = 0; i < p.xdt && ok; i++) {
        double sn, cs;
        sincos(nh, &sn, &cs);
        const double vx = cs * v + (-sn) * 0.0, vy = sn * v + cs * 0.0;
        nx = nx + vx * c.ig_dt;
        ny = ny + vy * c.ig_dt;
        nh = nh + w * c.ig_dt;
        if (v == 0.0) continue;
        const bool in_map = c.map_hw > nx && c.map_hh > ny && nx > -c.map_hw && ny > -c.map_hh;
        if (!in_map || !(edf_value(c, d2, nx, ny, cnt_clamped) > rad + 0.1)) ok = false;
      }
      if (!ok) continue; /* ig_mcts.py:237-243: the pose stays, nothing is observed */
      x = nx; y = ny; h = nh;
      s_pose[np_].x = x; s_pose[np_].y = y; s_pose[np_].cs = h; /* heading parked in .cs until the boxes are built */
      np_++;
    }
    s_nposes = np_;
    if (p.n_feasible) p.n_feasible[gk] = np_;
    if (p.end_pose) {
      p.end_pose[3 * (size_t)gk] = (float)x; p.end_pose[3 * (size_t)gk + 1] = (float)y;
      p.end_pose[3 * (size_t)gk + 2] = (float)h;
    }
  }
  __syncthreads();
  if (s_probe) { /* feasibility only: no visibility, no reward */
    if (tid == 0) p.rewards[gk] = 0.f;
    if (p.obsvd)
      for (int w = tid; w < XR_MASK_WORDS; w += XR_THREADS) p.obsvd[(size_t)gk * XR_MASK_WORDS + w] = 0u;
    return;
  }
  const int nposes = s_nposes;
  if (tid < nposes) { /* the boxes (three fp64 sincos each) do not depend on one another: one thread per pose */
    PoseBox b;
    pose_box(c, s_pose[tid].x, s_pose[tid].y, s_pose[tid].cs, b, cnt_clamped);
    b.feasible = 1;
    s_pose[tid] = b;
  }
  __syncthreads();

  int pstart = 0;
  while (pstart < nposes) {
    int pend = pstart + 1;
    int total = s_pose[pstart].w * s_pose[pstart].h;
    while (pend < nposes && pend - pstart < 16 && total + s_pose[pend].w * s_pose[pend].h <= p.qcap) {
      total += s_pose[pend].w * s_pose[pend].h;
      pend++;
    }
    int k = pstart, base = 0, cnt = s_pose[pstart].w * s_pose[pstart].h;
    for (int idx = tid; idx < total; idx += XR_THREADS) {
      while (idx - base >= cnt) {
        base += cnt;
        k++;
        cnt = s_pose[k].w * s_pose[k].h;
      }
      const int off = idx - base;
      const PoseBox& b = s_pose[k];
      const int ii = (off * b.inv_h) >> 18, jj = off - ii * b.h;
      const int i = b.x0 + ii, j = b.y0 + jj;
      if ((s_seen[(j * 64 + i) >> 5] >> (i & 31)) & 1u) continue; /* in the set since an earlier batch */
      const double dx = cell_centre(c, i, c.map_hw) - b.x, dy = cell_centre(c, j, c.map_hh) - b.y;
      const double r0 = b.cs * dx + b.sn * dy, r1 = -b.sn * dx + b.cs * dy;
      if (in_wedge(c, r0, r1)) {
        const int q = atomicAdd(&s_qn, 1);
        s_queue[q] = (unsigned short)(((unsigned)(k - pstart) << 12) | ((unsigned)j << 6) | (unsigned)i);
      }
    }
    __syncthreads();
    const int qn = s_qn;
    for (int q = tid; q < qn; q += XR_THREADS) {
      const unsigned en = s_queue[q];
      const int k = pstart + (int)(en >> 12), j = (int)((en >> 6) & 63u), i = (int)(en & 63u);
      const int bit = j * 64 + i;
      if ((s_seen[bit >> 5] >> (bit & 31)) & 1u) continue; /* already in the set (earlier batch, or a faster thread) */
      const PoseBox& b = s_pose[k];
      if (line_of_sight(c, d2, b.x, b.y, cell_centre(c, i, c.map_hw), cell_centre(c, j, c.map_hh), cnt_samples, cnt_clamped))
        atomicOr(&s_seen[bit >> 5], 1u << (bit & 31));
    }
    __syncthreads();
    if (tid == 0) s_qn = 0;
    __syncthreads();
    pstart = pend;
  }

  /* ---- mcts_reward: MI over obsvd_cells minus the other robots' cells (targetMap.py:130-143) ---- */
  if (p.mi_table) {
    /* The belief does not change while a planner grows its trees, so a cell's (fixed-point) mutual information is a
     * constant of the planning step: the caller tabulates it once (casim_explore_mi_table) and a rollout's reward is an
     * integer sum over its cells -- the same integers as below, no logarithms.  One mask word per warp pass, lane = bit. */
    const long long* tab = p.mi_table + (size_t)g * dim * dim;
    const int lane = tid & 31;
    long long part = 0;
    for (int w = tid >> 5; w < XR_MASK_WORDS; w += XR_THREADS / 32) {
      const unsigned int seen = s_seen[w];
      if (p.obsvd && lane == 0) p.obsvd[(size_t)gk * XR_MASK_WORDS + w] = seen;
      unsigned int m = seen;
      if (m && p.exclude) m &= ~p.exclude[((size_t)(p.exclude_per_rollout ? gk : g)) * XR_MASK_WORDS + w];
      if ((m >> lane) & 1u) part += tab[(w >> 1) * dim + (w & 1) * 32 + lane];
    }
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULL_MASK, part, o);
    if (lane == 0 && part) atomicAdd(&s_sum, (unsigned long long)part);
  } else {
    const double log_rocc = log(c.ig_rocc), log_remp = log(c.ig_remp);
    const float* belief = p.belief + (size_t)g * dim * dim;
    long long part = 0;
    for (int w = tid; w < XR_MASK_WORDS; w += XR_THREADS) { /* one word = 32 cells of one half row */
      const unsigned int seen = s_seen[w];
      if (p.obsvd) p.obsvd[(size_t)gk * XR_MASK_WORDS + w] = seen;
      unsigned int m = seen;
      if (p.exclude) m &= ~p.exclude[((size_t)(p.exclude_per_rollout ? gk : g)) * XR_MASK_WORDS + w];
      const int j = w >> 1, i0 = (w & 1) * 32;
      while (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        part += __double2ll_rn(cell_mi(c, (double)belief[j * dim + i0 + b], log_rocc, log_remp) * X_MI_SCALE);
      }
    }
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULL_MASK, part, o);
    if ((tid & 31) == 0 && part) atomicAdd(&s_sum, (unsigned long long)part);
  }
  __syncthreads();
  if (tid == 0) p.rewards[gk] = (float)((double)(long long)s_sum / X_MI_SCALE);
}

/* ---------------------------------------------------------------------------------------------------------
 * ig_mcts.update_belief (policies/ig_mcts.py:117-137): the team version of the target-map update.  One CTA per
 * explorer agent i; its belief is multiplied once per pose that sees a cell -- own pose with own detections, then
 * every other explorer j of the environment (ascending j) with j's obsvd_targets: the list j detected at its own
 * latest update, i.e. this step's when j < i and j is not done (recomputed here from j's observation rows rather
 * than waited for: the detector emulation is 9 rows), else the previous call's (targets_prev).  The poses are
 * processed one after the other (a cell seen from two poses is multiplied twice, targetMap.py:86-128); inside a
 * pose every cell is owned by one thread.  team_obsv_cells = union mask, team_reward = MI of the union on the
 * updated map (:136-137).
 * --------------------------------------------------------------------------------------------------------- */
struct TeamArgs {
  DevCfg c;
  const float4* state;
  const float* oas;            /* [E*A][rows][10] */
  float* belief;               /* [E*A][dim][dim] */
  const double* targets_prev;  /* [E*A][32][2] global frame */
  const int32_t* n_prev;       /* [E*A] */
  double* targets_new;
  int32_t* n_new;
  uint32_t* team_mask;         /* [E*A][128] or NULL */
  float* team_reward;          /* [E*A] or NULL */
  const int32_t* n_agents;
  const double* edf;
  int E, A, n_maps;
};

__global__ void __launch_bounds__(X_THREADS) k_team_update(const TeamArgs p) {
  const DevCfg& c = p.c;
  const int g = blockIdx.x;
  const int e = g / p.A, ai = g - e * p.A;
  const int n = p.n_agents ? min(p.n_agents[e], p.A) : p.A;
  if (ai >= n) return;
  const size_t N = (size_t)p.E * p.A;
  const int tid = threadIdx.x;
  const float4 p5 = p.state[5 * N + g];
  if ((__float_as_int(p5.w) & 0xff) != CASIM_POLICY_IG_GREEDY) return;
  const bool done = (__float_as_int(p5.z) & CASIM_F_IS_DONE) != 0;
  if (p.team_mask && done)
    for (int w = tid; w < XR_MASK_WORDS; w += X_THREADS) p.team_mask[(size_t)g * XR_MASK_WORDS + w] = 0u;
  if (done) { /* collision_avoidance_env.py:298-299: no update; the agent keeps the list of its last update */
    const int np_ = p.n_prev[g];
    if (tid == 0) {
      p.n_new[g] = np_;
      if (p.team_reward) p.team_reward[g] = 0.f;
    }
    if (tid < 2 * np_) p.targets_new[(size_t)g * 64 + tid] = p.targets_prev[(size_t)g * 64 + tid];
    return;
  }

  __shared__ PoseBox s_box;
  __shared__ double s_det[CASIM_MAX_AGENTS][2];
  __shared__ int s_ndet, s_qn, s_np;
  __shared__ int s_who[CASIM_MAX_AGENTS];
  __shared__ unsigned short s_queue[XQ_CAP];
  __shared__ unsigned int s_seen[XR_MASK_WORDS];
  __shared__ unsigned long long s_sum;

  const int dim = c.ig_dim;
  float* belief = p.belief + (size_t)g * dim * dim;
  const double* d2 = p.edf + (size_t)(p.n_maps == 1 ? 0 : e) * CASIM_MAP_DIM * CASIM_MAP_DIM;
  unsigned cnt_samples = 0, cnt_clamped = 0;
  for (int w = tid; w < XR_MASK_WORDS; w += X_THREADS) s_seen[w] = 0u;
  if (tid == 0) { /* poses: self first, then the other explorers in index order (ig_mcts.py:84, 125-130) */
    int m = 0;
    s_who[m++] = ai;
    for (int j = 0; j < n; j++)
      if (j != ai && (__float_as_int(p.state[5 * N + (size_t)e * p.A + j].w) & 0xff) == CASIM_POLICY_IG_GREEDY) s_who[m++] = j;
    s_np = m;
    s_sum = 0ull;
  }
  __syncthreads();
  const int nposes = s_np;

  for (int m = 0; m < nposes; m++) {
    const int j = s_who[m];
    const size_t gj = (size_t)e * p.A + j;
    const float4 q0 = p.state[gj];
    const double px = q0.x, py = q0.y, ph = q0.z;
    /* ---- this pose's detections, in its ego frame (targetMap.py:89-103) ---- */
    if (tid < 32) {
      const bool j_done = (__float_as_int(p.state[5 * N + gj].z) & CASIM_F_IS_DONE) != 0;
      const bool fresh = (m == 0) || (j < ai && !j_done);
      double tx = 0, ty = 0;
      bool hit = false;
      if (fresh) { /* find_targets_in_obs (ig_mcts.py:139-156) on j's own observation */
        if (tid < c.rows) {
          const float* row = p.oas + (gj * c.rows + tid) * 10;
          if (row[9] == 1.0f) {
            const double rx = row[0], ry = row[1];
            const double dphi = atan2(ry, rx) - ph;
            const bool in_fov = fabs(dphi) <= c.ig_detect_fov_half; /* degrees vs radians, sic */
            const double rn = sqrt(rx * rx + ry * ry);
            hit = in_fov && rn <= c.ig_range;
            tx = rx + px;
            ty = ry + py;
          }
        }
      } else if (tid < p.n_prev[gj]) {
        tx = p.targets_prev[gj * 64 + 2 * tid];
        ty = p.targets_prev[gj * 64 + 2 * tid + 1];
        hit = true;
      }
      const unsigned bal = __ballot_sync(FULL_MASK, hit);
      if (hit) {
        const int k = __popc(bal & ((1u << tid) - 1u));
        const double dx = tx - px, dy = ty - py;
        double cs, sn;
        sincos(ph, &sn, &cs);
        s_det[k][0] = cs * dx + sn * dy;
        s_det[k][1] = -sn * dx + cs * dy;
        if (m == 0) { /* becomes this agent's obsvd_targets */
          p.targets_new[(size_t)g * 64 + 2 * k] = tx;
          p.targets_new[(size_t)g * 64 + 2 * k + 1] = ty;
        }
      }
      if (tid == 0) {
        s_ndet = __popc(bal);
        if (m == 0) p.n_new[g] = __popc(bal);
        s_qn = 0;
        PoseBox b;
        pose_box(c, px, py, ph, b, cnt_clamped);
        s_box = b;
      }
    }
    __syncthreads();
    const PoseBox& b = s_box;
    const int total = b.w * b.h, ndet = s_ndet;
    for (int off = tid; off < total; off += X_THREADS) {
      const int ii = (off * b.inv_h) >> 18, jj = off - ii * b.h;
      const int i = b.x0 + ii, jc = b.y0 + jj;
      const double dx = cell_centre(c, i, c.map_hw) - b.x, dy = cell_centre(c, jc, c.map_hh) - b.y;
      const double r0 = b.cs * dx + b.sn * dy, r1 = -b.sn * dx + b.cs * dy;
      if (in_wedge(c, r0, r1)) s_queue[atomicAdd(&s_qn, 1)] = (unsigned short)(((unsigned)jc << 6) | (unsigned)i);
    }
    __syncthreads();
    const int qn = s_qn;
    for (int q = tid; q < qn; q += X_THREADS) {
      const unsigned en = s_queue[q];
      const int jc = (int)((en >> 6) & 63u), i = (int)(en & 63u);
      const double gx = cell_centre(c, i, c.map_hw), gy = cell_centre(c, jc, c.map_hh);
      if (!line_of_sight(c, d2, b.x, b.y, gx, gy, cnt_samples, cnt_clamped)) continue;
      double rs = c.ig_remp;
      if (ndet > 0) {
        const double dx = gx - b.x, dy = gy - b.y;
        const double r0 = b.cs * dx + b.sn * dy, r1 = -b.sn * dx + b.cs * dy;
        for (int t = 0; t < ndet; t++) {
          const double ex = s_det[t][0] - r0, ey = s_det[t][1] - r1;
          if (sqrt(ex * ex + ey * ey) < c.ig_det_thresh) {
            rs = c.ig_rocc;
            break;
          }
        }
      }
      float* cellp = belief + jc * dim + i;
      *cellp = fminf((float)((double)*cellp * rs), FLT_MAX);
      const int bit = jc * 64 + i;
      atomicOr(&s_seen[bit >> 5], 1u << (bit & 31));
    }
    __syncthreads();
  }

  /* ---- team_obsv_cells / team_reward (ig_mcts.py:136-137) ---- */
  {
    const double log_rocc = log(c.ig_rocc), log_remp = log(c.ig_remp);
    long long part = 0;
    for (int w = tid; w < XR_MASK_WORDS; w += X_THREADS) {
      unsigned int mk = s_seen[w];
      if (p.team_mask) p.team_mask[(size_t)g * XR_MASK_WORDS + w] = mk;
      const int jc = w >> 1, i0 = (w & 1) * 32;
      while (mk && p.team_reward) {
        const int bb = __ffs(mk) - 1;
        mk &= mk - 1;
        part += __double2ll_rn(cell_mi(c, (double)belief[jc * dim + i0 + bb], log_rocc, log_remp) * X_MI_SCALE);
      }
    }
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULL_MASK, part, o);
    if ((tid & 31) == 0 && part) atomicAdd(&s_sum, (unsigned long long)part);
  }
  __syncthreads();
  if (tid == 0 && p.team_reward) p.team_reward[g] = (float)((double)(long long)s_sum / X_MI_SCALE);
}

}  // namespace

cudaError_t casim_launch_explore(casim_handle* h, const float* state, const float* oas, float* belief,
                                 float* policy_actions, float* best_mi, float* cand_mi,
                                 unsigned long long* counters, cudaStream_t s) {
  ExploreArgs a;
  a.c = h->dc;
  a.state = reinterpret_cast<const float4*>(state);
  a.oas = oas;
  a.belief = belief;
  a.policy_actions = policy_actions;
  a.best_mi = best_mi;
  a.cand_mi = cand_mi;
  a.counters = counters;
  a.n_agents = h->n_agents;
  a.edf = h->edf;
  a.init = h->init;
  a.prev_game_over = h->prev_game_over;
  a.E = h->E;
  a.A = h->A;
  a.n_maps = h->n_maps;
  if (counters) {
    cudaError_t e = cudaMemsetAsync(counters, 0, 4 * sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
  }
  h->launches++;
  if (counters) k_explore<true><<<(unsigned)((size_t)h->E * h->A), X_THREADS, 0, s>>>(a);
  else k_explore<false><<<(unsigned)((size_t)h->E * h->A), X_THREADS, 0, s>>>(a);
  return cudaGetLastError();
}


namespace {
/* cell_mi of every target-map cell of every explorer, in the 2^-44 fixed point the rollout rewards are summed in */
__global__ void k_mi_table(DevCfg c, const float4* state, const float* belief, long long* table, int E, int A) {
  const size_t N = (size_t)E * A;
  const int cells = c.ig_dim * c.ig_dim;
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * cells) return;
  const size_t g = idx / cells;
  if ((__float_as_int(state[5 * N + g].w) & 0xff) != CASIM_POLICY_IG_GREEDY) return;
  const double log_rocc = log(c.ig_rocc), log_remp = log(c.ig_remp);
  table[idx] = __double2ll_rn(cell_mi(c, (double)belief[idx], log_rocc, log_remp) * X_MI_SCALE);
}
}  // namespace

cudaError_t casim_launch_mi_table(casim_handle* h, const float* state, const float* belief, long long* table, cudaStream_t s) {
  const size_t total = (size_t)h->E * h->A * h->dc.ig_dim * h->dc.ig_dim;
  h->launches++;
  k_mi_table<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(h->dc, reinterpret_cast<const float4*>(state), belief, table, h->E, h->A);
  return cudaGetLastError();
}

cudaError_t casim_launch_rollouts(casim_handle* h, const float* state, const float* belief, const uint8_t* actions, int K,
                                  int H, int xdt, const uint32_t* exclude, int exclude_per_rollout, float* rewards,
                                  float* end_pose, uint32_t* obsvd, int32_t* n_feasible, cudaStream_t s) {
  RolloutArgs a;
  a.c = h->dc;
  a.state = reinterpret_cast<const float4*>(state);
  a.belief = belief;
  a.actions = actions;
  a.exclude = exclude;
  a.rewards = rewards;
  a.end_pose = end_pose;
  a.obsvd = obsvd;
  a.n_feasible = n_feasible;
  a.n_agents = h->n_agents;
  a.edf = h->edf;
  a.mi_table = h->mi_table;
  a.slots = h->rollout_slots;
  a.E = h->E;
  a.A = h->A;
  a.n_maps = h->n_maps;
  a.K = K;
  a.H = H;
  a.xdt = xdt;
  a.exclude_per_rollout = exclude_per_rollout;
  h->launches++;
  static const int threads = casim_tuning_int("CASIM_XR_THREADS", 64);
  static const int qcap = [] { int q = casim_tuning_int("CASIM_XR_QCAP", 1024); return q < 1 ? 1 : (q > XR_QCAP ? XR_QCAP : q); }();
  a.qcap = qcap;
  const unsigned grid = (unsigned)((h->rollout_slots ? (size_t)h->n_rollout_slots : (size_t)h->E * h->A) * K);
  if (threads == 32) k_rollouts<32><<<grid, 32, 0, s>>>(a);
  else if (threads == 128) k_rollouts<128><<<grid, 128, 0, s>>>(a);
  else k_rollouts<64><<<grid, 64, 0, s>>>(a);
  return cudaGetLastError();
}


cudaError_t casim_launch_team_update(casim_handle* h, const float* state, const float* oas, float* belief,
                                     const double* targets_prev, const int32_t* n_prev, double* targets_new,
                                     int32_t* n_new, uint32_t* team_mask, float* team_reward, cudaStream_t s) {
  TeamArgs a;
  a.c = h->dc;
  a.state = reinterpret_cast<const float4*>(state);
  a.oas = oas;
  a.belief = belief;
  a.targets_prev = targets_prev;
  a.n_prev = n_prev;
  a.targets_new = targets_new;
  a.n_new = n_new;
  a.team_mask = team_mask;
  a.team_reward = team_reward;
  a.n_agents = h->n_agents;
  a.edf = h->edf;
  a.E = h->E;
  a.A = h->A;
  a.n_maps = h->n_maps;
  h->launches++;
  k_team_update<<<(unsigned)((size_t)h->E * h->A), X_THREADS, 0, s>>>(a);
  return cudaGetLastError();
}
