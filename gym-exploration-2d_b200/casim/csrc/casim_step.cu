/*
 * casim_step.cu -- the fused environment-step kernel.
 *
 * One launch does, for every environment of the shard (citations relative to
 * gym_collision_avoidance/envs/ of mlodel/gym-exploration-2d):
 *   _take_action            collision_avoidance_env.py:287-340   in-kernel policies, fp32 action rounding
 *   Agent.take_action       agent.py:147-190                     finished branch, dynamics, goal/time budget
 *   Dynamics.step           dynamics/{Unicycle,FirstOrder,UnicycleDynamicsMaxTurnRate,External}Dynamics.py
 *   update_ego_frame        dynamics/Dynamics.py:14-28 + agent.py:250-269
 *   _check_for_collisions   collision_avoidance_env.py:630-671   all pairs + wall disc
 *   _compute_rewards        collision_avoidance_env.py:502-567
 *   OtherAgentsStatesSensor sensors/OtherAgentsStatesSensor.py:11-77   ego transform + closest-N ordering
 *   _check_which_agents_done collision_avoidance_env.py:711-738
 *   + per-step episode statistics for the shard (last-block reduction, no atomics on the data path).
 *
 * Thread mapping: one lane per agent; a warp holds floor(32/A) whole environments, so lane L of warp w
 * is element (w*epw*A + L) of every [E,A] array: all state loads/stores are contiguous float4 per warp.
 * Agents of an environment exchange their post-move pose with warp shuffles (no shared memory, no
 * barrier).
 *
 * Arithmetic: the state is fp32, the reference is float64.  Everything a decision depends on (heading
 * wrap, pose update, goal / timeout tests, pair distances, cell indices) is computed in fp64 from the fp32
 * state exactly like the reference.  Two devices keep the fp64 instruction count low without giving up
 * exact decisions:
 *   - pair predicates are evaluated on SQUARED distances (no fp64 sqrt): with s = dx^2+dy^2 and the exact
 *     c = r_i + r_j, "sqrt(s) <= c" is decided by s <= c^2 / s > c^2(1+4.5e-16); only the (astronomically
 *     rare) band in between runs the exact sqrt.  The closest-N ordering and the GETTING_CLOSE test use
 *     fp32 keys with a guard band; a warp in which any lane falls inside the band (|key difference| < 1e-4,
 *     |nearest - 0.2| < 1e-5, or an exact tie) recomputes that decision with the reference's fp64 formula.
 *   - the wall test "some occupied cell lies inside the agent's disc of cells" is min-over-occupied-cells
 *     of the integer squared distance < (r/cell)^2, i.e. ONE lookup in the exact squared distance field
 *     built at map-set time (casim_map.cu) instead of a scan over the disc.
 * Quantities that are only reported (dist_to_goal, heading_ego, the observation rows, rewards) are
 * evaluated in fp32 from fp64 differences (error ~1e-7, tolerance 1e-5).
 * The observation rows are staged in shared memory and written out as contiguous 16-byte stores.
 */
#include <stdio.h>
#include <stdlib.h>

#include "casim_device.cuh"

namespace {

struct CoreArgs {
  DevCfg c;
  float4* state;               /* [6][E*A] float4 */
  const float* ext_actions;    /* [E*A][2] or NULL (device memory, or mapped host memory in the host-buffer step) */
  float2* ext_copy_dst;        /* [E*A] or NULL: device copy of ext_actions kept current when they are read from host */
  const float* policy_actions; /* [E*A][2] or NULL */
  const float* traj;           /* [E*A][T][4] or NULL */
  const float* init;           /* [E*A][7] */
  const int32_t* type0;        /* [E*A] */
  const int32_t* n_agents;     /* [E] or NULL */
  const uint32_t* map_planes;  /* [n_maps][4][300][10] or NULL */
  const int32_t* edf_d2;       /* [n_maps][300][300] exact squared distance (cells^2) to the nearest occupied cell */
  uint8_t* prev_game_over;     /* [E] */
  const uint8_t* reset_mask;   /* [E] or NULL.  RESET mode only: environments with mask[e] == 0 are left untouched
                                  (casim_reset_envs: DummyVecEnv-style restart of the finished environments) */
  float4* state0;              /* [6][E*A] float4 episode-start state (written by RESET, re-loaded by auto_reset) */
  double* pose64;              /* [E*A][4] or NULL */
  double* partials;            /* [ctas][8] */
  unsigned int* ticket;
  casim_obs o;
  double dt;
  int E, A, T, n_maps, epw, write_pose, diag;
  int n_steps; /* STEP mode: steps per launch (casim_step_n); rewards / done / game_over / stats are [n_steps, ...] rings */
};

struct Agent {
  /* decision-carrying words in fp64 (exact widening of the fp32 state), reported-only words in fp32 */
  double px, py, heading, turning, t, trem;
  double cs, sn; /* cos/sin of the heading when this step's dynamics computed them */
  float speed, vx, vy, dheading, gx, gy, radius, pref, dgoal, pdgoal, a0v, a0w, a1v, a1w, hego;
  int step_num, flags, type;
  bool have_cs;
};

__device__ __forceinline__ void load_agent(const float4* st, size_t plane, size_t g, Agent& a) {
  float4 p0 = st[g], p1 = st[plane + g], p2 = st[2 * plane + g], p3 = st[3 * plane + g],
         p4 = st[4 * plane + g], p5 = st[5 * plane + g];
  a.px = p0.x; a.py = p0.y; a.heading = p0.z; a.speed = p0.w;
  a.vx = p1.x; a.vy = p1.y; a.dheading = p1.z; a.turning = p1.w;
  a.gx = p2.x; a.gy = p2.y; a.radius = p2.z; a.pref = p2.w;
  a.t = p3.x; a.trem = p3.y; a.dgoal = p3.z; a.pdgoal = p3.w;
  a.a0v = p4.x; a.a0w = p4.y; a.a1v = p4.z; a.a1w = p4.w;
  a.hego = p5.x;
  a.step_num = __float_as_int(p5.y);
  a.flags = __float_as_int(p5.z);
  a.type = __float_as_int(p5.w);
  a.have_cs = false;
}

__device__ __forceinline__ void store_agent(float4* st, size_t plane, size_t g, const Agent& a, bool all) {
  st[g] = make_float4((float)a.px, (float)a.py, (float)a.heading, a.speed);
  st[plane + g] = make_float4(a.vx, a.vy, a.dheading, (float)a.turning);
  if (all) st[2 * plane + g] = make_float4(a.gx, a.gy, a.radius, a.pref);
  st[3 * plane + g] = make_float4((float)a.t, (float)a.trem, a.dgoal, a.pdgoal);
  st[4 * plane + g] = make_float4(a.a0v, a.a0w, a.a1v, a.a1w);
  st[5 * plane + g] = make_float4(a.hego, __int_as_float(a.step_num), __int_as_float(a.flags), __int_as_float(a.type));
}

/* util.py:27-32 wrap() for the common case (at most one turn off), same decisions as the loop */
__device__ __forceinline__ double wrap_pi1(double a) {
  if (a >= CASIM_PI) a -= CASIM_2PI;
  if (a < -CASIM_PI) a += CASIM_2PI;
  if (!(a >= -CASIM_PI && a < CASIM_PI)) a = wrap_pi(a);
  return a;
}

__device__ __forceinline__ float wrap_pif(float a) {
  const float PI = 3.14159265358979323846f, TWO_PI = 6.28318530717958647692f;
  if (a >= PI) a -= TWO_PI;
  if (a < -PI) a += TWO_PI;
  if (!(a >= -PI && a < PI)) a = (float)wrap_pi((double)a);
  return a;
}

/* agent.py:250-269 get_ref.  Reported quantities only: fp32 from the fp64 goal offset. */
__device__ __forceinline__ void ref_axes(const Agent& a, float& p0, float& p1) {
  float gdx = (float)((double)a.gx - a.px), gdy = (float)((double)a.gy - a.py);
  float dist = sqrtf(gdx * gdx + gdy * gdy);
  if (dist > 1e-8f) {
    float inv = 1.0f / dist;
    p0 = gdx * inv;
    p1 = gdy * inv;
  } else {
    p0 = gdx;
    p1 = gdy;
  }
}

/* dynamics/Dynamics.py:14-28 update_ego_frame */
__device__ __forceinline__ void update_ego_frame(Agent& a) {
  float gdx = (float)((double)a.gx - a.px), gdy = (float)((double)a.gy - a.py);
  float dist = sqrtf(gdx * gdx + gdy * gdy);
  a.pdgoal = a.dgoal;
  a.dgoal = dist;
  if (a.t == 0.0) a.pdgoal = dist;
  /* atan2(ref_prll) == atan2(goal offset): the normalisation (or its 1e-8 guard) does not change the angle */
  a.hego = wrap_pif((float)a.heading - atan2f(gdy, gdx));
}

/* agent.py:9-112 Agent.__init__ */
__device__ __forceinline__ void agent_init(const DevCfg& c, const float* in, int type, Agent& a) {
  double sx = in[0], sy = in[1], gx = in[2], gy = in[3], pref = in[5];
  float h = in[6];
  a.px = sx; a.py = sy; a.gx = in[2]; a.gy = in[3]; a.radius = in[4]; a.pref = in[5];
  a.heading = isnan(h) ? atan2(gy - sy, gx - sx) : (double)h;
  a.speed = 0; a.vx = 0; a.vy = 0; a.dheading = 0; a.turning = 0;
  a.t = 0; a.dgoal = 0; a.pdgoal = 0; a.a0v = a.a0w = a.a1v = a.a1w = 0; a.hego = 0;
  a.step_num = 0; a.flags = 0; a.type = type;
  a.have_cs = false;
  double dx = sx - gx, dy = sy - gy;
  double straight = (sqrt(dx * dx + dy * dy) - c.near_goal) / pref;
  a.trem = c.max_time_ratio * straight;
  update_ego_frame(a);
  /* With auto-reset the policy reads this heading_ego within the same launch, so it is evaluated in fp64 and
   * rounded once, exactly like the reference's stored value (a fresh episode is off the per-step fast path). */
  a.hego = (float)wrap_pi(a.heading - atan2(gy - sy, gx - sx));
}

/* collision_avoidance_env.py:298-320: the action each policy contributes, rounded to fp32 (:289) */
__device__ __forceinline__ void choose_action(const DevCfg& c, const Agent& a, bool ext, float2 ea, const float* pol,
                                              size_t g, float& a0, float& a1) {
  a0 = 0.f;
  a1 = 0.f;
  if (a.flags & CASIM_F_IS_DONE) return;
  switch (a.type & 0xff) {
    case CASIM_POLICY_NONCOOPERATIVE: /* NonCooperativePolicy.py:10-13 */
      a0 = a.pref;
      a1 = -a.hego;
      break;
    case CASIM_POLICY_EXTERNAL:
      if (ext) {
        a0 = ea.x;
        a1 = ea.y;
      }
      break;
    case CASIM_POLICY_LEARNING: /* LearningPolicy.py:11-16 */
      if (ext) {
        double heading = c.max_heading_change * (2. * (double)ea.y - 1.);
        double speed = (double)a.pref * (double)ea.x;
        a0 = (float)speed;
        a1 = (float)heading;
      }
      break;
    case CASIM_POLICY_CARRL: /* CARRLPolicy.py:7-15 */
      if (ext) {
        int k = (int)ea.x;
        if (k < 0) k += 11;
        k = min(max(k, 0), 10);
        const double mh = CASIM_PI / 6;
        double step = (mh - (-mh)) / 10.0;
        double w = (k == 10) ? mh : (double)k * step + (-mh);
        a0 = 1.0f;
        a1 = (float)w;
      }
      break;
    case CASIM_POLICY_IG_GREEDY:
    case CASIM_POLICY_GA3C_CADRL:
    case CASIM_POLICY_RVO:
      if (pol) {
        a0 = pol[2 * g];
        a1 = pol[2 * g + 1];
      }
      break;
    default: /* Static, PedestrianDataset: zeros */
      break;
  }
}

__device__ __forceinline__ void dynamics_step(const DevCfg& c, Agent& a, float a0, float a1, double dt,
                                              const float* traj, int T, size_t g) {
  int dyn = (a.type >> 8) & 0xff;
  if (dyn == CASIM_DYN_EXTERNAL) { /* ExternalDynamics.py:8-23 + agent.py:125-145 */
    if (!traj || T <= 0) return;
    int k = a.step_num < T - 1 ? a.step_num : T - 1;
    float4 row = *reinterpret_cast<const float4*>(traj + (g * (size_t)T + k) * 4);
    double heading = atan2((double)row.z, (double)row.w) - a.heading; /* x/y swapped, sic */
    double px = row.x, py = row.y;
    double vx = 0, vy = 0;
    if (a.step_num != 0) {
      vx = (px - a.px) / c.dt_nominal;
      vy = (py - a.py) / c.dt_nominal;
    }
    a.vx = (float)vx;
    a.vy = (float)vy;
    a.dheading = (float)wrap_pi(heading - a.heading);
    a.px = px;
    a.py = py;
    a.speed = (float)sqrt(vx * vx + vy * vy);
    a.heading = heading;
    return;
  }
  double speed = (double)a0, sel;
  float speed_word = a0; /* speed_global_frame */
  if (dyn == CASIM_DYN_UNICYCLE_MAX_ACC) { /* UnicycleDynamicsMaxAcc.py:17-41 */
    double cur_speed = (double)a.speed, cur_rate = a.turning;
    double rate = fmin(fmax((double)a1 / dt, -3.0), 3.0);
    double lin_acc = fmin(fmax(2.0 * ((double)a0 - cur_speed), -2.0), 2.0);
    double turn_acc = fmin(fmax(2.0 * (rate - cur_rate), -3.0), 3.0);
    cur_speed += lin_acc * dt;
    cur_speed = fmin(fmax(cur_speed, -1.1), 1.1);
    cur_rate += turn_acc * dt;
    a.turning = cur_rate;
    speed = cur_speed;
    speed_word = (float)cur_speed;
    sel = wrap_pi1(cur_rate * dt + a.heading);
  } else if (dyn == CASIM_DYN_UNICYCLE_SECOND_ORDER_EULER) { /* UnicycleSecondOrderEulerDynamics.py:12-31 */
    double vn = sqrt((double)a.vx * (double)a.vx + (double)a.vy * (double)a.vy);
    speed = fmin(fmax(vn + (double)a0 * dt, 0.0), 1.0);
    speed_word = (float)speed;
    double ang = fmin(fmax(a.turning + (double)a1 * dt, -3.0), 3.0);
    a.turning = ang;
    sel = wrap_pi1(ang * dt + a.heading);
  } else if (dyn == CASIM_DYN_FIRST_ORDER) /* FirstOrderDynamics.py:10-23 */
    sel = wrap_pi1((double)a1 * dt + a.heading);
  else if (dyn == CASIM_DYN_UNICYCLE_MAX_TURN_RATE) { /* UnicycleDynamicsMaxTurnRate.py:11-26 */
    double rate = (double)a1 / dt;
    rate = fmin(fmax(rate, -3.0), 3.0);
    sel = wrap_pi1(rate * dt + a.heading);
  } else /* UnicycleDynamics.py:10-31 */
    sel = wrap_pi1((double)a1 + a.heading);
  double sn, cs;
  sincos(sel, &sn, &cs);
  double vx = speed * cs, vy = speed * sn;
  a.px += vx * dt;
  a.py += vy * dt;
  a.vx = (float)vx;
  a.vy = (float)vy;
  a.speed = speed_word;
  a.dheading = (float)wrap_pi1(sel - a.heading);
  a.heading = sel;
  a.cs = cs;
  a.sn = sn;
  a.have_cs = true;
  if (dyn == CASIM_DYN_UNICYCLE) {
    double td = a.turning;
    if (fabs(td) < 1e-5)
      td = 0.11 * sgn_d(sel);
    else if (td * sel < 0)
      td = fmax(-CASIM_PI, fmin(CASIM_PI, -td + sel));
    else
      td = sgn_d(td) * fmax(0.0, fabs(td) - 0.1);
    a.turning = td;
  }
}

/* agent.py:147-190 */
__device__ __forceinline__ void take_action(const DevCfg& c, Agent& a, float a0, float a1, double dt,
                                            const float* traj, int T, size_t g) {
  int flags = a.flags;
  if (flags & (CASIM_F_IS_AT_GOAL | CASIM_F_RAN_OUT_OF_TIME | CASIM_F_IN_COLLISION)) {
    if (flags & CASIM_F_IS_AT_GOAL) flags |= CASIM_F_WAS_AT_GOAL_ALREADY;
    if (flags & CASIM_F_IN_COLLISION) flags |= CASIM_F_WAS_IN_COLLISION_ALREADY;
    if (!(flags & CASIM_F_IS_AT_GOAL)) a.t += dt;
    a.vx = 0;
    a.vy = 0;
    a.flags = flags;
    return;
  }
  a.a1v = a.a0v;
  a.a1w = a.a0w;
  a.a0v = a0;
  a.a0w = a1;
  dynamics_step(c, a, a0, a1, dt, traj, T, g);
  update_ego_frame(a);
  double ex = a.px - (double)a.gx, ey = a.py - (double)a.gy;
  if (ex * ex + ey * ey <= c.near_goal * c.near_goal)
    flags |= CASIM_F_IS_AT_GOAL;
  else
    flags &= ~CASIM_F_IS_AT_GOAL;
  a.trem -= dt;
  a.t += dt;
  a.step_num += 1;
  if (a.trem <= 0.0) flags |= CASIM_F_RAN_OUT_OF_TIME;
  a.flags = flags;
}

#define CASIM_CORE_MAX_WARPS 8
#define CASIM_STAT_BIAS (1ll << 31)
#define CASIM_CORE_THREADS (32 * CASIM_CORE_MAX_WARPS) /* launch bound; the launcher picks 1..8 warps per CTA */
#define CASIM_CORE_WARPS 4                             /* default warps per CTA (CASIM_WPC overrides) */

template <int MODE, int AP>
__global__ void __launch_bounds__(CASIM_CORE_THREADS) k_core(const CoreArgs p) {
  extern __shared__ __align__(16) float s_stage[];
  /* CASIM_MODE_STEP_N = the same step body run n_steps times per launch; its own instantiation, so that the loop-carried
   * state does not cost the single-step kernel registers */
  constexpr bool IS_STEP = MODE == CASIM_MODE_STEP || MODE == CASIM_MODE_STEP_N;
  casim_pdl_enter();
  const DevCfg& c = p.c;
  const int lane = threadIdx.x & 31, warp_in_cta = threadIdx.x >> 5;
  const int warps_per_cta = blockDim.x >> 5;
  const int w = blockIdx.x * warps_per_cta + warp_in_cta;
  const int A = p.A, epw = p.epw;
  const int le = lane / A, ai = lane - le * A;
  const int env0 = w * epw;
  const int e = env0 + le;
  bool slot = le < epw && e < p.E; /* lane maps to an [E,A] element */
  if (MODE == CASIM_MODE_RESET && p.reset_mask != nullptr) slot = slot && p.reset_mask[e] != 0;
  const int n = slot ? (p.n_agents ? min(p.n_agents[e], A) : A) : 0;
  const bool valid = slot && ai < n;
  const int base = le * A; /* first lane of my environment */
  const size_t plane = (size_t)p.E * A;
  const size_t g = (size_t)env0 * A + lane;

  /* caller-supplied actions: one coalesced 8-byte read per lane.  In the host-buffer step the source is mapped host
   * memory (the read crosses PCIe while the state loads are in flight) and the device copy is refreshed here. */
  float2 ea = make_float2(0.f, 0.f);
  const bool have_ext = IS_STEP && p.ext_actions != nullptr;
  if (have_ext && slot) {
    ea = reinterpret_cast<const float2*>(p.ext_actions)[g];
    if (p.ext_copy_dst) p.ext_copy_dst[g] = ea;
  }

  /* casim_step_n: the whole body below runs n_steps times per launch with the agent's state in registers (in-kernel
   * policies only); per-step rewards / done / game_over / stats go to [n_steps, ...] rings, the observations and the
   * state are those of the last step.  A single step is n_steps == 1. */
  const int nsteps = MODE == CASIM_MODE_STEP_N ? p.n_steps : 1;
  Agent a;
  bool wrote_all = false;
  int go_prev = 0; /* my environment's game_over of the previous step of this launch */
  /* The kernel is a latency chain on few warps, so its first global loads are issued together instead of one behind the other:
   * the state planes are read for every lane of the [E,A] range before n_agents (-> valid) and the restart flag have
   * arrived; a restarting environment re-reads the episode-start state below. */
  int pgo0 = 0;
  if (MODE != CASIM_MODE_RESET && slot) {
    if (IS_STEP && c.auto_reset && p.init) pgo0 = p.prev_game_over[e];
    load_agent(p.state, plane, g, a);
  }
#pragma unroll 1
  for (int it = 0; it < nsteps; it++) {
  const bool last = it == nsteps - 1;
  int old_flags = 0;
  if (valid) {
    if (MODE == CASIM_MODE_RESET) {
      agent_init(c, p.init + g * 7, p.type0[g], a);
      wrote_all = true;
    } else {
      /* auto_reset (the env's reset() folded into the next step): the episode-start state is the one casim_reset
       * stored, i.e. exactly what reset() followed by step() would read */
      const bool fresh = IS_STEP && c.auto_reset && p.init && (it == 0 ? pgo0 != 0 : go_prev != 0);
      if (fresh) load_agent(p.state0, plane, g, a);
      a.have_cs = false;
      wrote_all = wrote_all || fresh || it > 0;
    }
  } else if (it == 0) {
    a.px = a.py = a.heading = a.turning = a.t = a.trem = a.cs = a.sn = 0;
    a.speed = a.vx = a.vy = a.dheading = a.gx = a.gy = a.radius = a.pref = a.dgoal = a.pdgoal = 0.f;
    a.a0v = a.a0w = a.a1v = a.a1w = a.hego = 0.f;
    a.step_num = a.flags = a.type = 0;
    a.have_cs = false;
  }

  if (IS_STEP && valid) {
    old_flags = a.flags;
    float a0, a1;
    choose_action(c, a, have_ext, ea, p.policy_actions, g, a0, a1);
    take_action(c, a, a0, a1, p.dt, p.traj, p.T, g);
  }

  /* ---- all pairs of my environment ---- */
  const bool me_static = (a.type & 0xff) == CASIM_POLICY_STATIC;
  const int sensors = (a.type >> 16) & 0xff;
  const bool want_oas = p.o.other_agents_states != nullptr && last;
  const double rme = (double)a.radius;
  float d32[AP];
  const bool need_dist = want_oas || p.o.nearest_dist != nullptr;
  const double close_band = c.close_range + 1e-3;
  bool coll_agent = false, amb = false;
  float nearest = INFINITY;
  if (IS_STEP || want_oas) {
#pragma unroll
    for (int j = 0; j < AP; j++) {
      int src = (base + j) & 31;
      double ox = __shfl_sync(FULL_MASK, a.px, src), oy = __shfl_sync(FULL_MASK, a.py, src);
      float orad = __shfl_sync(FULL_MASK, a.radius, src);
      int ostatic = __shfl_sync(FULL_MASK, (int)me_static, src);
      double dx = a.px - ox, dy = a.py - oy;
      double s = dx * dx + dy * dy;
      float d = 0.f;
      if (need_dist) { /* distances themselves: only the sensor rows and the reported nearest_dist need every one */
        d = sqrtf((float)s);
        d32[j] = d;
      }
      if (IS_STEP && valid && j < n && j != ai) {
        /* pair (lo, hi): skipped when the HIGHER index is Static (collision_avoidance_env.py:643-644) */
        bool hi_static = j > ai ? (ostatic != 0) : me_static;
        if (!(hi_static && !c.coll_w_static)) {
          double cr = rme + (double)orad; /* exact: both radii are fp32 values */
          double t = cr * cr;             /* exact: <= 50 significant bits */
          if (s <= t)
            coll_agent = true;                 /* sqrt(s) <= cr, hence fl(sqrt(s)) <= cr */
          else if (s <= t * (1.0 + 4.5e-16))
            amb = true;                        /* fl(sqrt(s)) may still round down onto cr */
          if (j > ai) { /* :649 lower index only; value reported / used with a guard band */
            /* the reward only asks whether the smallest gap is <= GETTING_CLOSE_RANGE (and its value then): pairs whose
             * gap is beyond the range plus a guard cannot matter unless the value itself is reported */
            const double far = cr + close_band;
            if (need_dist || s <= far * far) {
              if (!need_dist) d = sqrtf((float)s);
              float gap = (float)(s - t) / fmaxf(d + (float)cr, 1e-30f);
              nearest = fminf(nearest, gap);
            }
          }
        }
      }
    }
  }
  bool close = nearest <= (float)c.close_range;
  if (IS_STEP) {
    if (fabsf(nearest - (float)c.close_range) < 1e-5f) amb = true;
    if (__any_sync(FULL_MASK, amb)) {
      /* exact path: the reference's own fp64 formulas (collision_avoidance_env.py:640-655) */
      bool ce = false;
      double ne = INFINITY;
#pragma unroll 1
      for (int j = 0; j < AP; j++) {
        int src = (base + j) & 31;
        double ox = __shfl_sync(FULL_MASK, a.px, src), oy = __shfl_sync(FULL_MASK, a.py, src);
        float orad = __shfl_sync(FULL_MASK, a.radius, src);
        int ostatic = __shfl_sync(FULL_MASK, (int)me_static, src);
        if (valid && j < n && j != ai) {
          bool hi_static = j > ai ? (ostatic != 0) : me_static;
          if (!(hi_static && !c.coll_w_static)) {
            double dx = a.px - ox, dy = a.py - oy;
            double dd = sqrt(dx * dx + dy * dy);
            double cr = rme + (double)orad;
            if (j > ai) ne = fmin(ne, dd - cr);
            if (dd <= cr) ce = true;
          }
        }
      }
      coll_agent = ce;
      nearest = (float)ne;
      close = ne <= c.close_range;
    }
  }

  bool coll_wall = false;
  float reward = 0.f;
  bool done = false;
  int st_new_coll = 0, st_new_goal = 0, st_new_to = 0, st_active = 0, st_agents = 0, st_go = 0;
  double st_rew = 0;
  if (IS_STEP) {
    if (valid && p.n_maps > 0) {
      /* :656-666 + Map.py:67-71: an occupied cell inside the disc (col-c0)^2+(row-r0)^2 < (r/cell)^2
       * <=> the exact squared distance from the agent's cell to the nearest occupied cell is < (r/cell)^2 */
      int r0, c0;
      if (world_to_map_mul(c, a.px, a.py, r0, c0)) {
        const int32_t* ed = p.edf_d2 + (size_t)(p.n_maps == 1 ? 0 : e) * CASIM_MAP_DIM * CASIM_MAP_DIM;
        double rc = rme * c.inv_cell;
        coll_wall = (double)__ldg(ed + r0 * CASIM_MAP_DIM + c0) < rc * rc;
      }
    }
    if (valid) { /* _compute_rewards :502-567 */
      int flags = a.flags;
      double r = c.r_time_step;
      if (flags & CASIM_F_IS_AT_GOAL) {
        if (!(flags & CASIM_F_WAS_AT_GOAL_ALREADY)) r = c.r_at_goal;
      } else {
        if (!(flags & CASIM_F_WAS_IN_COLLISION_ALREADY)) {
          if (coll_agent) {
            r = c.r_coll_agent;
            flags |= CASIM_F_IN_COLLISION;
          } else if (coll_wall) {
            r = c.r_coll_wall;
            flags |= CASIM_F_IN_COLLISION;
          } else {
            if (close) r += -0.1 - (double)nearest / 2.;
            if (c.r_wiggly != 0.0) {
              double wv = (double)a.a1v - (double)a.a0v, ww = (double)a.a1w - (double)a.a0w;
              if (sqrt(wv * wv + ww * ww) > c.wiggly_thr) r += c.r_wiggly;
            }
          }
        } else if (flags & CASIM_F_RAN_OUT_OF_TIME)
          r += c.r_timeout;
        if (flags & CASIM_F_IS_INFEASIBLE) r += c.r_infeasible;
        if (c.r_dist_goal != 0.0) r += c.r_dist_goal * ((double)a.pdgoal - (double)a.dgoal);
      }
      r = fmin(fmax(r, c.r_min), c.r_max);
      r = r * c.r_inv_span;
      /* _check_which_agents_done :711-720 */
      done = (flags & (CASIM_F_IS_AT_GOAL | CASIM_F_RAN_OUT_OF_TIME | CASIM_F_IN_COLLISION)) != 0;
      if (done)
        flags |= CASIM_F_IS_DONE;
      else
        flags &= ~CASIM_F_IS_DONE;
      a.flags = flags;
      reward = (float)r;
      int nf = flags & ~old_flags;
      st_new_coll = (nf & CASIM_F_IN_COLLISION) != 0;
      st_new_goal = (nf & CASIM_F_IS_AT_GOAL) != 0;
      st_new_to = (nf & CASIM_F_RAN_OUT_OF_TIME) != 0;
      st_rew = r;
      st_active = !(old_flags & CASIM_F_IS_DONE);
      st_agents = 1;
    }
    /* game over :722-736 */
    unsigned envmask = (A == 32 ? 0xffffffffu : ((1u << A) - 1u)) << (base & 31);
    bool learning = (a.type & 0xff) == CASIM_POLICY_LEARNING;
    unsigned b_all = __ballot_sync(FULL_MASK, done || !valid);
    unsigned b_learn = __ballot_sync(FULL_MASK, done || !valid || !learning);
    unsigned b_done = __ballot_sync(FULL_MASK, done);
    if (slot && ai == 0) {
      bool go;
      if (c.game_over_mode == CASIM_GAME_OVER_ALL)
        go = (b_all & envmask) == envmask;
      else if (c.game_over_mode == CASIM_GAME_OVER_ALL_LEARNING)
        go = (b_learn & envmask) == envmask;
      else
        go = n > 0 ? ((b_done >> base) & 1u) != 0 : true;
      if (p.o.game_over) p.o.game_over[(size_t)it * p.E + e] = (uint8_t)go;
      if (last) p.prev_game_over[e] = (uint8_t)go;
      st_go = go;
    }
    if (nsteps > 1) go_prev = __shfl_sync(FULL_MASK, st_go, base & 31); /* lane ai == 0 of my environment holds it */
    if (slot) {
      if (p.o.rewards) p.o.rewards[(size_t)it * plane + g] = reward;
      if (p.o.done) p.o.done[(size_t)it * plane + g] = (uint8_t)done;
      if (last) {
        if (p.o.collision_flags) p.o.collision_flags[g] = (uint8_t)((coll_agent ? 1 : 0) | (coll_wall ? 2 : 0));
        if (p.o.nearest_dist) p.o.nearest_dist[g] = valid ? nearest : 0.f;
      }
    }
  } else if (MODE == CASIM_MODE_RESET) {
    if (slot && ai == 0) p.prev_game_over[e] = 0;
  }

  /* ---- per-step statistics: warp totals -> CTA totals (shared memory) -> ONE returning 64-bit atomic add per
   * statistic and CTA on a single accumulator word that carries the arrival count in its top 16 bits and the
   * (biased, integer) running total in the low 48: whichever CTA's add brings the count to gridDim.x holds the
   * complete total of that statistic in (returned value + its own contribution), writes the float and clears the
   * word for the next launch.  One L2 round trip on the critical path, no fence, no ticket, no second pass; integer
   * sums, so the result does not depend on the order of arrival (the reward sum is 2^-20 fixed point per CTA). ---- */
  __shared__ long long s_stats_buf[2][CASIM_CORE_MAX_WARPS][CASIM_NUM_STATS]; /* ping-pong over the steps of a launch */
  long long (*s_stats)[CASIM_NUM_STATS] = s_stats_buf[it & 1];
  if (IS_STEP && p.o.stats != nullptr) {
    long long v[CASIM_NUM_STATS];
    v[0] = __reduce_add_sync(FULL_MASK, st_new_coll);
    v[1] = __reduce_add_sync(FULL_MASK, st_new_goal);
    v[2] = __reduce_add_sync(FULL_MASK, st_new_to);
    v[3] = __reduce_add_sync(FULL_MASK, st_go);
    v[4] = warp_sum(__double2ll_rn(st_rew * 4294967296.0));
    v[5] = __reduce_add_sync(FULL_MASK, st_active);
    v[6] = __reduce_add_sync(FULL_MASK, st_agents);
    v[7] = 0;
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < CASIM_NUM_STATS; k++) s_stats[warp_in_cta][k] = v[k];
    }
    __syncthreads();
    if (warp_in_cta == 0 && lane < CASIM_NUM_STATS && p.diag < 2) { /* lane k owns statistic k */
      long long t = 0;
      for (int q = 0; q < warps_per_cta; q++) t += s_stats[q][lane];
      if (lane == 4) t = (t + (1ll << 11)) >> 12; /* 2^-32 -> 2^-20 fixed point, rounded once per CTA */
      /* |t| < 2^31 for any CTA (<= 256 agents, |reward| <= 1): the bias keeps the 48-bit field non-negative */
      const unsigned long long mine = (1ull << 48) + (unsigned long long)(t + CASIM_STAT_BIAS);
      unsigned long long* acc = reinterpret_cast<unsigned long long*>(p.partials) + (it * CASIM_NUM_STATS + lane) * 8; /* 64 B apart */
      const unsigned long long tot = atomicAdd(acc, mine) + mine;
      if ((unsigned)(tot >> 48) == gridDim.x) {
        *acc = 0ull;
        const long long sum = (long long)(tot & ((1ull << 48) - 1)) - (long long)gridDim.x * CASIM_STAT_BIAS;
        p.o.stats[it * CASIM_NUM_STATS + lane] = lane == 4 ? (float)((double)sum * (1.0 / 1048576.0)) : (float)sum;
      }
    }
  }

  /* ---- state write-back ---- */
  if (IS_STEP) {
    if (valid && last) store_agent(p.state, plane, g, a, wrote_all);
  } else if (MODE == CASIM_MODE_RESET) {
    if (slot) {
      store_agent(p.state, plane, g, a, true);
      store_agent(p.state0, plane, g, a, true);
    }
  }
  if (p.write_pose && slot && last) { /* fp64 pose for the sensor kernels: px, py, cos(heading), sin(heading) */
    if (!a.have_cs) sincos(a.heading, &a.sn, &a.cs);
    double2* q = reinterpret_cast<double2*>(p.pose64 + 4 * g);
    q[0] = make_double2(a.px, a.py);
    q[1] = make_double2(a.cs, a.sn);
  }

  /* ---- OtherAgentsStatesSensor (OtherAgentsStatesSensor.py:11-77) ---- */
  int nobs = 0;
  if (want_oas) {
    const int rows = c.rows;
    const int rowlen = rows * 10;
    float* stage = s_stage + (size_t)warp_in_cta * 32 * rowlen;
    float* mine = stage + lane * rowlen;
    const bool sense = valid && (sensors & CASIM_SENSOR_OTHER_AGENTS);
    const float horizon = (float)c.sensing_horizon;
    /* ascending stable order of key = dist - r_h - r_o (ties -> lower index first): fp32 keys, guard band */
    float key[AP];
    int rank[AP];
    int m = 0;
    bool amb_rank = false;
#pragma unroll
    for (int j = 0; j < AP; j++) {
      float orad = __shfl_sync(FULL_MASK, a.radius, (base + j) & 31);
      bool cand = sense && j < n && j != ai;
      bool ok = cand && !(d32[j] > horizon);
      if (cand && fabsf(d32[j] - horizon) < 1e-4f + 2e-6f * horizon) amb_rank = true;
      key[j] = ok ? d32[j] - a.radius - orad : INFINITY;
      rank[j] = 0;
      m += ok;
    }
#pragma unroll
    for (int j = 0; j < AP; j++)
#pragma unroll
      for (int k = j + 1; k < AP; k++) {
        float diff = key[k] - key[j];
        bool kfirst = diff < 0.f;
        rank[j] += kfirst;
        rank[k] += !kfirst;
        if (fabsf(diff) < 1e-4f + 2e-6f * fabsf(key[j])) amb_rank = true;
      }
    if (__any_sync(FULL_MASK, amb_rank)) {
      /* exact path: the reference's fp64 keys (OtherAgentsStatesSensor.py:22-29) decide the order */
      double key64[AP];
      m = 0;
#pragma unroll
      for (int j = 0; j < AP; j++) {
        int src = (base + j) & 31;
        double ox = __shfl_sync(FULL_MASK, a.px, src), oy = __shfl_sync(FULL_MASK, a.py, src);
        float orad = __shfl_sync(FULL_MASK, a.radius, src);
        double dx = a.px - ox, dy = a.py - oy;
        double dd = sqrt(dx * dx + dy * dy);
        bool ok = sense && j < n && j != ai && !(dd > c.sensing_horizon);
        key64[j] = ok ? dd - rme - (double)orad : INFINITY;
        key[j] = ok ? key[j] : INFINITY;
        if (ok && !(key[j] < INFINITY)) key[j] = (float)key64[j];
        rank[j] = 0;
        m += ok;
      }
#pragma unroll
      for (int j = 0; j < AP; j++)
#pragma unroll
        for (int k = j + 1; k < AP; k++) {
          bool kfirst = key64[k] < key64[j];
          rank[j] += kfirst;
          rank[k] += !kfirst;
        }
    }
    const int K = min(m, c.max_obs);
    nobs = min(K, rows);
    float p0 = 0.f, p1 = 0.f;
    ref_axes(a, p0, p1);
    const float q0 = -p1, q1 = p0;
    if (slot)
      for (int r = nobs; r < rows; r++)
#pragma unroll
        for (int k = 0; k < 10; k++) mine[r * 10 + k] = 0.f;
#pragma unroll
    for (int j = 0; j < AP; j++) {
      int src = (base + j) & 31;
      double ox = __shfl_sync(FULL_MASK, a.px, src), oy = __shfl_sync(FULL_MASK, a.py, src);
      float ovx = __shfl_sync(FULL_MASK, a.vx, src), ovy = __shfl_sync(FULL_MASK, a.vy, src);
      float orad = __shfl_sync(FULL_MASK, a.radius, src);
      int ostatic = __shfl_sync(FULL_MASK, (int)me_static, src);
      int row = K - 1 - rank[j];
      if (key[j] < INFINITY && rank[j] < K && row < rows) {
        float rx = (float)(ox - a.px), ry = (float)(oy - a.py);
        float* wr = mine + row * 10;
        wr[0] = rx;
        wr[1] = ry;
        wr[2] = rx * p0 + ry * p1;
        wr[3] = rx * q0 + ry * q1;
        wr[4] = ovx * p0 + ovy * p1;
        wr[5] = ovx * q0 + ovy * q1;
        wr[6] = orad;
        wr[7] = a.radius + orad;
        wr[8] = key[j];
        wr[9] = ostatic ? 1.f : 2.f;
      }
    }
    __syncwarp();
    /* coalesced copy-out of the warp's contiguous [agents][rows][10] block */
    int nslots = min(epw, max(p.E - env0, 0)) * A;
    int total = nslots * rowlen;
    float* dst = p.o.other_agents_states + (size_t)env0 * A * rowlen;
    if (MODE == CASIM_MODE_RESET && p.reset_mask != nullptr) { /* masked reset: only the rows of the restarted envs */
      const unsigned live = __ballot_sync(FULL_MASK, slot);
      for (int i = lane; i < total; i += 32)
        if ((live >> (i / rowlen)) & 1u) dst[i] = stage[i];
    } else if (total > 0) {
      if ((((uintptr_t)dst) & 15) == 0 && (total & 3) == 0) {
        const float4* s4 = reinterpret_cast<const float4*>(stage);
        float4* d4 = reinterpret_cast<float4*>(dst);
        const int n4 = total >> 2;
        int i = lane;
        for (; i + 96 < n4; i += 128) { /* four independent 16-byte copies in flight per lane */
          const float4 v0 = s4[i], v1 = s4[i + 32], v2 = s4[i + 64], v3 = s4[i + 96];
          d4[i] = v0;
          d4[i + 32] = v1;
          d4[i + 64] = v2;
          d4[i + 96] = v3;
        }
        for (; i < n4; i += 32) d4[i] = s4[i];
      } else {
        for (int i = lane; i < total; i += 32) dst[i] = stage[i];
      }
    }
  }
  if (p.o.obs_misc && slot && last) { /* rel_goal (Dynamics.py:28), num_other_agents, use_ppo (config.py:208-213) */
    float4 mo = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid)
      mo = make_float4((float)((double)a.gx - a.px), (float)((double)a.gy - a.py), (float)nobs,
                       (a.type & 0xff) == CASIM_POLICY_LEARNING ? 1.f : 0.f);
    reinterpret_cast<float4*>(p.o.obs_misc)[g] = mo;
  }
  if (MODE == CASIM_MODE_STEP_N && !last) {
    /* the state between two steps is the STORED state: fp32 words (casim.h).  Round the fp64 working copies exactly as
     * a store + load would, so that n steps in one launch equal n launches bit for bit. */
    a.px = (double)(float)a.px;
    a.py = (double)(float)a.py;
    a.heading = (double)(float)a.heading;
    a.turning = (double)(float)a.turning;
    a.t = (double)(float)a.t;
    a.trem = (double)(float)a.trem;
  }
  } /* steps of this launch */
}

template <int MODE>
cudaError_t launch_mode(const CoreArgs& a, int ap, dim3 grid, dim3 block, size_t smem, cudaStream_t s) {
#define CASE(AP)                                                                                       \
  case AP: {                                                                                           \
    if (smem > 48 * 1024)                                                                              \
      cudaFuncSetAttribute(k_core<MODE, AP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);  \
    cudaError_t le = casim_launch_pdl<CoreArgs>(k_core<MODE, AP>, grid, block, smem, s, a);            \
    if (le != cudaSuccess) return le;                                                                  \
    break;                                                                                             \
  }
  switch (ap) {
    CASE(2)
    CASE(4)
    CASE(6)
    CASE(8)
    CASE(10)
    CASE(12)
    CASE(16)
    CASE(20)
    CASE(24)
    CASE(32)
    default:
      return cudaErrorInvalidValue;
  }
#undef CASE
  return cudaGetLastError();
}

}  // namespace

cudaError_t casim_launch_core(casim_handle* h, int mode, float* state, const float* ext_actions,
                              const float* policy_actions, double dt, const casim_obs* obs, cudaStream_t s,
                              const uint8_t* reset_mask, int n_steps) {
  CoreArgs a;
  a.reset_mask = mode == CASIM_MODE_RESET ? reset_mask : nullptr;
  a.n_steps = mode == CASIM_MODE_STEP ? n_steps : 1;
  if (mode == CASIM_MODE_STEP && n_steps > 1) mode = CASIM_MODE_STEP_N;
  a.c = h->dc;
  a.state = reinterpret_cast<float4*>(state);
  a.ext_actions = ext_actions;
  a.ext_copy_dst = reinterpret_cast<float2*>(h->ext_copy_dst);
  a.policy_actions = policy_actions;
  a.traj = h->traj;
  a.T = h->T;
  a.init = h->init;
  a.type0 = h->type0;
  a.n_agents = h->n_agents;
  a.map_planes = h->map_planes;
  a.edf_d2 = h->edf_d2;
  a.n_maps = h->n_maps;
  a.prev_game_over = h->prev_game_over;
  a.state0 = reinterpret_cast<float4*>(h->state0);
  a.pose64 = h->pose64;
  a.partials = h->partials;
  a.ticket = h->ticket;
  a.o = *obs;
  a.dt = dt;
  a.E = h->E;
  a.A = h->A;
  a.epw = 32 / h->A;
  {
    /* The step is latency-bound at the named batch sizes (a few thousand warps for the whole GPU): fewer
     * environments per warp means more warps in flight per scheduler.  CASIM_EPW overrides (tuning aid). */
    static const int epw_env = casim_tuning_int("CASIM_EPW", 0);
    if (epw_env > 0 && epw_env < a.epw) a.epw = epw_env;
    static const int diag_env = casim_tuning_int("CASIM_STATS_DIAG", 0);
    a.diag = diag_env;
  }
  a.write_pose = (obs->laserscan != nullptr || obs->local_grid != nullptr) ? 1 : 0;
  static const int aps[] = {2, 4, 6, 8, 10, 12, 16, 20, 24, 32};
  int ap = 32;
  for (int k = 0; k < 10; k++)
    if (aps[k] >= h->A) {
      ap = aps[k];
      break;
    }
  int warps = (h->E + a.epw - 1) / a.epw;
  size_t per_warp = obs->other_agents_states ? (size_t)32 * h->dc.rows * 10 * sizeof(float) : 0;
  static const int wpc_env = casim_tuning_int("CASIM_WPC", 0);
  int wpc = (wpc_env >= 1 && wpc_env <= CASIM_CORE_MAX_WARPS) ? wpc_env : CASIM_CORE_WARPS;
  while (wpc > 1 && per_warp * wpc > 96 * 1024) wpc >>= 1;
  /* the statistics word counts CTA arrivals in 16 bits: widen the CTAs rather than exceed it */
  while (wpc < CASIM_CORE_MAX_WARPS && (warps + wpc - 1) / wpc > 65535 && per_warp * wpc * 2 <= 96 * 1024) wpc <<= 1;
  dim3 block(32 * wpc), grid((warps + wpc - 1) / wpc);
  if (obs->stats && grid.x > 65535) return cudaErrorInvalidConfiguration;
  if ((int)grid.x > h->max_ctas) return cudaErrorInvalidConfiguration;
  size_t smem = per_warp * wpc;
  h->launches++;
  switch (mode) {
    case CASIM_MODE_STEP:
      return launch_mode<CASIM_MODE_STEP>(a, ap, grid, block, smem, s);
    case CASIM_MODE_STEP_N:
      return launch_mode<CASIM_MODE_STEP_N>(a, ap, grid, block, smem, s);
    case CASIM_MODE_RESET:
      return launch_mode<CASIM_MODE_RESET>(a, ap, grid, block, smem, s);
    default:
      return launch_mode<CASIM_MODE_OBSERVE>(a, ap, grid, block, smem, s);
  }
}
