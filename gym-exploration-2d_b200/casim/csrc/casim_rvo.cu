/*
 * casim_rvo.cu -- the RVO policy (ORCA) for every agent whose policy id is CASIM_POLICY_RVO, one thread per agent.
 *
 * Replaces (gym_collision_avoidance/envs/policies/ of mlodel/gym-exploration-2d):
 *   RVOPolicy.find_next_action            RVOPolicy.py:53-117
 *   rvo2.PyRVOSimulator setters           Python-RVO2/src/rvo2.pyx:212-228 (double -> float)
 *   RVOSimulator::doStep (ego agent only) Python-RVO2/src/RVOSimulator.cpp:163-183
 *   Agent::computeNeighbors / insertAgentNeighbor / computeNewVelocity / update
 *                                         Python-RVO2/src/Agent.cpp:43-56, 449-473, 286-447, 516-521
 *   linearProgram1/2/3                    Python-RVO2/src/Agent.cpp:477-631
 *
 * The reference gives every RVO agent its own simulator, feeds it ALL agents of the environment each step and reads
 * back only the ego agent's new position; here each thread is one ego agent and the other agents of its
 * environment are its neighbour candidates.  The simulator arithmetic is binary32 in the reference's operation
 * order (the library is built with -fmad=false, IEEE division and square root), so the new position is bit-identical
 * to the C++ library; the host-side arithmetic of RVOPolicy.py around it (preferred velocity, heading change, the
 * pi/6 turn limit that zeroes the speed) is fp64 like the reference.
 * Not covered: obstacle ORCA lines (Agent.cpp:66-284) -- casim_rvo_step refuses a handle that has a static map.
 * Neighbour ties at exactly equal squared distance are broken by agent index (reference: k-d tree visiting order).
 *
 * Mapping note (measured on B200, C2-shaped all-RVO batch of 40 960 agents): one thread per agent takes 22.8 us;
 * a sub-warp cooperative variant (16 lanes per agent, one ORCA line per lane, ballot / shuffle linear programs --
 * kept for reference in scratch/casim_rvo_subwarp.cu.txt, parity-green) took 30.8 us: it issues 2.3x more warp
 * instructions (23 M vs ~10 M) and the kernel is issue-bound once enough warps are resident.
 *
 * Bound: latency / fp32 issue (a few hundred dependent flops per neighbour, data-dependent linear programs); the
 * HBM traffic is 48 B read + 8 B written per agent (neighbour reads hit L1/L2 within the environment).
 */
#include "casim_device.cuh"

namespace {

#define RVO_EPS 0.00001f
#define RVO_MAXN CASIM_MAX_AGENTS

struct RvoArgs {
  const float4* state; /* [6][E*A] */
  const float* coop;   /* [E*A] Agent.cooperation_coef or NULL (= 1.0, agent.py:10) */
  const int32_t* n_agents;
  float* policy_actions; /* [E*A][2] */
  double dt, max_delta_heading, radius_scale;
  float time_step, neighbor_dist, time_horizon;
  int max_neighbors, E, A;
};

struct Line {
  float2 point, direction;
};

__device__ __forceinline__ float2 V(float x, float y) { return make_float2(x, y); }
__device__ __forceinline__ float2 add(float2 a, float2 b) { return V(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 sub(float2 a, float2 b) { return V(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 neg(float2 a) { return V(-a.x, -a.y); }
__device__ __forceinline__ float2 mulvs(float2 a, float s) { return V(a.x * s, a.y * s); }
__device__ __forceinline__ float2 mulsv(float s, float2 a) { return V(s * a.x, s * a.y); }
__device__ __forceinline__ float2 divs(float2 a, float s) { /* Vector2.h:121-126 */
  float inv = 1.0f / s;
  return V(a.x * inv, a.y * inv);
}
__device__ __forceinline__ float dot(float2 a, float2 b) { return a.x * b.x + a.y * b.y; }
__device__ __forceinline__ float det(float2 a, float2 b) { return a.x * b.y - a.y * b.x; }
__device__ __forceinline__ float absSq(float2 a) { return dot(a, a); }
__device__ __forceinline__ float2 normalize(float2 a) { return divs(a, sqrtf(dot(a, a))); }
__device__ __forceinline__ float sqr(float a) { return a * a; }

/* Agent.cpp:477-546 */
__device__ bool lp1(const Line* lines, int lineNo, float radius, float2 opt, bool dirOpt, float2& result) {
  const Line L = lines[lineNo];
  const float dotProduct = dot(L.point, L.direction);
  const float discriminant = sqr(dotProduct) + sqr(radius) - absSq(L.point);
  if (discriminant < 0.0f) return false;
  const float sq = sqrtf(discriminant);
  float tLeft = -dotProduct - sq;
  float tRight = -dotProduct + sq;
  for (int i = 0; i < lineNo; i++) {
    const float denominator = det(L.direction, lines[i].direction);
    const float numerator = det(lines[i].direction, sub(L.point, lines[i].point));
    if (fabsf(denominator) <= RVO_EPS) {
      if (numerator < 0.0f) return false;
      continue;
    }
    const float t = numerator / denominator;
    if (denominator >= 0.0f)
      tRight = (t < tRight) ? t : tRight;
    else
      tLeft = (tLeft < t) ? t : tLeft;
    if (tLeft > tRight) return false;
  }
  if (dirOpt) {
    if (dot(opt, L.direction) > 0.0f)
      result = add(L.point, mulsv(tRight, L.direction));
    else
      result = add(L.point, mulsv(tLeft, L.direction));
  } else {
    const float t = dot(L.direction, sub(opt, L.point));
    if (t < tLeft)
      result = add(L.point, mulsv(tLeft, L.direction));
    else if (t > tRight)
      result = add(L.point, mulsv(tRight, L.direction));
    else
      result = add(L.point, mulsv(t, L.direction));
  }
  return true;
}

/* Agent.cpp:548-578 */
__device__ int lp2(const Line* lines, int nl, float radius, float2 opt, bool dirOpt, float2& result) {
  if (dirOpt)
    result = mulvs(opt, radius);
  else if (absSq(opt) > sqr(radius))
    result = mulvs(normalize(opt), radius);
  else
    result = opt;
  for (int i = 0; i < nl; i++) {
    if (det(lines[i].direction, sub(lines[i].point, result)) > 0.0f) {
      const float2 temp = result;
      if (!lp1(lines, i, radius, opt, dirOpt, result)) {
        result = temp;
        return i;
      }
    }
  }
  return nl;
}

/* Agent.cpp:580-630 with numObstLines = 0 */
__device__ void lp3(const Line* lines, int nl, int beginLine, float radius, float2& result, Line* proj) {
  float distance = 0.0f;
  for (int i = beginLine; i < nl; i++) {
    if (det(lines[i].direction, sub(lines[i].point, result)) > distance) {
      int np = 0;
      for (int j = 0; j < i; j++) {
        Line l;
        const float determinant = det(lines[i].direction, lines[j].direction);
        if (fabsf(determinant) <= RVO_EPS) {
          if (dot(lines[i].direction, lines[j].direction) > 0.0f) continue;
          l.point = mulsv(0.5f, add(lines[i].point, lines[j].point));
        } else {
          l.point = add(lines[i].point,
                        mulsv(det(lines[j].direction, sub(lines[i].point, lines[j].point)) / determinant, lines[i].direction));
        }
        l.direction = normalize(sub(lines[j].direction, lines[i].direction));
        proj[np++] = l;
      }
      const float2 temp = result;
      if (lp2(proj, np, radius, V(-lines[i].direction.y, lines[i].direction.x), true, result) < np) result = temp;
      distance = det(lines[i].direction, sub(lines[i].point, result));
    }
  }
}

#define RVO_THREADS 128

__global__ void __launch_bounds__(RVO_THREADS) k_rvo(const RvoArgs p) {
  casim_pdl_enter();
  const size_t N = (size_t)p.E * p.A;
  const size_t g = (size_t)blockIdx.x * RVO_THREADS + threadIdx.x;
  if (g >= N) return;
  const int A = p.A;
  const int e = (int)(g / A), ego = (int)(g - (size_t)e * A);
  const int n = p.n_agents ? min(p.n_agents[e], A) : A;
  if (ego >= n) return;
  const float4* s0 = p.state + (size_t)e * A; /* px, py, heading, speed */
  const float4* s1 = s0 + N;                  /* vx, vy, ... */
  const float4* s2 = s1 + N;                  /* gx, gy, radius, pref_speed */
  const float4* s5 = s2 + 3 * N;
  const int type = __float_as_int(s5[ego].w);
  if ((type & 0xff) != CASIM_POLICY_RVO) return;

  const float4 me0 = s0[ego], me1 = s1[ego], me2 = s2[ego];
  const float2 position = V(me0.x, me0.y), velocity = V(me1.x, me1.y);
  const float radius_ = (float)(p.radius_scale * (double)me2.z); /* RVOPolicy.py:78 + rvo2.pyx:218 */
  const float maxSpeed = me2.w;                                  /* :77 */
  const float collab = p.coop ? p.coop[g] : 1.0f;
  float2 prefVel;
  { /* RVOPolicy.py:73-74 in fp64, then the binding's float conversion */
    double dx = (double)me2.x - (double)me0.x, dy = (double)me2.y - (double)me0.y;
    double k = (double)me2.w / sqrt(dx * dx + dy * dy);
    prefVel = V((float)(k * dx), (float)(k * dy));
  }

  /* Agent.cpp:43-56 + 449-473: neighbours sorted by squared distance, at most maxNeighbors */
  float ndist[RVO_MAXN];
  int nidx[RVO_MAXN];
  int nn = 0;
  if (p.max_neighbors > 0) {
    float rangeSq = sqr(p.neighbor_dist);
    for (int j = 0; j < n; j++) {
      if (j == ego) continue;
      const float4 o0 = s0[j];
      const float distSq = absSq(sub(position, V(o0.x, o0.y)));
      if (distSq < rangeSq) {
        if (nn < p.max_neighbors) nn++;
        int i = nn - 1;
        while (i != 0 && distSq < ndist[i - 1]) {
          ndist[i] = ndist[i - 1];
          nidx[i] = nidx[i - 1];
          --i;
        }
        ndist[i] = distSq;
        nidx[i] = j;
        if (nn == p.max_neighbors) rangeSq = ndist[nn - 1];
      }
    }
  }

  Line lines[RVO_MAXN + 2], proj[RVO_MAXN + 2];
  int nl = 0;
  const float invTimeHorizon = 1.0f / p.time_horizon;
  if (collab == 0.0f) {
    /* Agent.cpp:290-292 */
  } else if (collab > 0.0f) {
    for (int k = 0; k < nn; k++) { /* Agent.cpp:298-364 */
      const int o = nidx[k];
      const float4 o0 = s0[o], o1 = s1[o], o2 = s2[o];
      const float2 relativePosition = sub(V(o0.x, o0.y), position);
      const float2 relativeVelocity = sub(velocity, V(o1.x, o1.y));
      const float distSq = absSq(relativePosition);
      const float combinedRadius = radius_ + (float)(p.radius_scale * (double)o2.z);
      const float combinedRadiusSq = sqr(combinedRadius);
      Line line;
      float2 u;
      if (distSq > combinedRadiusSq) {
        const float2 w = sub(relativeVelocity, mulsv(invTimeHorizon, relativePosition));
        const float wLengthSq = absSq(w);
        const float dotProduct1 = dot(w, relativePosition);
        if (dotProduct1 < 0.0f && sqr(dotProduct1) > combinedRadiusSq * wLengthSq) {
          const float wLength = sqrtf(wLengthSq);
          const float2 unitW = divs(w, wLength);
          line.direction = V(unitW.y, -unitW.x);
          u = mulsv(combinedRadius * invTimeHorizon - wLength, unitW);
        } else {
          const float leg = sqrtf(distSq - combinedRadiusSq);
          if (det(relativePosition, w) > 0.0f)
            line.direction = divs(V(relativePosition.x * leg - relativePosition.y * combinedRadius,
                                    relativePosition.x * combinedRadius + relativePosition.y * leg), distSq);
          else
            line.direction = divs(neg(V(relativePosition.x * leg + relativePosition.y * combinedRadius,
                                        -relativePosition.x * combinedRadius + relativePosition.y * leg)), distSq);
          const float dotProduct2 = dot(relativeVelocity, line.direction);
          u = sub(mulsv(dotProduct2, line.direction), relativeVelocity);
        }
      } else {
        const float invTimeStep = 1.0f / p.time_step;
        const float2 w = sub(relativeVelocity, mulsv(invTimeStep, relativePosition));
        const float wLength = sqrtf(dot(w, w));
        const float2 unitW = divs(w, wLength);
        line.direction = V(unitW.y, -unitW.x);
        u = mulsv(combinedRadius * invTimeStep - wLength, unitW);
      }
      line.point = add(velocity, mulsv(collab, u)); /* Agent.cpp:361 */
      lines[nl++] = line;
    }
  } else if (nn > 0) { /* Agent.cpp:367-416 */
    float minDistSq = INFINITY;
    int closest = 0;
    for (int k = 0; k < nn; k++) {
      const float4 o0 = s0[nidx[k]];
      const float d = absSq(sub(V(o0.x, o0.y), position));
      if (d < minDistSq) {
        minDistSq = d;
        closest = k;
      }
    }
    const int o = nidx[closest];
    const float4 o0 = s0[o], o1 = s1[o], o2 = s2[o];
    const float2 relativePosition = sub(V(o0.x, o0.y), position);
    const float2 ovel = V(o1.x, o1.y);
    const float distSq = absSq(relativePosition);
    const float combinedRadius = radius_ + (float)(p.radius_scale * (double)o2.z);
    const float combinedRadiusSq = sqr(combinedRadius);
    const float leg = sqrtf(distSq - combinedRadiusSq);
    Line line;
    line.direction = divs(neg(V(relativePosition.x * leg - relativePosition.y * combinedRadius,
                                relativePosition.x * combinedRadius + relativePosition.y * leg)), distSq);
    line.point = add(ovel, relativePosition);
    lines[nl++] = line;
    line.direction = divs(V(relativePosition.x * leg + relativePosition.y * combinedRadius,
                            -relativePosition.x * combinedRadius + relativePosition.y * leg), distSq);
    line.point = add(ovel, relativePosition);
    lines[nl++] = line;
    prefVel = add(ovel, mulsv(-collab, normalize(relativePosition)));
  }
  float2 newVel = V(0.f, 0.f);
  const int lineFail = lp2(lines, nl, maxSpeed, prefVel, false, newVel);
  if (lineFail < nl) lp3(lines, nl, lineFail, maxSpeed, newVel, proj);
  const float2 newPos = add(position, mulvs(newVel, p.time_step)); /* Agent.cpp:516-521 */

  /* RVOPolicy.py:91-110 (fp64) */
  const double ddx = (double)newPos.x - (double)me0.x, ddy = (double)newPos.y - (double)me0.y;
  const double ang1 = atan2(ddy, ddx);
  double new_heading = fmod(ang1, CASIM_2PI); /* np.mod(ang1 - 0, 2 pi) */
  if (new_heading != 0.0 && new_heading < 0.0) new_heading += CASIM_2PI;
  double delta_heading = wrap_pi(new_heading - (double)me0.z);
  double pref_speed = 1 / p.dt * sqrt(ddx * ddx + ddy * ddy);
  if (fabs(delta_heading) > p.max_delta_heading) {
    delta_heading = (delta_heading > 0 ? 1.0 : (delta_heading < 0 ? -1.0 : 0.0)) * p.max_delta_heading;
    pref_speed = 0.;
  }
  reinterpret_cast<float2*>(p.policy_actions)[g] = make_float2((float)pref_speed, (float)delta_heading);
}

}  // namespace

cudaError_t casim_launch_rvo(casim_handle* h, const float* state, const float* coop, const casim_rvo_params* prm,
                             float* policy_actions, cudaStream_t s) {
  RvoArgs a;
  a.state = reinterpret_cast<const float4*>(state);
  a.coop = coop;
  a.n_agents = h->n_agents;
  a.policy_actions = policy_actions;
  a.dt = h->cfg.dt;
  a.max_delta_heading = prm->max_delta_heading;
  a.radius_scale = prm->radius_scale;
  a.time_step = (float)h->cfg.dt;
  a.neighbor_dist = (float)prm->neighbor_dist;
  a.time_horizon = (float)prm->time_horizon;
  a.max_neighbors = prm->max_neighbors;
  a.E = h->E;
  a.A = h->A;
  const size_t N = (size_t)h->E * h->A;
  h->launches++;
  return casim_launch_pdl<RvoArgs>(k_rvo, dim3((unsigned)((N + RVO_THREADS - 1) / RVO_THREADS)), dim3(RVO_THREADS), 0, s, a);
}
