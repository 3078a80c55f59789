/*
 * casim_oracle_rvo.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see casim_oracle.c).
 *
 * Plain-C restatement of the RVO policy of mlodel/gym-exploration-2d (paths relative to
 * gym_collision_avoidance/envs/policies/):
 *   RVOPolicy.py:53-117                     find_next_action (fp64 host arithmetic around the simulator)
 *   Python-RVO2/src/rvo2.pyx:212-228        double -> float at the binding
 *   Python-RVO2/src/RVOSimulator.cpp:163-183 doStep
 *   Python-RVO2/src/Agent.cpp:43-56         computeNeighbors (agents; obstacle neighbours: see below)
 *   Python-RVO2/src/Agent.cpp:286-447       computeNewVelocity, agent ORCA lines incl. the collabCoeff variants
 *   Python-RVO2/src/Agent.cpp:449-473       insertAgentNeighbor
 *   Python-RVO2/src/Agent.cpp:477-631       linearProgram1/2/3
 *   Python-RVO2/src/Agent.cpp:516-521       update
 * All simulator arithmetic is IEEE binary32 in the reference's operation order (built with -ffp-contract=off).
 *
 * Parity status: PINNED.  The reference's own C++ sources compile here (oracle/Makefile `ref` ->
 * oracle/_ref/librvo2_ref.so); tests/test_oracle_golden.py checks this file bit for bit against that library when
 * it is present and against tests/golden/rvo.npz (generated from it by oracle/gen_golden_rvo.py) everywhere.
 * Scope: agent-agent ORCA lines.  Static obstacles reach the reference simulator only through the first
 * find_next_action call (RVOPolicy.py:56-57 adds them on every call but processObstacles runs once, :44); obstacle
 * ORCA lines (Agent.cpp:66-284) are not restated -- the entry point refuses RVO agents together with obstacles.
 * Neighbour ties (two agents at exactly the same squared distance) are broken by agent index; the reference
 * breaks them by k-d tree visiting order.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/casim.h"

#define NF CASIM_STATE_WORDS
#define RVO_EPS 0.00001f /* Definitions.h:52 */
#define MAXN CASIM_MAX_AGENTS

typedef struct { float x, y; } v2;
typedef struct { v2 point, direction; } line_t;

static v2 V(float x, float y) { v2 r = {x, y}; return r; }
static v2 add(v2 a, v2 b) { return V(a.x + b.x, a.y + b.y); }
static v2 sub(v2 a, v2 b) { return V(a.x - b.x, a.y - b.y); }
static v2 neg(v2 a) { return V(-a.x, -a.y); }
static v2 scale(v2 a, float s) { return V(a.x * s, a.y * s); }   /* Vector2 * float */
static v2 scalel(float s, v2 a) { return V(s * a.x, s * a.y); }  /* float * Vector2 */
static v2 divs(v2 a, float s) { float inv = 1.0f / s; return V(a.x * inv, a.y * inv); } /* Vector2.h:121-126 */
static float dot(v2 a, v2 b) { return a.x * b.x + a.y * b.y; }
static float det(v2 a, v2 b) { return a.x * b.y - a.y * b.x; }
static float absSq(v2 a) { return dot(a, a); }
static float vabs(v2 a) { return sqrtf(dot(a, a)); }
static v2 normalize(v2 a) { return divs(a, vabs(a)); }
static float sqr(float a) { return a * a; }

/* Agent.cpp:477-546 */
static int lp1(const line_t* lines, int lineNo, float radius, v2 opt, int dirOpt, v2* result) {
  const float dotProduct = dot(lines[lineNo].point, lines[lineNo].direction);
  const float discriminant = sqr(dotProduct) + sqr(radius) - absSq(lines[lineNo].point);
  if (discriminant < 0.0f) return 0;
  const float sq = sqrtf(discriminant);
  float tLeft = -dotProduct - sq;
  float tRight = -dotProduct + sq;
  for (int i = 0; i < lineNo; i++) {
    const float denominator = det(lines[lineNo].direction, lines[i].direction);
    const float numerator = det(lines[i].direction, sub(lines[lineNo].point, lines[i].point));
    if (fabsf(denominator) <= RVO_EPS) {
      if (numerator < 0.0f) return 0;
      continue;
    }
    const float t = numerator / denominator;
    if (denominator >= 0.0f)
      tRight = (t < tRight) ? t : tRight; /* std::min(tRight, t) */
    else
      tLeft = (tLeft < t) ? t : tLeft; /* std::max(tLeft, t) */
    if (tLeft > tRight) return 0;
  }
  if (dirOpt) {
    if (dot(opt, lines[lineNo].direction) > 0.0f)
      *result = add(lines[lineNo].point, scalel(tRight, lines[lineNo].direction));
    else
      *result = add(lines[lineNo].point, scalel(tLeft, lines[lineNo].direction));
  } else {
    const float t = dot(lines[lineNo].direction, sub(opt, lines[lineNo].point));
    if (t < tLeft)
      *result = add(lines[lineNo].point, scalel(tLeft, lines[lineNo].direction));
    else if (t > tRight)
      *result = add(lines[lineNo].point, scalel(tRight, lines[lineNo].direction));
    else
      *result = add(lines[lineNo].point, scalel(t, lines[lineNo].direction));
  }
  return 1;
}

/* Agent.cpp:548-578 */
static int lp2(const line_t* lines, int nl, float radius, v2 opt, int dirOpt, v2* result) {
  if (dirOpt)
    *result = scale(opt, radius);
  else if (absSq(opt) > sqr(radius))
    *result = scale(normalize(opt), radius);
  else
    *result = opt;
  for (int i = 0; i < nl; i++) {
    if (det(lines[i].direction, sub(lines[i].point, *result)) > 0.0f) {
      const v2 temp = *result;
      if (!lp1(lines, i, radius, opt, dirOpt, result)) {
        *result = temp;
        return i;
      }
    }
  }
  return nl;
}

/* Agent.cpp:580-630 (numObstLines = 0 here) */
static void lp3(const line_t* lines, int nl, int numObst, int beginLine, float radius, v2* result) {
  float distance = 0.0f;
  line_t proj[2 * MAXN];
  for (int i = beginLine; i < nl; i++) {
    if (det(lines[i].direction, sub(lines[i].point, *result)) > distance) {
      int np = 0;
      for (int j = 0; j < numObst; j++) proj[np++] = lines[j];
      for (int j = numObst; j < i; j++) {
        line_t l;
        const float determinant = det(lines[i].direction, lines[j].direction);
        if (fabsf(determinant) <= RVO_EPS) {
          if (dot(lines[i].direction, lines[j].direction) > 0.0f) continue;
          l.point = scalel(0.5f, add(lines[i].point, lines[j].point));
        } else {
          l.point = add(lines[i].point,
                        scalel(det(lines[j].direction, sub(lines[i].point, lines[j].point)) / determinant, lines[i].direction));
        }
        l.direction = normalize(sub(lines[j].direction, lines[i].direction));
        proj[np++] = l;
      }
      const v2 temp = *result;
      if (lp2(proj, np, radius, V(-lines[i].direction.y, lines[i].direction.x), 1, result) < np) *result = temp;
      distance = det(lines[i].direction, sub(lines[i].point, *result));
    }
  }
}

/* One ego agent of RVOSimulator::doStep (RVOSimulator.cpp:163-183): computeNeighbors + computeNewVelocity + update.
 * pos/vel/pref/radius/max_speed are the values AFTER the binding's float conversion. */
static void rvo_ego_step(int n, const v2* pos, const v2* vel, const v2* pref, const float* radius, const float* max_speed,
                         int ego, float collab, float timeStep, float neighborDist, int maxNeighbors, float timeHorizon,
                         v2* new_pos) {
  /* Agent.cpp:43-56 + 449-473: neighbours sorted by squared distance, at most maxNeighbors */
  float ndist[MAXN];
  int nidx[MAXN];
  int nn = 0;
  if (maxNeighbors > 0) {
    float rangeSq = sqr(neighborDist);
    for (int j = 0; j < n; j++) {
      if (j == ego) continue;
      const float distSq = absSq(sub(pos[ego], pos[j]));
      if (distSq < rangeSq) {
        if (nn < maxNeighbors) nn++;
        int i = nn - 1;
        while (i != 0 && distSq < ndist[i - 1]) {
          ndist[i] = ndist[i - 1];
          nidx[i] = nidx[i - 1];
          --i;
        }
        ndist[i] = distSq;
        nidx[i] = j;
        if (nn == maxNeighbors) rangeSq = ndist[nn - 1];
      }
    }
  }
  line_t lines[MAXN + 2];
  int nl = 0;
  v2 prefVel = pref[ego];
  const v2 position = pos[ego], velocity = vel[ego];
  const float radius_ = radius[ego];
  const float invTimeHorizon = 1.0f / timeHorizon;
  if (collab == 0.0f) {
    /* Agent.cpp:290-292: ignores the other agents */
  } else if (collab > 0.0f) {
    for (int k = 0; k < nn; k++) { /* Agent.cpp:298-364 */
      const int o = nidx[k];
      const v2 relativePosition = sub(pos[o], position);
      const v2 relativeVelocity = sub(velocity, vel[o]);
      const float distSq = absSq(relativePosition);
      const float combinedRadius = radius_ + radius[o];
      const float combinedRadiusSq = sqr(combinedRadius);
      line_t line;
      v2 u;
      if (distSq > combinedRadiusSq) {
        const v2 w = sub(relativeVelocity, scalel(invTimeHorizon, relativePosition));
        const float wLengthSq = absSq(w);
        const float dotProduct1 = dot(w, relativePosition);
        if (dotProduct1 < 0.0f && sqr(dotProduct1) > combinedRadiusSq * wLengthSq) {
          const float wLength = sqrtf(wLengthSq);
          const v2 unitW = divs(w, wLength);
          line.direction = V(unitW.y, -unitW.x);
          u = scalel(combinedRadius * invTimeHorizon - wLength, unitW);
        } else {
          const float leg = sqrtf(distSq - combinedRadiusSq);
          if (det(relativePosition, w) > 0.0f)
            line.direction = divs(V(relativePosition.x * leg - relativePosition.y * combinedRadius,
                                    relativePosition.x * combinedRadius + relativePosition.y * leg), distSq);
          else
            line.direction = divs(neg(V(relativePosition.x * leg + relativePosition.y * combinedRadius,
                                        -relativePosition.x * combinedRadius + relativePosition.y * leg)), distSq);
          const float dotProduct2 = dot(relativeVelocity, line.direction);
          u = sub(scalel(dotProduct2, line.direction), relativeVelocity);
        }
      } else {
        const float invTimeStep = 1.0f / timeStep;
        const v2 w = sub(relativeVelocity, scalel(invTimeStep, relativePosition));
        const float wLength = vabs(w);
        const v2 unitW = divs(w, wLength);
        line.direction = V(unitW.y, -unitW.x);
        u = scalel(combinedRadius * invTimeStep - wLength, unitW);
      }
      line.point = add(velocity, scalel(collab, u)); /* Agent.cpp:361 */
      lines[nl++] = line;
    }
  } else if (nn > 0) { /* Agent.cpp:367-416: negative collaboration, aim at the closest agent */
    float minDistSq = INFINITY;
    int closest = 0;
    for (int k = 0; k < nn; k++) {
      const float d = absSq(sub(pos[nidx[k]], position));
      if (d < minDistSq) {
        minDistSq = d;
        closest = k;
      }
    }
    const int o = nidx[closest];
    const v2 relativePosition = sub(pos[o], position);
    const float distSq = absSq(relativePosition);
    const float combinedRadius = radius_ + radius[o];
    const float combinedRadiusSq = sqr(combinedRadius);
    const float leg = sqrtf(distSq - combinedRadiusSq);
    line_t line;
    line.direction = divs(neg(V(relativePosition.x * leg - relativePosition.y * combinedRadius,
                                relativePosition.x * combinedRadius + relativePosition.y * leg)), distSq);
    line.point = add(vel[o], relativePosition);
    lines[nl++] = line;
    line.direction = divs(V(relativePosition.x * leg + relativePosition.y * combinedRadius,
                            -relativePosition.x * combinedRadius + relativePosition.y * leg), distSq);
    line.point = add(vel[o], relativePosition);
    lines[nl++] = line;
    prefVel = add(vel[o], scalel(-collab, normalize(relativePosition)));
  }
  v2 newVel = V(0.f, 0.f);
  const int lineFail = lp2(lines, nl, max_speed[ego], prefVel, 0, &newVel);
  if (lineFail < nl) lp3(lines, nl, 0, lineFail, max_speed[ego], &newVel);
  /* Agent.cpp:516-521 update */
  *new_pos = add(position, scale(newVel, timeStep));
}

/* Same signature as oracle/rvo2_ref_shim.cpp:rvo2_ref_step, so that the two can be compared call for call. */
int casim_oracle_rvo_ref_step(int n, const double* pos, const double* vel, const double* pref_vel, const double* radius,
                              const double* max_speed, int ego, double collab, double dt, double neighbor_dist,
                              int max_neighbors, double time_horizon, double* new_pos) {
  if (n > MAXN) return 1;
  v2 p[MAXN], v[MAXN], pv[MAXN];
  float r[MAXN], ms[MAXN];
  for (int a = 0; a < n; a++) {
    p[a] = V((float)pos[2 * a], (float)pos[2 * a + 1]);
    v[a] = V((float)vel[2 * a], (float)vel[2 * a + 1]);
    pv[a] = V((float)pref_vel[2 * a], (float)pref_vel[2 * a + 1]);
    r[a] = (float)radius[a];
    ms[a] = (float)max_speed[a];
  }
  v2 np_;
  rvo_ego_step(n, p, v, pv, r, ms, ego, (float)collab, (float)dt, (float)neighbor_dist, max_neighbors, (float)time_horizon, &np_);
  new_pos[0] = np_.x;
  new_pos[1] = np_.y;
  return 0;
}

/* util.py:27-32 */
static double wrap_angle(double a) {
  while (a >= M_PI) a -= 2 * M_PI;
  while (a < -M_PI) a += 2 * M_PI;
  return a;
}

/* np.mod(a, b) for b > 0 */
static double np_mod(double a, double b) {
  double r = fmod(a, b);
  if (r != 0.0 && r < 0.0) r += b;
  return r;
}

/* RVOPolicy.find_next_action (RVOPolicy.py:53-117) for every agent whose policy is CASIM_POLICY_RVO.
 * state: [E,A,24] doubles (oracle layout); coop: [E,A] Agent.cooperation_coef (agent.py:103) or NULL = 1.0;
 * actions: [E,A,2] = [pref_speed, delta_heading] (zeros for other agents).  The simulator instance of each policy
 * is fed ALL agents of its environment (:66-82) and only the ego agent's new position is read back (:91). */
typedef struct {
  const casim_config* cfg; const casim_rvo_params* prm; int A; const int32_t* n_agents;
  const double* state; const double* coop; double* actions; int e0, e1;
} rvo_job;

static void* rvo_job_main(void* arg) {
  const rvo_job* J = (const rvo_job*)arg;
  const casim_config* cfg = J->cfg; const casim_rvo_params* prm = J->prm;
  const int A = J->A; const int32_t* n_agents = J->n_agents;
  const double* state = J->state; const double* coop = J->coop; double* actions = J->actions;
  const double dt = cfg->dt;
  for (int e = J->e0; e < J->e1; e++) {
    const int n = n_agents ? (n_agents[e] < A ? n_agents[e] : A) : A;
    const double* S = state + (size_t)e * A * NF;
    v2 p[MAXN], v[MAXN], pv[MAXN];
    float r[MAXN], ms[MAXN];
    for (int a = 0; a < n; a++) { /* :66-82 */
      const double* s = S + (size_t)a * NF;
      double dx = s[CASIM_GX] - s[CASIM_PX], dy = s[CASIM_GY] - s[CASIM_PY];
      double k = s[CASIM_PREF_SPEED] / sqrt(dx * dx + dy * dy); /* pref_speed / np.linalg.norm(.) * vec */
      p[a] = V((float)s[CASIM_PX], (float)s[CASIM_PY]);
      v[a] = V((float)s[CASIM_VX], (float)s[CASIM_VY]);
      pv[a] = V((float)(k * dx), (float)(k * dy));
      r[a] = (float)(prm->radius_scale * s[CASIM_RADIUS]);
      ms[a] = (float)s[CASIM_PREF_SPEED];
    }
    for (int a = 0; a < A; a++) {
      double* act = actions + ((size_t)e * A + a) * 2;
      act[0] = act[1] = 0.0;
      if (a >= n) continue;
      const double* s = S + (size_t)a * NF;
      if ((((int)s[CASIM_TYPE]) & 0xff) != CASIM_POLICY_RVO) continue;
      v2 np_;
      rvo_ego_step(n, p, v, pv, r, ms, a, (float)(coop ? coop[(size_t)e * A + a] : 1.0), (float)dt,
                   (float)prm->neighbor_dist, prm->max_neighbors, (float)prm->time_horizon, &np_);
      /* :91-110 */
      double ddx = (double)np_.x - s[CASIM_PX], ddy = (double)np_.y - s[CASIM_PY];
      double ang1 = atan2(ddy, ddx);
      double ang2 = atan2(0.0, 1.0);
      double new_heading = np_mod(ang1 - ang2, 2 * M_PI);
      double delta_heading = wrap_angle(new_heading - s[CASIM_HEADING]);
      double pref_speed = 1 / dt * sqrt(ddx * ddx + ddy * ddy);
      if (fabs(delta_heading) > prm->max_delta_heading) {
        delta_heading = (delta_heading > 0 ? 1.0 : (delta_heading < 0 ? -1.0 : 0.0)) * prm->max_delta_heading;
        pref_speed = 0.;
      }
      act[0] = pref_speed;
      act[1] = delta_heading;
    }
  }
  return NULL;
}

int casim_oracle_rvo_actions_mt(const casim_config* cfg, const casim_rvo_params* prm, int E, int A, const int32_t* n_agents,
                                const double* state, const double* coop, double* actions, int n_threads) {
  if (A > MAXN) return 1;
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  rvo_job jobs[64];
  pthread_t th[64];
  for (int t = 0; t < n_threads; t++) {
    rvo_job* j = &jobs[t];
    j->cfg = cfg; j->prm = prm; j->A = A; j->n_agents = n_agents; j->state = state; j->coop = coop; j->actions = actions;
    j->e0 = (int)((long long)E * t / n_threads);
    j->e1 = (int)((long long)E * (t + 1) / n_threads);
    if (n_threads == 1) rvo_job_main(j); else pthread_create(&th[t], NULL, rvo_job_main, j);
  }
  if (n_threads > 1) for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
  return 0;
}

int casim_oracle_rvo_actions(const casim_config* cfg, const casim_rvo_params* prm, int E, int A, const int32_t* n_agents,
                             const double* state, const double* coop, double* actions) {
  return casim_oracle_rvo_actions_mt(cfg, prm, E, A, n_agents, state, coop, actions, 1);
}
