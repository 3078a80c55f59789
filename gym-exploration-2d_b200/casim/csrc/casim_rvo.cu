/*
 * ATTRIBUTION.  The ORCA-line construction and linearProgram1/2/3 below restate, operation for operation, the RVO2
 * Library (Python-RVO2/src/Agent.cpp, vendored by the reference):
 *   Copyright 2008 University of North Carolina at Chapel Hill -- Jur van den Berg, Stephen J. Guy, Jamie Snape,
 *   Ming C. Lin, Dinesh Manocha -- Licensed under the Apache License, Version 2.0
 *   (http://www.apache.org/licenses/LICENSE-2.0), <http://gamma.cs.unc.edu/RVO2/>.  See NOTICE at the repository root.
 * Bit-identical new velocities require the library's binary32 operation order, so that arithmetic is the library's
 * by necessity; the GPU mapping around it is this repository's.
 */
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
 * Static obstacles: addObstacle + buildObstacleTree (RVOSimulator.cpp:124-161, KdTree.cpp:113-245) run on the host in
 * casim_rvo_set_obstacles; k_rvo<true> queries the tree (KdTree.cpp:300-329) and emits the obstacle ORCA lines
 * (Agent.cpp:66-284).
 * Neighbour ties at exactly equal squared distance are broken by agent index (reference: k-d tree visiting order).
 *
 * Mapping note (measured on B200, C2-shaped all-RVO batch of 40 960 agents): one thread per agent takes 22.8 us;
 * a sub-warp cooperative variant (16 lanes per agent, one ORCA line per lane, ballot / shuffle linear programs --
 * kept for reference in profiles/parked/casim_rvo_subwarp16.cu.txt, parity-green) took 30.8 us: it issues 2.3x more warp
 * instructions (23 M vs ~10 M) and the kernel is issue-bound once enough warps are resident.
 *
 * Bound: latency / fp32 issue (a few hundred dependent flops per neighbour, data-dependent linear programs); the
 * HBM traffic is 48 B read + 8 B written per agent (neighbour reads hit L1/L2 within the environment).
 */
#include <algorithm>
#include <utility>
#include <vector>

#include "casim_device.cuh"

namespace {

#define RVO_EPS 0.00001f
#define RVO_MAXN CASIM_MAX_AGENTS

#define RVO_MAX_OBST 64 /* obstacle vertices after edge splitting (the crossing map has 16) */

struct RvoObst { /* device image of the processed obstacle set, struct of arrays */
  const float2* point;   /* Obstacle::point_ */
  const float2* unitDir; /* Obstacle::unitDir_ */
  const int* next;       /* nextObstacle_ */
  const int* prev;       /* prevObstacle_ */
  const int* convex;     /* isConvex_ */
  const int4* node;      /* obstacle tree: (obstacle, left, right, -) */
  int n, root;
};

struct RvoArgs {
  RvoObst ob;
  float time_horizon_obst;
  const float4* state; /* [6][E*A] */
  const float4* state0;          /* episode-start state kept by casim_reset, or NULL */
  const uint8_t* prev_game_over; /* [E] or NULL: with auto_reset, an environment that ended last step restarts inside the
                                    coming casim_step from state0 -- the policy must read that state, like k_core does */
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
__device__ __forceinline__ float leftOf(float2 a, float2 b, float2 c) { return det(sub(a, c), sub(b, a)); } /* Definitions.h:93 */

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

/* Agent.cpp:580-630 */
__device__ void lp3(const Line* lines, int nl, int numObst, int beginLine, float radius, float2& result, Line* proj) {
  float distance = 0.0f;
  for (int i = beginLine; i < nl; i++) {
    if (det(lines[i].direction, sub(lines[i].point, result)) > distance) {
      int np = 0;
      for (int j = 0; j < numObst; j++) proj[np++] = lines[j];
      for (int j = numObst; j < i; j++) {
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

/* Definitions.h:68-82 */
__device__ __forceinline__ float distSqPointLineSegment(float2 a, float2 b, float2 c) {
  const float r = dot(sub(c, a), sub(b, a)) / absSq(sub(b, a));
  if (r < 0.0f) return absSq(sub(c, a));
  if (r > 1.0f) return absSq(sub(c, b));
  return absSq(sub(c, add(a, mulsv(r, sub(b, a)))));
}

/* KdTree::queryObstacleTreeRecursive (KdTree.cpp:300-329) with an explicit stack + Agent::insertObstacleNeighbor
 * (Agent.cpp:475-493): obstacle edges within rangeSq that the agent sees from their right side, ascending distance. */
__device__ int obstacle_neighbors(const RvoObst& O, float2 position, float rangeSq, float* nd, int* no) {
  int nn = 0;
  int stack[2 * RVO_MAX_OBST + 2]; /* node << 1 | stage */
  int sp = 0;
  if (O.root >= 0) stack[sp++] = O.root << 1;
  while (sp > 0) {
    const int en = stack[--sp];
    const int4 nd4 = O.node[en >> 1];
    const int o1 = nd4.x, o2 = O.next[o1];
    const float2 p1 = O.point[o1], p2 = O.point[o2];
    const float agentLeftOfLine = leftOf(p1, p2, position);
    const int first = agentLeftOfLine >= 0.0f ? nd4.y : nd4.z, second = agentLeftOfLine >= 0.0f ? nd4.z : nd4.y;
    if ((en & 1) == 0) { /* descend into the agent's side first, come back to this node afterwards */
      stack[sp++] = en | 1;
      if (first >= 0) stack[sp++] = first << 1;
    } else {
      const float distSqLine = sqr(agentLeftOfLine) / absSq(sub(p2, p1));
      if (distSqLine < rangeSq) {
        if (agentLeftOfLine < 0.0f) {
          const float distSq = distSqPointLineSegment(p1, p2, position);
          if (distSq < rangeSq && nn < RVO_MAX_OBST) {
            int i = nn++;
            while (i != 0 && distSq < nd[i - 1]) {
              nd[i] = nd[i - 1];
              no[i] = no[i - 1];
              --i;
            }
            nd[i] = distSq;
            no[i] = o1;
          }
        }
        if (second >= 0) stack[sp++] = second << 1;
      }
    }
  }
  return nn;
}

/* Agent.cpp:66-284: ORCA lines of the obstacle neighbours */
__device__ int obstacle_lines(const RvoObst& O, const int* no, int nn, float2 position_, float2 velocity_, float radius_,
                              float timeHorizonObst_, Line* lines) {
  int nl = 0;
  const float invTimeHorizonObst = 1.0f / timeHorizonObst_;
  for (int i = 0; i < nn; i++) {
    int obstacle1 = no[i];
    int obstacle2 = O.next[obstacle1];
    const float2 P1 = O.point[obstacle1], P2 = O.point[obstacle2];
    const float2 relativePosition1 = sub(P1, position_);
    const float2 relativePosition2 = sub(P2, position_);
    bool alreadyCovered = false;
    for (int j = 0; j < nl; j++) {
      if (det(sub(mulsv(invTimeHorizonObst, relativePosition1), lines[j].point), lines[j].direction) - invTimeHorizonObst * radius_ >= -RVO_EPS &&
          det(sub(mulsv(invTimeHorizonObst, relativePosition2), lines[j].point), lines[j].direction) - invTimeHorizonObst * radius_ >= -RVO_EPS) {
        alreadyCovered = true;
        break;
      }
    }
    if (alreadyCovered) continue;
    const float distSq1 = absSq(relativePosition1);
    const float distSq2 = absSq(relativePosition2);
    const float radiusSq = sqr(radius_);
    const float2 obstacleVector = sub(P2, P1);
    const float s = dot(neg(relativePosition1), obstacleVector) / absSq(obstacleVector);
    const float distSqLine = absSq(sub(neg(relativePosition1), mulsv(s, obstacleVector)));
    const float2 unitDir1 = O.unitDir[obstacle1], unitDir2o = O.unitDir[obstacle2];
    const bool convex1 = O.convex[obstacle1] != 0, convex2o = O.convex[obstacle2] != 0;
    Line line;
    if (s < 0.0f && distSq1 <= radiusSq) {
      if (convex1) {
        line.point = V(0.0f, 0.0f);
        line.direction = normalize(V(-relativePosition1.y, relativePosition1.x));
        lines[nl++] = line;
      }
      continue;
    } else if (s > 1.0f && distSq2 <= radiusSq) {
      if (convex2o && det(relativePosition2, unitDir2o) >= 0.0f) {
        line.point = V(0.0f, 0.0f);
        line.direction = normalize(V(-relativePosition2.y, relativePosition2.x));
        lines[nl++] = line;
      }
      continue;
    } else if (s >= 0.0f && s < 1.0f && distSqLine <= radiusSq) {
      line.point = V(0.0f, 0.0f);
      line.direction = neg(unitDir1);
      lines[nl++] = line;
      continue;
    }
    float2 leftLegDirection, rightLegDirection;
    if (s < 0.0f && distSqLine <= radiusSq) {
      if (!convex1) continue;
      obstacle2 = obstacle1;
      const float leg1 = sqrtf(distSq1 - radiusSq);
      leftLegDirection = divs(V(relativePosition1.x * leg1 - relativePosition1.y * radius_, relativePosition1.x * radius_ + relativePosition1.y * leg1), distSq1);
      rightLegDirection = divs(V(relativePosition1.x * leg1 + relativePosition1.y * radius_, -relativePosition1.x * radius_ + relativePosition1.y * leg1), distSq1);
    } else if (s > 1.0f && distSqLine <= radiusSq) {
      if (!convex2o) continue;
      obstacle1 = obstacle2;
      const float leg2 = sqrtf(distSq2 - radiusSq);
      leftLegDirection = divs(V(relativePosition2.x * leg2 - relativePosition2.y * radius_, relativePosition2.x * radius_ + relativePosition2.y * leg2), distSq2);
      rightLegDirection = divs(V(relativePosition2.x * leg2 + relativePosition2.y * radius_, -relativePosition2.x * radius_ + relativePosition2.y * leg2), distSq2);
    } else {
      if (convex1) {
        const float leg1 = sqrtf(distSq1 - radiusSq);
        leftLegDirection = divs(V(relativePosition1.x * leg1 - relativePosition1.y * radius_, relativePosition1.x * radius_ + relativePosition1.y * leg1), distSq1);
      } else {
        leftLegDirection = neg(unitDir1);
      }
      if (convex2o) {
        const float leg2 = sqrtf(distSq2 - radiusSq);
        rightLegDirection = divs(V(relativePosition2.x * leg2 + relativePosition2.y * radius_, -relativePosition2.x * radius_ + relativePosition2.y * leg2), distSq2);
      } else {
        rightLegDirection = unitDir1;
      }
    }
    /* obstacle1 / obstacle2 may have been merged above: re-read their attributes */
    const float2 uD1 = O.unitDir[obstacle1], uD2 = O.unitDir[obstacle2];
    const bool cv1 = O.convex[obstacle1] != 0, cv2 = O.convex[obstacle2] != 0;
    const float2 uLeftNb = O.unitDir[O.prev[obstacle1]];
    bool isLeftLegForeign = false, isRightLegForeign = false;
    if (cv1 && det(leftLegDirection, neg(uLeftNb)) >= 0.0f) {
      leftLegDirection = neg(uLeftNb);
      isLeftLegForeign = true;
    }
    if (cv2 && det(rightLegDirection, uD2) <= 0.0f) {
      rightLegDirection = uD2;
      isRightLegForeign = true;
    }
    const float2 leftCutoff = mulsv(invTimeHorizonObst, sub(O.point[obstacle1], position_));
    const float2 rightCutoff = mulsv(invTimeHorizonObst, sub(O.point[obstacle2], position_));
    const float2 cutoffVec = sub(rightCutoff, leftCutoff);
    const bool same = obstacle1 == obstacle2;
    const float t = same ? 0.5f : dot(sub(velocity_, leftCutoff), cutoffVec) / absSq(cutoffVec);
    const float tLeft = dot(sub(velocity_, leftCutoff), leftLegDirection);
    const float tRight = dot(sub(velocity_, rightCutoff), rightLegDirection);
    if ((t < 0.0f && tLeft < 0.0f) || (same && tLeft < 0.0f && tRight < 0.0f)) {
      const float2 unitW = normalize(sub(velocity_, leftCutoff));
      line.direction = V(unitW.y, -unitW.x);
      line.point = add(leftCutoff, mulsv(radius_ * invTimeHorizonObst, unitW));
      lines[nl++] = line;
      continue;
    } else if (t > 1.0f && tRight < 0.0f) {
      const float2 unitW = normalize(sub(velocity_, rightCutoff));
      line.direction = V(unitW.y, -unitW.x);
      line.point = add(rightCutoff, mulsv(radius_ * invTimeHorizonObst, unitW));
      lines[nl++] = line;
      continue;
    }
    const float distSqCutoff = (t < 0.0f || t > 1.0f || same) ? INFINITY : absSq(sub(velocity_, add(leftCutoff, mulsv(t, cutoffVec))));
    const float distSqLeft = tLeft < 0.0f ? INFINITY : absSq(sub(velocity_, add(leftCutoff, mulsv(tLeft, leftLegDirection))));
    const float distSqRight = tRight < 0.0f ? INFINITY : absSq(sub(velocity_, add(rightCutoff, mulsv(tRight, rightLegDirection))));
    if (distSqCutoff <= distSqLeft && distSqCutoff <= distSqRight) {
      line.direction = neg(uD1);
      line.point = add(leftCutoff, mulsv(radius_ * invTimeHorizonObst, V(-line.direction.y, line.direction.x)));
      lines[nl++] = line;
    } else if (distSqLeft <= distSqRight) {
      if (isLeftLegForeign) continue;
      line.direction = leftLegDirection;
      line.point = add(leftCutoff, mulsv(radius_ * invTimeHorizonObst, V(-line.direction.y, line.direction.x)));
      lines[nl++] = line;
    } else {
      if (isRightLegForeign) continue;
      line.direction = neg(rightLegDirection);
      line.point = add(rightCutoff, mulsv(radius_ * invTimeHorizonObst, V(-line.direction.y, line.direction.x)));
      lines[nl++] = line;
    }
  }
  return nl;
}

/* Agent.cpp:298-364: the ORCA half-plane one neighbour induces on the ego agent's velocity (o0 = px, py, ..; o1 = vx, vy,
 * ..; o2 = .., .., radius, ..).  Binary32, the library's operation order. */
__device__ __forceinline__ Line agent_orca_line(float2 position, float2 velocity, float radius_, float4 o0, float4 o1, float4 o2,
                                                double radius_scale, float invTimeHorizon, float time_step, float collab) {
  const float2 relativePosition = sub(V(o0.x, o0.y), position);
  const float2 relativeVelocity = sub(velocity, V(o1.x, o1.y));
  const float distSq = absSq(relativePosition);
  const float combinedRadius = radius_ + (float)(radius_scale * (double)o2.z);
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
    const float invTimeStep = 1.0f / time_step;
    const float2 w = sub(relativeVelocity, mulsv(invTimeStep, relativePosition));
    const float wLength = sqrtf(dot(w, w));
    const float2 unitW = divs(w, wLength);
    line.direction = V(unitW.y, -unitW.x);
    u = mulsv(combinedRadius * invTimeStep - wLength, unitW);
  }
  line.point = add(velocity, mulsv(collab, u)); /* Agent.cpp:361 */
  return line;
}

#ifndef RVO_THREADS
#define RVO_THREADS 128
#endif

template <bool OBST>
__global__ void __launch_bounds__(RVO_THREADS) k_rvo(const RvoArgs p) {
  casim_pdl_enter();
  const size_t N = (size_t)p.E * p.A;
  const size_t g = (size_t)blockIdx.x * RVO_THREADS + threadIdx.x;
  if (g >= N) return;
  const int A = p.A;
  const int e = (int)(g / A), ego = (int)(g - (size_t)e * A);
  const int n = p.n_agents ? min(p.n_agents[e], A) : A;
  if (ego >= n) return;
  const float4* s0 = ((p.prev_game_over && p.prev_game_over[e]) ? p.state0 : p.state) + (size_t)e * A; /* px, py, heading, speed */
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

  constexpr int NOB = OBST ? RVO_MAX_OBST : 0;
  Line lines[NOB + RVO_MAXN + 2], proj[NOB + RVO_MAXN + 2];
  int nl = 0;
  if (OBST) { /* Agent.cpp:45-47 (obstacle neighbours) + 66-284 (their ORCA lines come first) */
    float od[RVO_MAX_OBST];
    int oo[RVO_MAX_OBST];
    const int on = obstacle_neighbors(p.ob, position, sqr(p.time_horizon_obst * maxSpeed + radius_), od, oo);
    nl = obstacle_lines(p.ob, oo, on, position, velocity, radius_, p.time_horizon_obst, lines);
  }
  const int numObstLines = nl;
  const float invTimeHorizon = 1.0f / p.time_horizon;
  if (collab == 0.0f) {
    /* Agent.cpp:290-292 */
  } else if (collab > 0.0f) {
    for (int k = 0; k < nn; k++) { /* Agent.cpp:298-364 */
      const int o = nidx[k];
      lines[nl++] = agent_orca_line(position, velocity, radius_, s0[o], s1[o], s2[o], p.radius_scale, invTimeHorizon,
                                    p.time_step, collab);
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
  if (lineFail < nl) lp3(lines, nl, numObstLines, lineFail, maxSpeed, newVel, proj);
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


/* ------------------------------------------------------------------------------------------------------------------
 * k_rvo4 -- the agents-only case (no static obstacles) mapped onto the GPU: R4_LPA (8) LANES PER EGO AGENT, several agents per warp.
 * One thread per agent leaves the machine with ~9 warps per SM and a serial chain of a few thousand dependent fp32
 * operations per thread (52.9 us for 40 960 agents).  Here the chain is cut in R4_LPA pieces and there are that many times as many
 * warps:
 *   neighbours   lane q evaluates the candidates j = q, q+4, ..; the reference's "sorted insertion with the range
 *                shrinking to the farthest kept neighbour" (Agent.cpp:449-473) is a stable rank by (distSq, index)
 *                cut at maxNeighbors, so each lane RANKS its candidates by counting -- no serial sort;
 *   ORCA lines   built by the lane that owns the candidate (agent_orca_line), stored at slot = rank in shared memory;
 *   LP 2         "the next violated line at or after i" is a 4-lane min over per-lane scans (the result only changes
 *                at a repair); LP 1's scan over the earlier lines only tightens tLeft (max) / tRight (min) and fails
 *                on an infeasible parallel line, so each lane scans a quarter of them and the group combines with
 *                min / max / or -- exact, order-free (tLeft never decreases and tRight never increases, hence "some
 *                prefix has tLeft > tRight" == "the final one has");
 *   LP 3         the projected lines of one violated line are built four at a time and compacted in order with a
 *                group ballot, then LP 2 runs on them.
 * All lanes of a group carry the same scalars (result, tLeft, ...), computed redundantly from the same inputs, so every
 * branch on them is uniform in the group; groups of one warp run independently (group masks on every shuffle).
 * The arithmetic per (line, line) pair is the library's binary32 sequence: new velocities are bit-identical to it. */
#define R4_MAXL 32 /* ORCA lines per agent (A - 1 <= 31) */
#define R4_THREADS 128
#ifndef R4_LPA
#define R4_LPA 8 /* lanes per ego agent (power of two <= 32).  Measured at 4096 x 10 (profiles/sweep_rvo_lanes.sh): 4 lanes 38.7 us, 8 lanes 36.6 us, 16 lanes 46.1 us; one thread per agent 52.9 us */
#endif
#define R4_GROUPS (R4_THREADS / R4_LPA)

struct R4Group {
  unsigned mask; /* the lanes of my group */
  int q;         /* my lane inside the group */
  __device__ __forceinline__ float gminf(float v) const {
#pragma unroll
    for (int o = 1; o < R4_LPA; o <<= 1) {
      const float t = __shfl_xor_sync(mask, v, o);
      v = (t < v) ? t : v;
    }
    return v;
  }
  __device__ __forceinline__ float gmaxf(float v) const {
#pragma unroll
    for (int o = 1; o < R4_LPA; o <<= 1) {
      const float t = __shfl_xor_sync(mask, v, o);
      v = (v < t) ? t : v;
    }
    return v;
  }
  __device__ __forceinline__ int gmini(int v) const {
#pragma unroll
    for (int o = 1; o < R4_LPA; o <<= 1) v = min(v, __shfl_xor_sync(mask, v, o));
    return v;
  }
  __device__ __forceinline__ int gsumi(int v) const {
#pragma unroll
    for (int o = 1; o < R4_LPA; o <<= 1) v += __shfl_xor_sync(mask, v, o);
    return v;
  }
  __device__ __forceinline__ unsigned gballot(bool b, int lane) const {
    return (__ballot_sync(mask, b) >> (lane & ~(R4_LPA - 1))) & (R4_LPA == 32 ? 0xFFFFFFFFu : ((1u << R4_LPA) - 1u));
  }
  __device__ __forceinline__ void sync() const { __syncwarp(mask); }
};

__device__ __forceinline__ Line r4_load(const float4* L, int i) {
  const float4 v = L[i];
  Line l;
  l.point = V(v.x, v.y);
  l.direction = V(v.z, v.w);
  return l;
}

/* Agent.cpp:477-546 with the scan over the earlier lines split over the four lanes */
__device__ bool r4_lp1(const R4Group& G, const float4* lines, int lineNo, float radius, float2 opt, bool dirOpt, float2& result) {
  const Line L = r4_load(lines, lineNo);
  const float dotProduct = dot(L.point, L.direction);
  const float discriminant = sqr(dotProduct) + sqr(radius) - absSq(L.point);
  if (discriminant < 0.0f) return false;
  const float sq = sqrtf(discriminant);
  float tLeft = -dotProduct - sq;
  float tRight = -dotProduct + sq;
  bool infeasible = false;
  for (int i = G.q; i < lineNo; i += R4_LPA) {
    const Line Li = r4_load(lines, i);
    const float denominator = det(L.direction, Li.direction);
    const float numerator = det(Li.direction, sub(L.point, Li.point));
    if (fabsf(denominator) <= RVO_EPS) {
      if (numerator < 0.0f) infeasible = true;
      continue;
    }
    const float t = numerator / denominator;
    if (denominator >= 0.0f)
      tRight = (t < tRight) ? t : tRight;
    else
      tLeft = (tLeft < t) ? t : tLeft;
  }
  tRight = G.gminf(tRight);
  tLeft = G.gmaxf(tLeft);
  if (__any_sync(G.mask, infeasible) || tLeft > tRight) return false;
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
__device__ int r4_lp2(const R4Group& G, const float4* lines, int nl, float radius, float2 opt, bool dirOpt, float2& result) {
  if (dirOpt)
    result = mulvs(opt, radius);
  else if (absSq(opt) > sqr(radius))
    result = mulvs(normalize(opt), radius);
  else
    result = opt;
  int i = 0;
  while (i < nl) {
    /* the first violated line at or after i: each lane scans its own lines (k = q mod 4) */
    int k = i + ((G.q - i) & (R4_LPA - 1));
    for (; k < nl; k += R4_LPA) {
      const Line Lk = r4_load(lines, k);
      if (det(Lk.direction, sub(Lk.point, result)) > 0.0f) break;
    }
    const int v = G.gmini(k);
    if (v >= nl) break;
    const float2 temp = result;
    if (!r4_lp1(G, lines, v, radius, opt, dirOpt, result)) {
      result = temp;
      return v;
    }
    i = v + 1;
  }
  return nl;
}

/* Agent.cpp:580-630 (no obstacle lines in this kernel: numObstLines = 0) */
__device__ void r4_lp3(const R4Group& G, int lane, const float4* lines, int nl, int beginLine, float radius, float2& result,
                       float4* proj) {
  float distance = 0.0f;
  for (int i = beginLine; i < nl; i++) {
    const Line Li = r4_load(lines, i);
    if (det(Li.direction, sub(Li.point, result)) > distance) {
      int np = 0;
      for (int j0 = 0; j0 < i; j0 += R4_LPA) { /* one projected line per lane and round, compacted in order */
        const int j = j0 + G.q;
        bool keep = false;
        Line l;
        l.point = V(0.f, 0.f);
        l.direction = V(0.f, 0.f);
        if (j < i) {
          const Line Lj = r4_load(lines, j);
          const float determinant = det(Li.direction, Lj.direction);
          keep = true;
          if (fabsf(determinant) <= RVO_EPS) {
            if (dot(Li.direction, Lj.direction) > 0.0f)
              keep = false;
            else
              l.point = mulsv(0.5f, add(Li.point, Lj.point));
          } else {
            l.point = add(Li.point, mulsv(det(Lj.direction, sub(Li.point, Lj.point)) / determinant, Li.direction));
          }
          if (keep) l.direction = normalize(sub(Lj.direction, Li.direction));
        }
        const unsigned b = G.gballot(keep, lane);
        if (keep) proj[np + __popc(b & ((1u << G.q) - 1u))] = make_float4(l.point.x, l.point.y, l.direction.x, l.direction.y);
        np += __popc(b);
      }
      G.sync();
      const float2 temp = result;
      if (r4_lp2(G, proj, np, radius, V(-Li.direction.y, Li.direction.x), true, result) < np) result = temp;
      distance = det(Li.direction, sub(Li.point, result));
      G.sync(); /* proj is rebuilt for the next violated line */
    }
  }
}


__global__ void __launch_bounds__(R4_THREADS) k_rvo4(const RvoArgs p) {
  __shared__ __align__(16) float4 s_line[R4_GROUPS][R4_MAXL];
  __shared__ __align__(16) float4 s_proj[R4_GROUPS][R4_MAXL];
  __shared__ float s_cd[R4_GROUPS][R4_MAXL];
  casim_pdl_enter();
  const size_t N = (size_t)p.E * p.A;
  const int lane = threadIdx.x & 31, grp = threadIdx.x / R4_LPA;
  const size_t g = (size_t)blockIdx.x * R4_GROUPS + grp;
  if (g >= N) return; /* whole groups leave together */
  R4Group G;
  G.q = lane & (R4_LPA - 1);
  G.mask = (R4_LPA == 32 ? 0xFFFFFFFFu : ((1u << R4_LPA) - 1u)) << (lane & ~(R4_LPA - 1));
  const int A = p.A;
  const int e = (int)(g / A), ego = (int)(g - (size_t)e * A);
  const int n = p.n_agents ? min(p.n_agents[e], A) : A;
  if (ego >= n) return;
  const float4* s0 = ((p.prev_game_over && p.prev_game_over[e]) ? p.state0 : p.state) + (size_t)e * A; /* px, py, heading, speed */
  const float4* s1 = s0 + N;                                                                              /* vx, vy, ... */
  const float4* s2 = s1 + N;                                                                              /* gx, gy, radius, pref_speed */
  const float4* s5 = s2 + 3 * N;
  if ((__float_as_int(s5[ego].w) & 0xff) != CASIM_POLICY_RVO) return;
  float4* lines = s_line[grp];
  float* cd = s_cd[grp];

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

  /* ---- neighbours (Agent.cpp:43-56, 449-473): squared distances, then a stable rank by counting ---- */
  const float rangeSq = sqr(p.neighbor_dist);
  int mine = 0;
  for (int j = G.q; j < n; j += R4_LPA) {
    float d = INFINITY;
    if (j != ego && p.max_neighbors > 0) {
      const float4 o0 = s0[j];
      const float distSq = absSq(sub(position, V(o0.x, o0.y)));
      if (distSq < rangeSq) {
        d = distSq;
        mine++;
      }
    }
    cd[j] = d;
  }
  const int nn = min(G.gsumi(mine), p.max_neighbors);
  G.sync();
  const float invTimeHorizon = 1.0f / p.time_horizon;
  int nl = 0;
  int j_closest = INT_MAX; /* the candidate of rank 0 */
  for (int j = G.q; j < n; j += R4_LPA) {
    const float d = cd[j];
    if (!(d < INFINITY)) continue;
    int rank = 0;
    for (int k = 0; k < n; k++) {
      const float dk = cd[k];
      rank += (dk < d || (dk == d && k < j)) ? 1 : 0;
    }
    if (rank == 0) j_closest = j;
    if (rank < nn && collab > 0.0f) { /* Agent.cpp:298-364, slot = position in the sorted neighbour list */
      const Line l = agent_orca_line(position, velocity, radius_, s0[j], s1[j], s2[j], p.radius_scale, invTimeHorizon, p.time_step,
                                     collab);
      lines[rank] = make_float4(l.point.x, l.point.y, l.direction.x, l.direction.y);
    }
  }
  if (collab > 0.0f) nl = nn;
  j_closest = G.gmini(j_closest);
  if (collab < 0.0f && nn > 0) { /* Agent.cpp:367-416: two lines around the closest neighbour (every lane the same values) */
    const int o = j_closest;
    const float4 o0 = s0[o], o1 = s1[o], o2 = s2[o];
    const float2 relativePosition = sub(V(o0.x, o0.y), position);
    const float2 ovel = V(o1.x, o1.y);
    const float distSq = absSq(relativePosition);
    const float combinedRadius = radius_ + (float)(p.radius_scale * (double)o2.z);
    const float combinedRadiusSq = sqr(combinedRadius);
    const float leg = sqrtf(distSq - combinedRadiusSq);
    const float2 d1 = divs(neg(V(relativePosition.x * leg - relativePosition.y * combinedRadius,
                                 relativePosition.x * combinedRadius + relativePosition.y * leg)), distSq);
    const float2 d2 = divs(V(relativePosition.x * leg + relativePosition.y * combinedRadius,
                             -relativePosition.x * combinedRadius + relativePosition.y * leg), distSq);
    const float2 pt = add(ovel, relativePosition);
    if (G.q == 0) {
      lines[0] = make_float4(pt.x, pt.y, d1.x, d1.y);
      lines[1] = make_float4(pt.x, pt.y, d2.x, d2.y);
    }
    nl = 2;
    prefVel = add(ovel, mulsv(-collab, normalize(relativePosition)));
  }
  G.sync();

  float2 newVel = V(0.f, 0.f);
  const int lineFail = r4_lp2(G, lines, nl, maxSpeed, prefVel, false, newVel);
  if (lineFail < nl) r4_lp3(G, lane, lines, nl, lineFail, maxSpeed, newVel, s_proj[grp]);
  if (G.q != 0) return;
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

/* ---- host side: RVOSimulator::addObstacle (RVOSimulator.cpp:124-161) + KdTree::buildObstacleTree (KdTree.cpp:113-245),
 * run once per casim_rvo_set_obstacles; the processed set is uploaded as flat arrays ---- */
namespace {

struct HostObstacles {
  std::vector<float2> point, unitDir;
  std::vector<int> next, prev, convex;
  std::vector<int4> node;

  static float2 hsub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
  static float hdet(float2 a, float2 b) { return a.x * b.y - a.y * b.x; }
  static float hleft(float2 a, float2 b, float2 c) { return hdet(hsub(a, c), hsub(b, a)); }
  static float2 hnorm(float2 a) {
    const float inv = 1.0f / sqrtf(a.x * a.x + a.y * a.y);
    return make_float2(a.x * inv, a.y * inv);
  }

  void add_polygon(const std::vector<float2>& v) {
    const int nv = (int)v.size(), first = (int)point.size();
    for (int i = 0; i < nv; i++) {
      const int k = first + i;
      point.push_back(v[i]);
      unitDir.push_back(hnorm(hsub(v[(i + 1) % nv], v[i])));
      next.push_back(i == nv - 1 ? first : k + 1);
      prev.push_back(i == 0 ? first + nv - 1 : k - 1);
      convex.push_back(nv == 2 ? 1 : (hleft(v[(i + nv - 1) % nv], v[i], v[(i + 1) % nv]) >= 0.0f ? 1 : 0));
    }
  }

  /* side of edge j relative to the line through edge i: +1 left, -1 right, 0 straddles */
  int side(int i1, int j1, float& j1Left) const {
    const int i2 = next[i1], j2 = next[j1];
    j1Left = hleft(point[i1], point[i2], point[j1]);
    const float j2Left = hleft(point[i1], point[i2], point[j2]);
    if (j1Left >= -RVO_EPS && j2Left >= -RVO_EPS) return 1;
    if (j1Left <= RVO_EPS && j2Left <= RVO_EPS) return -1;
    return 0;
  }

  int build(const std::vector<int>& list) {
    if (list.empty()) return -1;
    const size_t cnt = list.size();
    /* the split edge minimises (max(left, right), min(left, right)) lexicographically; first best wins */
    size_t best = 0, bestL = cnt, bestR = cnt;
    for (size_t i = 0; i < cnt; i++) {
      size_t l = 0, r = 0;
      const std::pair<size_t, size_t> bound(std::max(bestL, bestR), std::min(bestL, bestR));
      for (size_t j = 0; j < cnt; j++) {
        if (i == j) continue;
        float dummy;
        const int sd = side(list[i], list[j], dummy);
        if (sd >= 0) ++l;
        if (sd <= 0) ++r;
        if (std::make_pair(std::max(l, r), std::min(l, r)) >= bound) break;
      }
      if (std::make_pair(std::max(l, r), std::min(l, r)) < bound) {
        best = i;
        bestL = l;
        bestR = r;
      }
    }
    std::vector<int> L, R;
    const int i1 = list[best];
    for (size_t j = 0; j < cnt; j++) {
      if (j == best) continue;
      const int j1 = list[j];
      float j1Left;
      const int sd = side(i1, j1, j1Left);
      if (sd > 0) L.push_back(j1);
      else if (sd < 0) R.push_back(j1);
      else { /* edge j crosses the line: a new vertex splits it */
        const int i2 = next[i1], j2 = next[j1];
        const float t = hdet(hsub(point[i2], point[i1]), hsub(point[j1], point[i1])) /
                        hdet(hsub(point[i2], point[i1]), hsub(point[j1], point[j2]));
        const float2 d = hsub(point[j2], point[j1]);
        const int nw = (int)point.size();
        point.push_back(make_float2(point[j1].x + t * d.x, point[j1].y + t * d.y));
        unitDir.push_back(unitDir[j1]);
        prev.push_back(j1);
        next.push_back(j2);
        convex.push_back(1);
        next[j1] = nw;
        prev[j2] = nw;
        if (j1Left > 0.0f) { L.push_back(j1); R.push_back(nw); }
        else { R.push_back(j1); L.push_back(nw); }
      }
    }
    const int me = (int)node.size();
    node.push_back(make_int4(i1, -1, -1, 0));
    const int l = build(L);
    const int r = build(R);
    node[me].y = l;
    node[me].z = r;
    return me;
  }
};

template <typename T>
cudaError_t upload(T** dst, const std::vector<T>& v) {
  cudaFree(*dst);
  *dst = nullptr;
  if (v.empty()) return cudaSuccess;
  cudaError_t e = cudaMalloc((void**)dst, v.size() * sizeof(T));
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*dst, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
}

}  // namespace

int casim_rvo_obstacles_build(casim_handle* h, const double* verts, const int32_t* poly_sizes, int n_polys) {
  HostObstacles H;
  int off = 0;
  for (int p = 0; p < n_polys; p++) {
    if (poly_sizes[p] < 2) return -1;
    std::vector<float2> v;
    for (int k = 0; k < poly_sizes[p]; k++) v.push_back(make_float2((float)verts[2 * (off + k)], (float)verts[2 * (off + k) + 1]));
    H.add_polygon(v);
    off += poly_sizes[p];
  }
  std::vector<int> all(H.point.size());
  for (size_t k = 0; k < all.size(); k++) all[k] = (int)k;
  const int root = H.build(all);
  if ((int)H.point.size() > RVO_MAX_OBST) return -2;
  cudaError_t e = upload(&h->rvo_point, H.point);
  if (e == cudaSuccess) e = upload(&h->rvo_unitdir, H.unitDir);
  if (e == cudaSuccess) e = upload(&h->rvo_next, H.next);
  if (e == cudaSuccess) e = upload(&h->rvo_prev, H.prev);
  if (e == cudaSuccess) e = upload(&h->rvo_convex, H.convex);
  if (e == cudaSuccess) e = upload(&h->rvo_node, H.node);
  if (e != cudaSuccess) return (int)e;
  h->rvo_n = (int)H.point.size();
  h->rvo_root = n_polys > 0 ? root : -1;
  return 0;
}

void casim_rvo_obstacles_free(casim_handle* h) {
  cudaFree(h->rvo_point); cudaFree(h->rvo_unitdir); cudaFree(h->rvo_next); cudaFree(h->rvo_prev);
  cudaFree(h->rvo_convex); cudaFree(h->rvo_node);
  h->rvo_point = h->rvo_unitdir = nullptr;
  h->rvo_next = h->rvo_prev = h->rvo_convex = nullptr;
  h->rvo_node = nullptr;
  h->rvo_n = 0;
  h->rvo_root = -1;
}

cudaError_t casim_launch_rvo(casim_handle* h, const float* state, const float* coop, const casim_rvo_params* prm,
                             float* policy_actions, cudaStream_t s) {
  RvoArgs a;
  a.ob.point = h->rvo_point;
  a.ob.unitDir = h->rvo_unitdir;
  a.ob.next = h->rvo_next;
  a.ob.prev = h->rvo_prev;
  a.ob.convex = h->rvo_convex;
  a.ob.node = h->rvo_node;
  a.ob.n = h->rvo_n;
  a.ob.root = h->rvo_n > 0 ? h->rvo_root : -1;
  a.time_horizon_obst = (float)prm->time_horizon; /* RVOPolicy.py:26-27: timeHorizonObst = time_horizon */
  a.state = reinterpret_cast<const float4*>(state);
  a.state0 = reinterpret_cast<const float4*>(h->state0);
  a.prev_game_over = (h->dc.auto_reset && h->init) ? h->prev_game_over : nullptr;
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
  const dim3 grid((unsigned)((N + RVO_THREADS - 1) / RVO_THREADS)), block(RVO_THREADS);
  if (a.ob.root >= 0) return casim_launch_pdl<RvoArgs>(k_rvo<true>, grid, block, 0, s, a);
  static const int one_thread = casim_tuning_int("CASIM_RVO_THREAD_PER_AGENT", 0);
  if (one_thread) return casim_launch_pdl<RvoArgs>(k_rvo<false>, grid, block, 0, s, a);
  /* agents only: four lanes per ego agent */
  return casim_launch_pdl<RvoArgs>(k_rvo4, dim3((unsigned)((N + R4_GROUPS - 1) / R4_GROUPS)), dim3(R4_THREADS), 0, s, a);
}
