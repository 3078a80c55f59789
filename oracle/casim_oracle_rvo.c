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
 * Scope: agent-agent and static-obstacle ORCA lines.  Static obstacles reach the reference simulator only through the
 * first find_next_action call (RVOPolicy.py:56-57 adds them on every call but processObstacles runs once, :44): the
 * obstacle set is an episode constant here (RVOSimulator::addObstacle, KdTree::buildObstacleTree with edge splitting,
 * queryObstacleTreeRecursive, Agent.cpp:66-284 are all restated below).
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
#define MAXOBST 1024 /* obstacle vertices incl. the ones edge splitting creates */

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
  static __thread line_t proj[MAXOBST + 2 * MAXN];
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

/* ---------------------------------------------------------------------------------------------------------
 * Static obstacles: RVOSimulator::addObstacle (RVOSimulator.cpp:124-161), KdTree::buildObstacleTree(Recursive)
 * (KdTree.cpp:113-245, including the edge splitting), KdTree::queryObstacleTreeRecursive (:300-329),
 * Agent::insertObstacleNeighbor (Agent.cpp:475-493) and the obstacle ORCA lines (Agent.cpp:66-284).
 * --------------------------------------------------------------------------------------------------------- */
typedef struct {
  int n;                       /* vertices ("obstacles" in RVO2's vocabulary), including the ones splitting created */
  v2 point[MAXOBST], unitDir[MAXOBST];
  int next[MAXOBST], prev[MAXOBST], convex[MAXOBST];
  int n_nodes, root;
  int node_obst[MAXOBST], node_left[MAXOBST], node_right[MAXOBST];
} obst_set;

static float leftOf(v2 a, v2 b, v2 c) { return det(sub(a, c), sub(b, a)); } /* Definitions.h:93-96 */

static int obst_add_polygon(obst_set* O, const v2* vert, int nv) { /* RVOSimulator.cpp:124-161 */
  if (nv < 2 || O->n + nv > MAXOBST) return -1;
  const int first = O->n;
  for (int i = 0; i < nv; i++) {
    const int k = O->n;
    O->point[k] = vert[i];
    if (i != 0) {
      O->prev[k] = k - 1;
      O->next[k - 1] = k;
    }
    if (i == nv - 1) {
      O->next[k] = first;
      O->prev[first] = k;
    }
    O->unitDir[k] = normalize(sub(vert[i == nv - 1 ? 0 : i + 1], vert[i]));
    O->convex[k] = nv == 2 ? 1 : (leftOf(vert[i == 0 ? nv - 1 : i - 1], vert[i], vert[i == nv - 1 ? 0 : i + 1]) >= 0.0f);
    O->n++;
  }
  return first;
}

static int pair_lt(size_t a1, size_t a2, size_t b1, size_t b2) { return a1 < b1 || (a1 == b1 && a2 < b2); }

static int obst_build(obst_set* O, const int* list, int count) { /* KdTree.cpp:127-245 */
  if (count == 0) return -1;
  const int node = O->n_nodes++;
  size_t optimalSplit = 0, minLeft = (size_t)count, minRight = (size_t)count;
  for (int i = 0; i < count; i++) {
    size_t leftSize = 0, rightSize = 0;
    const int i1 = list[i], i2 = O->next[i1];
    for (int j = 0; j < count; j++) {
      if (i == j) continue;
      const int j1 = list[j], j2 = O->next[j1];
      const float j1LeftOfI = leftOf(O->point[i1], O->point[i2], O->point[j1]);
      const float j2LeftOfI = leftOf(O->point[i1], O->point[i2], O->point[j2]);
      if (j1LeftOfI >= -RVO_EPS && j2LeftOfI >= -RVO_EPS) ++leftSize;
      else if (j1LeftOfI <= RVO_EPS && j2LeftOfI <= RVO_EPS) ++rightSize;
      else { ++leftSize; ++rightSize; }
      size_t a1 = leftSize > rightSize ? leftSize : rightSize, a2 = leftSize > rightSize ? rightSize : leftSize;
      size_t b1 = minLeft > minRight ? minLeft : minRight, b2 = minLeft > minRight ? minRight : minLeft;
      if (!pair_lt(a1, a2, b1, b2)) break; /* >= */
    }
    size_t a1 = leftSize > rightSize ? leftSize : rightSize, a2 = leftSize > rightSize ? rightSize : leftSize;
    size_t b1 = minLeft > minRight ? minLeft : minRight, b2 = minLeft > minRight ? minRight : minLeft;
    if (pair_lt(a1, a2, b1, b2)) {
      minLeft = leftSize;
      minRight = rightSize;
      optimalSplit = (size_t)i;
    }
  }
  int* leftL = (int*)malloc(sizeof(int) * (minLeft + 1));
  int* rightL = (int*)malloc(sizeof(int) * (minRight + 1));
  size_t lc = 0, rc = 0;
  const int i = (int)optimalSplit;
  const int i1 = list[i], i2 = O->next[i1];
  for (int j = 0; j < count; j++) {
    if (i == j) continue;
    const int j1 = list[j], j2 = O->next[j1];
    const float j1LeftOfI = leftOf(O->point[i1], O->point[i2], O->point[j1]);
    const float j2LeftOfI = leftOf(O->point[i1], O->point[i2], O->point[j2]);
    if (j1LeftOfI >= -RVO_EPS && j2LeftOfI >= -RVO_EPS) leftL[lc++] = j1;
    else if (j1LeftOfI <= RVO_EPS && j2LeftOfI <= RVO_EPS) rightL[rc++] = j1;
    else { /* split edge j at the line through edge i */
      const float t = det(sub(O->point[i2], O->point[i1]), sub(O->point[j1], O->point[i1])) /
                      det(sub(O->point[i2], O->point[i1]), sub(O->point[j1], O->point[j2]));
      const v2 splitpoint = add(O->point[j1], scalel(t, sub(O->point[j2], O->point[j1])));
      const int nw = O->n++;
      O->point[nw] = splitpoint;
      O->prev[nw] = j1;
      O->next[nw] = j2;
      O->convex[nw] = 1;
      O->unitDir[nw] = O->unitDir[j1];
      O->next[j1] = nw;
      O->prev[j2] = nw;
      if (j1LeftOfI > 0.0f) { leftL[lc++] = j1; rightL[rc++] = nw; }
      else { rightL[rc++] = j1; leftL[lc++] = nw; }
    }
  }
  O->node_obst[node] = i1;
  const int l = obst_build(O, leftL, (int)lc);
  const int r = obst_build(O, rightL, (int)rc);
  O->node_left[node] = l;
  O->node_right[node] = r;
  free(leftL);
  free(rightL);
  return node;
}

static void obst_process(obst_set* O) { /* KdTree.cpp:113-124 */
  int* all = (int*)malloc(sizeof(int) * (O->n + 1));
  const int n0 = O->n;
  for (int k = 0; k < n0; k++) all[k] = k;
  O->n_nodes = 0;
  O->root = obst_build(O, all, n0);
  free(all);
}

static float distSqPointLineSegment(v2 a, v2 b, v2 c) { /* Definitions.h:68-82 */
  const float r = dot(sub(c, a), sub(b, a)) / absSq(sub(b, a));
  if (r < 0.0f) return absSq(sub(c, a));
  if (r > 1.0f) return absSq(sub(c, b));
  return absSq(sub(c, add(a, scalel(r, sub(b, a)))));
}

typedef struct { float d[MAXOBST]; int o[MAXOBST]; int n; } obst_nb;

static void obst_query(const obst_set* O, v2 position, float rangeSq, int node, obst_nb* nb) { /* KdTree.cpp:300-329 */
  if (node < 0) return;
  const int o1 = O->node_obst[node], o2 = O->next[o1];
  const float agentLeftOfLine = leftOf(O->point[o1], O->point[o2], position);
  obst_query(O, position, rangeSq, agentLeftOfLine >= 0.0f ? O->node_left[node] : O->node_right[node], nb);
  const float distSqLine = sqr(agentLeftOfLine) / absSq(sub(O->point[o2], O->point[o1]));
  if (distSqLine < rangeSq) {
    if (agentLeftOfLine < 0.0f) { /* Agent.cpp:475-493 insertObstacleNeighbor */
      const float distSq = distSqPointLineSegment(O->point[o1], O->point[O->next[o1]], position);
      if (distSq < rangeSq) {
        int i = nb->n++;
        while (i != 0 && distSq < nb->d[i - 1]) {
          nb->d[i] = nb->d[i - 1];
          nb->o[i] = nb->o[i - 1];
          --i;
        }
        nb->d[i] = distSq;
        nb->o[i] = o1;
      }
    }
    obst_query(O, position, rangeSq, agentLeftOfLine >= 0.0f ? O->node_right[node] : O->node_left[node], nb);
  }
}

/* Agent.cpp:66-284: ORCA lines of the obstacle neighbours (sorted by distance); returns the number of lines */
static int obstacle_lines(const obst_set* O, const obst_nb* nb, v2 position_, v2 velocity_, float radius_,
                          float timeHorizonObst_, line_t* lines) {
  int nl = 0;
  const float invTimeHorizonObst = 1.0f / timeHorizonObst_;
  for (int i = 0; i < nb->n; i++) {
    int obstacle1 = nb->o[i];
    int obstacle2 = O->next[obstacle1];
    const v2 relativePosition1 = sub(O->point[obstacle1], position_);
    const v2 relativePosition2 = sub(O->point[obstacle2], position_);
    int alreadyCovered = 0;
    for (int j = 0; j < nl; j++) {
      if (det(sub(scalel(invTimeHorizonObst, relativePosition1), lines[j].point), lines[j].direction) - invTimeHorizonObst * radius_ >= -RVO_EPS &&
          det(sub(scalel(invTimeHorizonObst, relativePosition2), lines[j].point), lines[j].direction) - invTimeHorizonObst * radius_ >= -RVO_EPS) {
        alreadyCovered = 1;
        break;
      }
    }
    if (alreadyCovered) continue;
    const float distSq1 = absSq(relativePosition1);
    const float distSq2 = absSq(relativePosition2);
    const float radiusSq = sqr(radius_);
    const v2 obstacleVector = sub(O->point[obstacle2], O->point[obstacle1]);
    const float s = dot(neg(relativePosition1), obstacleVector) / absSq(obstacleVector);
    const float distSqLine = absSq(sub(neg(relativePosition1), scalel(s, obstacleVector)));
    line_t line;
    if (s < 0.0f && distSq1 <= radiusSq) {
      if (O->convex[obstacle1]) {
        line.point = V(0.0f, 0.0f);
        line.direction = normalize(V(-relativePosition1.y, relativePosition1.x));
        lines[nl++] = line;
      }
      continue;
    } else if (s > 1.0f && distSq2 <= radiusSq) {
      if (O->convex[obstacle2] && det(relativePosition2, O->unitDir[obstacle2]) >= 0.0f) {
        line.point = V(0.0f, 0.0f);
        line.direction = normalize(V(-relativePosition2.y, relativePosition2.x));
        lines[nl++] = line;
      }
      continue;
    } else if (s >= 0.0f && s < 1.0f && distSqLine <= radiusSq) {
      line.point = V(0.0f, 0.0f);
      line.direction = neg(O->unitDir[obstacle1]);
      lines[nl++] = line;
      continue;
    }
    v2 leftLegDirection, rightLegDirection;
    if (s < 0.0f && distSqLine <= radiusSq) {
      if (!O->convex[obstacle1]) continue;
      obstacle2 = obstacle1;
      const float leg1 = sqrtf(distSq1 - radiusSq);
      leftLegDirection = divs(V(relativePosition1.x * leg1 - relativePosition1.y * radius_, relativePosition1.x * radius_ + relativePosition1.y * leg1), distSq1);
      rightLegDirection = divs(V(relativePosition1.x * leg1 + relativePosition1.y * radius_, -relativePosition1.x * radius_ + relativePosition1.y * leg1), distSq1);
    } else if (s > 1.0f && distSqLine <= radiusSq) {
      if (!O->convex[obstacle2]) continue;
      obstacle1 = obstacle2;
      const float leg2 = sqrtf(distSq2 - radiusSq);
      leftLegDirection = divs(V(relativePosition2.x * leg2 - relativePosition2.y * radius_, relativePosition2.x * radius_ + relativePosition2.y * leg2), distSq2);
      rightLegDirection = divs(V(relativePosition2.x * leg2 + relativePosition2.y * radius_, -relativePosition2.x * radius_ + relativePosition2.y * leg2), distSq2);
    } else {
      if (O->convex[obstacle1]) {
        const float leg1 = sqrtf(distSq1 - radiusSq);
        leftLegDirection = divs(V(relativePosition1.x * leg1 - relativePosition1.y * radius_, relativePosition1.x * radius_ + relativePosition1.y * leg1), distSq1);
      } else {
        leftLegDirection = neg(O->unitDir[obstacle1]);
      }
      if (O->convex[obstacle2]) {
        const float leg2 = sqrtf(distSq2 - radiusSq);
        rightLegDirection = divs(V(relativePosition2.x * leg2 + relativePosition2.y * radius_, -relativePosition2.x * radius_ + relativePosition2.y * leg2), distSq2);
      } else {
        rightLegDirection = O->unitDir[obstacle1];
      }
    }
    const int leftNeighbor = O->prev[obstacle1];
    int isLeftLegForeign = 0, isRightLegForeign = 0;
    if (O->convex[obstacle1] && det(leftLegDirection, neg(O->unitDir[leftNeighbor])) >= 0.0f) {
      leftLegDirection = neg(O->unitDir[leftNeighbor]);
      isLeftLegForeign = 1;
    }
    if (O->convex[obstacle2] && det(rightLegDirection, O->unitDir[obstacle2]) <= 0.0f) {
      rightLegDirection = O->unitDir[obstacle2];
      isRightLegForeign = 1;
    }
    const v2 leftCutoff = scalel(invTimeHorizonObst, sub(O->point[obstacle1], position_));
    const v2 rightCutoff = scalel(invTimeHorizonObst, sub(O->point[obstacle2], position_));
    const v2 cutoffVec = sub(rightCutoff, leftCutoff);
    const float t = obstacle1 == obstacle2 ? 0.5f : dot(sub(velocity_, leftCutoff), cutoffVec) / absSq(cutoffVec);
    const float tLeft = dot(sub(velocity_, leftCutoff), leftLegDirection);
    const float tRight = dot(sub(velocity_, rightCutoff), rightLegDirection);
    if ((t < 0.0f && tLeft < 0.0f) || (obstacle1 == obstacle2 && tLeft < 0.0f && tRight < 0.0f)) {
      const v2 unitW = normalize(sub(velocity_, leftCutoff));
      line.direction = V(unitW.y, -unitW.x);
      line.point = add(leftCutoff, scalel(radius_ * invTimeHorizonObst, unitW));
      lines[nl++] = line;
      continue;
    } else if (t > 1.0f && tRight < 0.0f) {
      const v2 unitW = normalize(sub(velocity_, rightCutoff));
      line.direction = V(unitW.y, -unitW.x);
      line.point = add(rightCutoff, scalel(radius_ * invTimeHorizonObst, unitW));
      lines[nl++] = line;
      continue;
    }
    const float distSqCutoff = (t < 0.0f || t > 1.0f || obstacle1 == obstacle2) ? INFINITY : absSq(sub(velocity_, add(leftCutoff, scalel(t, cutoffVec))));
    const float distSqLeft = tLeft < 0.0f ? INFINITY : absSq(sub(velocity_, add(leftCutoff, scalel(tLeft, leftLegDirection))));
    const float distSqRight = tRight < 0.0f ? INFINITY : absSq(sub(velocity_, add(rightCutoff, scalel(tRight, rightLegDirection))));
    if (distSqCutoff <= distSqLeft && distSqCutoff <= distSqRight) {
      line.direction = neg(O->unitDir[obstacle1]);
      line.point = add(leftCutoff, scalel(radius_ * invTimeHorizonObst, V(-line.direction.y, line.direction.x)));
      lines[nl++] = line;
      continue;
    } else if (distSqLeft <= distSqRight) {
      if (isLeftLegForeign) continue;
      line.direction = leftLegDirection;
      line.point = add(leftCutoff, scalel(radius_ * invTimeHorizonObst, V(-line.direction.y, line.direction.x)));
      lines[nl++] = line;
      continue;
    } else {
      if (isRightLegForeign) continue;
      line.direction = neg(rightLegDirection);
      line.point = add(rightCutoff, scalel(radius_ * invTimeHorizonObst, V(-line.direction.y, line.direction.x)));
      lines[nl++] = line;
      continue;
    }
  }
  return nl;
}

/* One ego agent of RVOSimulator::doStep (RVOSimulator.cpp:163-183): computeNeighbors + computeNewVelocity + update.
 * pos/vel/pref/radius/max_speed are the values AFTER the binding's float conversion. */
static void rvo_ego_step(int n, const v2* pos, const v2* vel, const v2* pref, const float* radius, const float* max_speed,
                         int ego, float collab, float timeStep, float neighborDist, int maxNeighbors, float timeHorizon,
                         const obst_set* O, float timeHorizonObst, v2* new_pos) {
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
  static __thread line_t lines[MAXOBST + MAXN + 2];
  int nl = 0;
  v2 prefVel = pref[ego];
  const v2 position = pos[ego], velocity = vel[ego];
  const float radius_ = radius[ego];
  if (O && O->root >= 0) { /* Agent.cpp:45-47 + 66-284 */
    static __thread obst_nb nb;
    nb.n = 0;
    obst_query(O, position, sqr(timeHorizonObst * max_speed[ego] + radius_), O->root, &nb);
    nl = obstacle_lines(O, &nb, position, velocity, radius_, timeHorizonObst, lines);
  }
  const int numObstLines = nl;
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
  if (lineFail < nl) lp3(lines, nl, numObstLines, lineFail, max_speed[ego], &newVel);
  /* Agent.cpp:516-521 update */
  *new_pos = add(position, scale(newVel, timeStep));
}

static obst_set* obst_from_polys(int n_polys, const int32_t* poly_sizes, const double* verts) {
  if (n_polys <= 0) return NULL;
  obst_set* O = (obst_set*)calloc(1, sizeof(obst_set));
  O->root = -1;
  int off = 0;
  for (int p = 0; p < n_polys; p++) {
    v2 tmp[64];
    const int nv = poly_sizes[p];
    if (nv > 64) { free(O); return NULL; }
    for (int k = 0; k < nv; k++) tmp[k] = V((float)verts[2 * (off + k)], (float)verts[2 * (off + k) + 1]); /* rvo2.pyx:119-120 */
    if (obst_add_polygon(O, tmp, nv) < 0) { free(O); return NULL; }
    off += nv;
  }
  obst_process(O);
  return O;
}

/* Same signatures as oracle/rvo2_ref_shim.cpp, so that the two can be compared call for call. */
int casim_oracle_rvo_ref_step_obst(int n, const double* pos, const double* vel, const double* pref_vel, const double* radius,
                                   const double* max_speed, int ego, double collab, double dt, double neighbor_dist,
                                   int max_neighbors, double time_horizon, int n_polys, const int32_t* poly_sizes,
                                   const double* poly_verts, double* new_pos) {
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
  obst_set* O = obst_from_polys(n_polys, poly_sizes, poly_verts);
  if (n_polys > 0 && !O) return 2;
  v2 np_;
  rvo_ego_step(n, p, v, pv, r, ms, ego, (float)collab, (float)dt, (float)neighbor_dist, max_neighbors, (float)time_horizon,
               O, (float)time_horizon, &np_);
  free(O);
  new_pos[0] = np_.x;
  new_pos[1] = np_.y;
  return 0;
}

int casim_oracle_rvo_ref_step(int n, const double* pos, const double* vel, const double* pref_vel, const double* radius,
                              const double* max_speed, int ego, double collab, double dt, double neighbor_dist,
                              int max_neighbors, double time_horizon, double* new_pos) {
  return casim_oracle_rvo_ref_step_obst(n, pos, vel, pref_vel, radius, max_speed, ego, collab, dt, neighbor_dist,
                                        max_neighbors, time_horizon, 0, NULL, NULL, new_pos);
}

/* The processed obstacle set (vertices incl. split points, links, tree) for cross-checking the product's own
 * preprocessing: returns the number of vertices; arrays must hold MAXOBST entries. */
int casim_oracle_rvo_obstacles(int n_polys, const int32_t* poly_sizes, const double* poly_verts, float* points /*[n][2]*/,
                               int32_t* next, int32_t* prev, int32_t* convex, int32_t* n_nodes, int32_t* node_obst,
                               int32_t* node_left, int32_t* node_right) {
  obst_set* O = obst_from_polys(n_polys, poly_sizes, poly_verts);
  if (!O) return -1;
  for (int k = 0; k < O->n; k++) {
    points[2 * k] = O->point[k].x; points[2 * k + 1] = O->point[k].y;
    next[k] = O->next[k]; prev[k] = O->prev[k]; convex[k] = O->convex[k];
  }
  *n_nodes = O->n_nodes;
  for (int k = 0; k < O->n_nodes; k++) { node_obst[k] = O->node_obst[k]; node_left[k] = O->node_left[k]; node_right[k] = O->node_right[k]; }
  int n = O->n;
  free(O);
  return n;
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
  const double* state; const double* coop; double* actions; int e0, e1; const obst_set* O;
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
                   (float)prm->neighbor_dist, prm->max_neighbors, (float)prm->time_horizon, J->O,
                   (float)prm->time_horizon, &np_);
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

/* obstacles: the polygons RVOPolicy hands to sim.addObstacle on its first call (RVOPolicy.py:56-57, :44), shared by
 * all environments; n_polys = 0 for none. */
int casim_oracle_rvo_actions_obst(const casim_config* cfg, const casim_rvo_params* prm, int E, int A, const int32_t* n_agents,
                                  const double* state, const double* coop, double* actions, int n_threads, int n_polys,
                                  const int32_t* poly_sizes, const double* poly_verts) {
  if (A > MAXN) return 1;
  obst_set* O = obst_from_polys(n_polys, poly_sizes, poly_verts);
  if (n_polys > 0 && !O) return 2;
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  rvo_job jobs[64];
  pthread_t th[64];
  for (int t = 0; t < n_threads; t++) {
    rvo_job* j = &jobs[t];
    j->cfg = cfg; j->prm = prm; j->A = A; j->n_agents = n_agents; j->state = state; j->coop = coop; j->actions = actions;
    j->O = O;
    j->e0 = (int)((long long)E * t / n_threads);
    j->e1 = (int)((long long)E * (t + 1) / n_threads);
    if (n_threads == 1) rvo_job_main(j); else pthread_create(&th[t], NULL, rvo_job_main, j);
  }
  if (n_threads > 1) for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
  free(O);
  return 0;
}

int casim_oracle_rvo_actions_mt(const casim_config* cfg, const casim_rvo_params* prm, int E, int A, const int32_t* n_agents,
                                const double* state, const double* coop, double* actions, int n_threads) {
  return casim_oracle_rvo_actions_obst(cfg, prm, E, A, n_agents, state, coop, actions, n_threads, 0, NULL, NULL);
}

int casim_oracle_rvo_actions(const casim_config* cfg, const casim_rvo_params* prm, int E, int A, const int32_t* n_agents,
                             const double* state, const double* coop, double* actions) {
  return casim_oracle_rvo_actions_mt(cfg, prm, E, A, n_agents, state, coop, actions, 1);
}
