/*
 * rvo2_ref_shim.cpp -- TEST INFRASTRUCTURE.  A C entry point over the reference's OWN RVO2 sources
 * (gym_collision_avoidance/envs/policies/Python-RVO2/src/{Agent,KdTree,Obstacle,RVOSimulator}.cpp, compiled where
 * they lie under /root/reference by oracle/Makefile into oracle/_ref/librvo2_ref.so; no reference source is copied).
 * It drives RVOSimulator through exactly the calls RVOPolicy.find_next_action makes
 * (envs/policies/RVOPolicy.py:24-28 constructor, :47-48 addAgent, :76-85 setters, :88 doStep, :91 getAgentPosition),
 * with the same double -> float conversions the Cython binding performs (Python-RVO2/src/rvo2.pyx:212-228).
 * Used only to pin oracle/casim_oracle_rvo.c and to generate tests/golden/rvo.npz.
 */
#include <stddef.h>
#include <stdint.h>
#include <vector>

#include "RVO.h"

/* With obstacles: RVOPolicy.find_next_action adds them before the simulator's one processObstacles call of the first
 * step (RVOPolicy.py:56-57, :61-63, :44); Cython converts each vertex to Vector2(float, float) (rvo2.pyx:119-120). */
extern "C" int rvo2_ref_step_obst(int n, const double* pos, const double* vel, const double* pref_vel, const double* radius,
                                  const double* max_speed, int ego, double collab, double dt, double neighbor_dist,
                                  int max_neighbors, double time_horizon, int n_polys, const int32_t* poly_sizes,
                                  const double* poly_verts, double* new_pos /* [2] */) {
  RVO::RVOSimulator sim((float)dt, (float)neighbor_dist, (size_t)max_neighbors, (float)time_horizon, (float)time_horizon,
                        0.0f, 0.0f);
  int off = 0;
  for (int p = 0; p < n_polys; p++) {
    std::vector<RVO::Vector2> v;
    for (int k = 0; k < poly_sizes[p]; k++) v.push_back(RVO::Vector2((float)poly_verts[2 * (off + k)], (float)poly_verts[2 * (off + k) + 1]));
    sim.addObstacle(v);
    off += poly_sizes[p];
  }
  sim.processObstacles();
  for (int a = 0; a < n; a++) sim.addAgent(RVO::Vector2(0.0f, 0.0f));
  for (int a = 0; a < n; a++) {
    sim.setAgentMaxSpeed(a, (float)max_speed[a]);
    sim.setAgentRadius(a, (float)radius[a]);
    sim.setAgentPosition(a, RVO::Vector2((float)pos[2 * a], (float)pos[2 * a + 1]));
    sim.setAgentVelocity(a, RVO::Vector2((float)vel[2 * a], (float)vel[2 * a + 1]));
    sim.setAgentPrefVelocity(a, RVO::Vector2((float)pref_vel[2 * a], (float)pref_vel[2 * a + 1]));
  }
  sim.setAgentCollabCoeff(ego, (float)collab);
  sim.doStep();
  RVO::Vector2 p = sim.getAgentPosition(ego);
  new_pos[0] = p.x();
  new_pos[1] = p.y();
  return 0;
}

extern "C" int rvo2_ref_step(int n, const double* pos, const double* vel, const double* pref_vel, const double* radius,
                             const double* max_speed, int ego, double collab, double dt, double neighbor_dist,
                             int max_neighbors, double time_horizon, double* new_pos /* [2] */) {
  return rvo2_ref_step_obst(n, pos, vel, pref_vel, radius, max_speed, ego, collab, dt, neighbor_dist, max_neighbors,
                            time_horizon, 0, nullptr, nullptr, new_pos);
}
