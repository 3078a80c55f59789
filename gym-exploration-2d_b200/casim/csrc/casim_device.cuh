/*
 * casim_device.cuh -- device helpers shared by the kernels: angle wrap, world->cell transform,
 * bit-packed map access, warp reductions.  Reference citations are relative to
 * gym_collision_avoidance/envs/ of mlodel/gym-exploration-2d.
 */
#ifndef CASIM_DEVICE_CUH
#define CASIM_DEVICE_CUH

#include "casim_internal.h"

#define CASIM_PI 3.14159265358979323846
#define CASIM_2PI (2.0 * CASIM_PI)
#define FULL_MASK 0xffffffffu

/* util.py:27-32 wrap(): repeated +-2pi into [-pi, pi).  The loop is what the reference does; the
 * closed form only guards against absurd caller-supplied angles. */
__device__ __forceinline__ double wrap_pi(double a) {
  if (!(fabs(a) < 400.0)) {
    if (!isfinite(a)) return a;
    a = a - CASIM_2PI * floor((a + CASIM_PI) / CASIM_2PI);
  }
  while (a >= CASIM_PI) a -= CASIM_2PI;
  while (a < -CASIM_PI) a += CASIM_2PI;
  return a;
}

__device__ __forceinline__ double sgn_d(double x) { return (double)((x > 0) - (x < 0)); }

/* Map.py:40-47 world_coordinates_to_map_indices: row = floor(ox - y/cell), col = floor(oy + x/cell),
 * true fp64 divisions.  Returns in-map flag. */
__device__ __forceinline__ bool world_to_map_div(const DevCfg& c, double x, double y, int& row, int& col) {
  double fr = floor(c.map_ox - y / c.cell), fc = floor(c.map_oy + x / c.cell);
  if (!(fr > -1e9 && fr < 1e9)) fr = -1e9;
  if (!(fc > -1e9 && fc < 1e9)) fc = -1e9;
  row = (int)fr;
  col = (int)fc;
  return row >= 0 && col >= 0 && row < CASIM_MAP_DIM && col < CASIM_MAP_DIM;
}

/* Same transform with the division replaced by a multiplication by 1/cell (exact 10.0 for the
 * reference's 0.1 m cells).  x*10 and x/0.1 differ by at most one ulp, so the floor differs only for
 * points within 1e-14 cells of a cell edge; used for the laser lattice (256 points per agent). */
__device__ __forceinline__ bool world_to_map_mul(const DevCfg& c, double x, double y, int& row, int& col) {
  double fr = floor(c.map_ox - y * c.inv_cell), fc = floor(c.map_oy + x * c.inv_cell);
  if (!(fr > -1e9 && fr < 1e9)) fr = -1e9;
  if (!(fc > -1e9 && fc < 1e9)) fc = -1e9;
  row = (int)fr;
  col = (int)fc;
  return row >= 0 && col >= 0 && row < CASIM_MAP_DIM && col < CASIM_MAP_DIM;
}

/* bit-packed occupancy plane: word (row*10 + col/32), bit col%32 */
__device__ __forceinline__ int map_bit(const uint32_t* __restrict__ plane, int row, int col) {
  return (__ldg(plane + row * CASIM_MAP_WORDS + (col >> 5)) >> (col & 31)) & 1u;
}

/* Largest k >= 0 with k*k < rem (rem > 0), exact for the small values used (disc radius <= ~30 cells). */
__device__ __forceinline__ int chord_halfwidth(double rem) {
  int k = (int)sqrt(rem);
  while ((double)k * (double)k >= rem) k--;
  while ((double)(k + 1) * (double)(k + 1) < rem) k++;
  return k;
}

/* collision_avoidance_env.py:656-666 + Map.py:67-71: does the disc of cells
 * (col-c0)^2 + (row-r0)^2 < (radius/cell)^2 around the agent's (in-map) cell touch an occupied cell? */
__device__ __forceinline__ bool disc_hits_map(const DevCfg& c, const uint32_t* __restrict__ plane, int r0,
                                              int c0, double radius) {
  double rc = radius / c.cell;
  double rc2 = rc * rc;
  int span = (int)ceil(rc) + 1;
  if (span > CASIM_MAP_DIM) span = CASIM_MAP_DIM;
  int rlo = max(r0 - span, 0), rhi = min(r0 + span, CASIM_MAP_DIM - 1);
  for (int row = rlo; row <= rhi; row++) {
    double dy = (double)(row - r0);
    double rem = rc2 - dy * dy;
    if (!(rem > 0.0)) continue;
    int k = chord_halfwidth(rem);
    int lo = max(c0 - k, 0), hi = min(c0 + k, CASIM_MAP_DIM - 1);
    if (lo > hi) continue;
    const uint32_t* wrow = plane + row * CASIM_MAP_WORDS;
    for (int w = lo >> 5; w <= (hi >> 5); w++) {
      int b0 = max(lo - (w << 5), 0), b1 = min(hi - (w << 5), 31);
      uint32_t m = (b1 - b0 == 31) ? 0xffffffffu : (((1u << (b1 - b0 + 1)) - 1u) << b0);
      if (__ldg(wrow + w) & m) return true;
    }
  }
  return false;
}

__device__ __forceinline__ bool in_disc(int row, int col, int r0, int c0, double rc2) {
  double dx = (double)(col - c0), dy = (double)(row - r0);
  return dx * dx + dy * dy < rc2;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
  return v;
}

#endif
