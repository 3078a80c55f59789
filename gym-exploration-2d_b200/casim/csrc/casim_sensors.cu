/*
 * casim_sensors.cu -- LaserScanSensor and OccupancyGridSensor kernels.
 *
 *   k_laser : sensors/LaserScanSensor.py:27-58 (+ Map.py:49-59 lattice -> cell indices, Map.py:73-79 ego disc).
 *             One warp per agent ("ray fan"): lanes split the beams x range-samples lattice, each lane builds
 *             a hit bit-mask for its part of a beam, masks are OR-combined with shuffles, and the
 *             "last sample whose running hit count is 1" rule is evaluated with bit tricks.
 *   k_grid  : sensors/OccupancyGridSensor.py:70-98,131-143 + Map.py:81-105 + cv2.getRotationMatrix2D /
 *             cv2.warpAffine (bilinear, zero border) restated in OpenCV's fixed point (AB_BITS 10, INTER_BITS 5).
 *             Because the result is cast to bool, a destination cell is the OR of the source cells with a
 *             non-zero bilinear weight; the four possible footprints are precomputed as bit planes
 *             (casim_map.cu), so a cell is ONE bit lookup.
 *
 * Both read the fp64 post-move pose scratch written by the step kernel (pose64), so that their cell
 * indices are computed from the same float64 numbers the reference would use inside one step.
 */
#include <cuda.h>

#include "casim_device.cuh"

namespace {

struct LaserArgs {
  DevCfg c;
  const float4* state;
  const double* pose64;
  const int32_t* n_agents;
  const uint32_t* map_planes;
  const uint32_t* grid_maps; /* [n_maps][CASIM_GMAP_WORDS]: plane 0 = the map padded by CASIM_GPAD free cells */
  float* out;
  int E, A, n_maps;
  int lg2P; /* lanes per beam = 1 << lg2P = 32 / next_pow2(num_beams), computed by the launcher */
};

#ifndef LASER_WPC
#define LASER_WPC 8 /* warps (= agents) per CTA; needs 32 * LASER_WPC >= CASIM_MAX_BEAMS + CASIM_MAX_RANGES + 1 threads for the table fill */
#endif
/* PADDED: one shared map -- samples read the padded copy (no bounds test); per-environment maps keep the 12 KB unpadded
 * planes (the padded copies are 2.6x the bytes per environment: 35.3 vs 32.2 us at 4096 maps) */
template <bool PADDED>
__global__ void __launch_bounds__(32 * LASER_WPC) k_laser(const LaserArgs p) {
  /* Per-beam / per-sample tables: indexed by lane-dependent b and k, so they live in shared memory (a divergent
   * index into the kernel-parameter constant bank is serialised, one address per issue).  s_out[k] is the value
   * reported for "range = sample k" (k == nr: no hit, max range), divided once on the host (DevCfg.laser_out). */
  __shared__ double s_cos[CASIM_MAX_BEAMS], s_sin[CASIM_MAX_BEAMS], s_r[CASIM_MAX_RANGES];
  __shared__ float s_out[CASIM_MAX_RANGES + 1];
  const DevCfg& c = p.c;
  const int nb = c.nb, nr = c.nr;
  if (threadIdx.x < CASIM_MAX_BEAMS) {
    s_cos[threadIdx.x] = c.laser_cos[threadIdx.x];
    s_sin[threadIdx.x] = c.laser_sin[threadIdx.x];
  } else if (threadIdx.x < CASIM_MAX_BEAMS + CASIM_MAX_RANGES) {
    const int k = threadIdx.x - CASIM_MAX_BEAMS;
    s_r[k] = c.laser_r[k];
    s_out[k] = c.laser_out[k];
  } else if (threadIdx.x == CASIM_MAX_BEAMS + CASIM_MAX_RANGES) {
    s_out[CASIM_MAX_RANGES] = c.laser_out[CASIM_MAX_RANGES];
  }
  casim_pdl_enter();
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const unsigned g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); /* E*A < 2^31 (checked by the launcher) */
  const unsigned N = (unsigned)p.E * (unsigned)p.A;
  if (g >= N) return;
  const int e = (int)(g / (unsigned)p.A), ai = (int)(g - (unsigned)e * (unsigned)p.A);
  const int n = p.n_agents ? min(p.n_agents[e], p.A) : p.A;
  float* out = p.out + (size_t)g * nb;
  const int type = __float_as_int(p.state[5 * (size_t)N + g].w);
  const bool on = ai < n && (((type >> 16) & 0xff) & CASIM_SENSOR_LASERSCAN);
  if (!on) {
    for (int b = lane; b < nb; b += 32) out[b] = 0.f;
    return;
  }
  const double2 q0 = reinterpret_cast<const double2*>(p.pose64 + 4 * (size_t)g)[0];
  const double2 q1 = reinterpret_cast<const double2*>(p.pose64 + 4 * (size_t)g)[1];
  const double px = q0.x, py = q0.y, ch = q1.x, sh = q1.y;
  const double radius = (double)p.state[2 * (size_t)N + g].z;
  const uint32_t* pl =
      p.n_maps > 0 ? p.map_planes + (size_t)(p.n_maps == 1 ? 0 : e) * CASIM_NUM_PLANES * CASIM_PLANE_WORDS : nullptr;
  int r0, c0;
  const bool ego_in = world_to_map_mul(c, px, py, r0, c0);
  const double rc = radius * c.inv_cell, rc2 = rc * rc;
  /* integer form of the ego-disc test d^2 < rc2 (d^2 integer): d^2 <= ceil(rc2) - 1; no disc if out of map */
  const int disc_thr = ego_in ? (int)fmin(ceil(rc2), 2.0e9) - 1 : -1;
  /* lanes per beam: P = 32 / next_pow2(nb), nb <= 32 */
  const int P = 1 << p.lg2P;
  const int b = lane >> p.lg2P, part = lane & (P - 1);
  /* kd = samples that may fall into the ego disc: r_k / cell < rc + 1.5 (a cell index is within one cell of the point it
   * holds); a prefix when the r_k increase (np.arange, LaserScanSensor.py:18-20), every sample otherwise */
  const int kd = !c.laser_sorted ? nr
                                 : __popc(__ballot_sync(FULL_MASK, lane < nr && disc_thr >= 0 &&
                                                                       s_r[min(lane, CASIM_MAX_RANGES - 1)] * c.inv_cell < rc + 1.5));
  unsigned mask = 0;
  if (b < nb && pl) {
    /* beam direction cos/sin(angle_b + heading) by angle addition from the step kernel's cos/sin(heading);
     * lattice point k in CELL units: col = floor(oy + (px + r_k cs)/cell), row = floor(ox - (py + r_k sn)/cell).
     * The samples read plane 0 of the PADDED copy of the map (casim_map.cu k_grid_pad: 96 free cells on every side), so an
     * out-of-map sample needs no bounds test: its (unsigned-)clamped address is a free padding cell. */
    const uint32_t* gp = p.grid_maps + (size_t)(p.n_maps == 1 ? 0 : e) * CASIM_GMAP_WORDS;
    const double cs = s_cos[b] * ch - s_sin[b] * sh, sn = s_sin[b] * ch + s_cos[b] * sh;
    const double u0 = c.map_oy + px * c.inv_cell, du = cs * c.inv_cell;  /* column coordinate, cells */
    const double v0 = c.map_ox - py * c.inv_cell, dv = -(sn * c.inv_cell); /* row coordinate, cells */
    auto sample = [&](int row, int col) -> uint32_t {
      if (PADDED) {
        const uint32_t pr = min((uint32_t)row + (uint32_t)CASIM_GPAD, (uint32_t)(CASIM_GROWS - 1));
        const uint32_t pc = min((uint32_t)col + (uint32_t)CASIM_GPAD, (uint32_t)(32 * CASIM_GWORDS - 1));
        return (__ldg(gp + pr * CASIM_GWORDS + (pc >> 5)) >> (pc & 31u)) & 1u;
      }
      const bool in = (unsigned)row < (unsigned)CASIM_MAP_DIM && (unsigned)col < (unsigned)CASIM_MAP_DIM;
      return in ? (uint32_t)map_bit(pl, row, col) : 0u;
    };
    int k = part;
    for (; k < kd; k += P) { /* near the agent: with the ego-disc test (Map.py:73-79) */
      const double rk = s_r[k];
      const int col = __double2int_rd(fma(rk, du, u0)), row = __double2int_rd(fma(rk, dv, v0));
      const int dr = row - r0, dc = col - c0;
      mask |= ((dr * dr + dc * dc > disc_thr) ? sample(row, col) : 0u) << k;
    }
#pragma unroll 4
    for (; k < nr; k += P) {
      const double rk = s_r[k];
      mask |= sample(__double2int_rd(fma(rk, dv, v0)), __double2int_rd(fma(rk, du, u0))) << k;
    }
  }
  for (int o = 1; o < P; o <<= 1) mask |= __shfl_xor_sync(FULL_MASK, mask, o);
  if (b < nb && part == 0) {
    int idx = CASIM_MAX_RANGES; /* no hit: max range */
    if (mask) {
      unsigned rest = mask & (mask - 1); /* drop the first hit */
      idx = rest ? (__ffs(rest) - 2) : (nr - 1);
    }
    out[b] = s_out[idx];
  }
}

/* ------------------------------------------------------------------------------------------------ */
struct GridArgs {
  DevCfg c;
  const float4* state;
  const double* pose64;
  const int32_t* n_agents;
  const uint32_t* map_planes;
  uint8_t* out;
  const uint32_t* grid_maps;
  int E, A, n_maps;
};

/* Generic-shape fallback (submap sides that are not multiples of 4, unaligned output).
 * One CTA per agent: warps 0-1 build the per-row / per-column fixed-point tables in fp64, then every thread
 * produces 4-cell output words with integer arithmetic and one bit lookup per cell. */
__global__ void __launch_bounds__(256) k_grid_generic(const GridArgs p) {
  casim_pdl_enter();
  const DevCfg& c = p.c;
  const long long g = blockIdx.x;
  const long long N = (long long)p.E * p.A;
  const int e = (int)(g / p.A), ai = (int)(g - (long long)e * p.A);
  const int n = p.n_agents ? min(p.n_agents[e], p.A) : p.A;
  const int H = c.submap_h, W = c.submap_w; /* out[oy*H + ox], oy < W rows, ox < H cols */
  const int cells = H * W;
  uint8_t* out = p.out + g * cells;
  const int type = __float_as_int(p.state[5 * N + g].w);
  const bool on = ai < n && (((type >> 16) & 0xff) & CASIM_SENSOR_OCCUPANCY_GRID) && p.n_maps > 0;
  __shared__ int sX0[128], sY0[128], sAd[128], sBd[128];
  __shared__ int s_geom[4];
  if (!on) {
    if ((cells & 3) == 0 && (((uintptr_t)out) & 3) == 0)
      for (int i = threadIdx.x; i < (cells >> 2); i += blockDim.x) reinterpret_cast<uint32_t*>(out)[i] = 0u;
    else
      for (int i = threadIdx.x; i < cells; i += blockDim.x) out[i] = 0;
    return;
  }
  const double2 q0 = reinterpret_cast<const double2*>(p.pose64 + 4 * g)[0];
  const double2 q1 = reinterpret_cast<const double2*>(p.pose64 + 4 * g)[1];
  const double px = q0.x, py = q0.y, ch = q1.x, sh = q1.y;
  int row0, col0;
  world_to_map_mul(c, px, py, row0, col0);
  /* Map.getSubmapByIndices (Map.py:81-105) */
  int sx = max(row0 - W / 2, 0), sy = max(col0 - H / 2, 0);
  int ex = sx + W, ey = sy + H;
  if (ex > CASIM_MAP_DIM - 1) {
    ex = CASIM_MAP_DIM - 1;
    sx = ex - W;
  }
  if (ey > CASIM_MAP_DIM - 1) {
    ey = CASIM_MAP_DIM - 1;
    sy = ey - H;
  }
  const int nrows = min(ex - sx, W), ncols = min(ey - sy, H);
  if (threadIdx.x < 128) {
    /* cv2.getRotationMatrix2D(center=(col0,row0), angle=-heading*180/pi, 1) then warpAffine's inversion */
    /* angle = -heading (degrees and back in the reference): alpha = cos(heading), beta = -sin(heading) */
    double alpha = ch, beta = -sh;
    double cx = (double)(float)col0, cy = (double)(float)row0;
    double M0 = alpha, M1 = beta, M2 = (1 - alpha) * cx - beta * cy;
    double M3 = -beta, M4 = alpha, M5 = beta * cx + (1 - alpha) * cy;
    double D = M0 * M4 - M1 * M3;
    D = D != 0 ? 1. / D : 0;
    double A11 = M4 * D, A22 = M0 * D;
    M0 = A11;
    M1 *= -D;
    M3 *= -D;
    M4 = A22;
    double b1 = -M0 * M2 - M1 * M5;
    double b2 = -M3 * M2 - M4 * M5;
    M2 = b1;
    M5 = b2;
    int t = threadIdx.x;
    if (t < 64) {
      if (t < nrows) {
        int y = sx + t;
        sX0[t] = __double2int_rn((M1 * y + M2) * 1024) + 16;
        sY0[t] = __double2int_rn((M4 * y + M5) * 1024) + 16;
      }
    } else {
      t -= 64;
      if (t < ncols) {
        int x = sy + t;
        sAd[t] = __double2int_rn(M0 * x * 1024);
        sBd[t] = __double2int_rn(M3 * x * 1024);
      }
    }
  }
  __syncthreads();
  const uint32_t* pl = p.map_planes + (size_t)(p.n_maps == 1 ? 0 : e) * CASIM_NUM_PLANES * CASIM_PLANE_WORDS;
  const bool vec = (H & 3) == 0 && (((uintptr_t)out) & 3) == 0;
  const int wpr = vec ? (H >> 2) : H; /* output words (or bytes) per row */
  const int step = vec ? 4 : 1;
  for (int i = threadIdx.x; i < W * wpr; i += blockDim.x) {
    int oy = i / wpr, ox0 = (i - oy * wpr) * step;
    uint32_t word = 0;
#pragma unroll 4
    for (int k = 0; k < step; k++) {
      int ox = ox0 + k;
      int v = 0;
      if (oy < nrows && ox < ncols) {
        int X = (sX0[oy] + sAd[ox]) >> 5, Y = (sY0[oy] + sBd[ox]) >> 5;
        int sxi = X >> 5, syi = Y >> 5, fx = X & 31, fy = Y & 31;
        if (sxi >= 0 && sxi < CASIM_MAP_DIM && syi >= 0 && syi < CASIM_MAP_DIM) {
          int plane = (fx > 0 ? 1 : 0) + (fy > 0 ? 2 : 0);
          v = map_bit(pl + plane * CASIM_PLANE_WORDS, syi, sxi);
        } else { /* border: the footprint may still reach into the map from outside */
          auto S = [&](int r, int cc) -> int {
            return (r >= 0 && r < CASIM_MAP_DIM && cc >= 0 && cc < CASIM_MAP_DIM) ? map_bit(pl, r, cc) : 0;
          };
          v = S(syi, sxi);
          if (fx > 0) v |= S(syi, sxi + 1);
          if (fy > 0) v |= S(syi + 1, sxi);
          if (fx > 0 && fy > 0) v |= S(syi + 1, sxi + 1);
        }
      }
      word |= (uint32_t)v << (8 * k);
    }
    if (vec)
      reinterpret_cast<uint32_t*>(out)[i] = word;
    else
      out[i] = (uint8_t)word;
  }
}


/* ------------------------------------------------------------------------------------------------ */
/* k_grid_gather (per-env maps; the shared-map case runs k_grid_tma below): one WARP per agent, three phases.
 *   tables   per-row / per-column fixed-point terms of OpenCV's warpAffine (fp64, once per agent), biased by the
 *            96-cell padding of the fast-path planes so that no lookup needs a bounds check;
 *   phase 1  one lane per output word (4 cells of a destination row).  The source cells of the word lie between
 *            those of its end cells (the per-axis terms are monotone), i.e. inside the 6x6 box anchored at the
 *            component-wise minimum: the 2-bit classification plane says whether that box is entirely free
 *            (word = 0), entirely occupied (word = 0x01010101) or mixed (queued);
 *   phase 2  the queued words only: exact per-cell lookup in the footprint plane chosen by (fx > 0, fy > 0);
 *   phase 3  the 3600-byte grid leaves shared memory as 16-byte coalesced stores.
 * Maps are unions of rectangles (Map.get_occupancy_grid), so most words are uniform.                 */
#ifndef GRID_WARPS
#define GRID_WARPS 4
#endif
#ifndef GRID_MINB
#define GRID_MINB 8 /* resident CTAs per SM asked of the compiler: 8-10 measured 108 us at C-full, 12 (40-register cap) 115 us, 16 121 us */
#endif
#ifndef GRID_U
#define GRID_U 5
#endif
#define GRID_MAXW 16  /* words per row (submap side <= 64) */
#define GRID_MAXWORDS 1024 /* words per grid: the fast path takes grids of up to 4096 cells */

__device__ __forceinline__ uint32_t grid_cell_exact(const uint32_t* __restrict__ gp, int X, int Y) {
  /* X, Y: biased 1/32-cell fixed point (already >> 5) */
  int xi = X >> 5, yi = Y >> 5;
  int plane = ((X & 31) ? 1 : 0) + ((Y & 31) ? 2 : 0);
  uint32_t w = __ldg(gp + plane * CASIM_GPLANE_WORDS + yi * CASIM_GWORDS + (xi >> 5));
  return (w >> (xi & 31)) & 1u;
}

__global__ void __launch_bounds__(32 * GRID_WARPS, GRID_MINB) k_grid_gather(const GridArgs p) {
  casim_pdl_enter();
  const DevCfg& c = p.c;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const long long N = (long long)p.E * p.A;
  const long long g = (long long)blockIdx.x * GRID_WARPS + wid;
  __shared__ __align__(16) uint8_t s_out[GRID_WARPS][GRID_MAXWORDS]; /* one nibble per output word: bit k = cell k */
  __shared__ __align__(16) int2 s_r[GRID_WARPS][64];                 /* per destination row: (X, Y) terms */
  __shared__ __align__(16) int s_cx[GRID_WARPS][64], s_cy[GRID_WARPS][64]; /* per destination column */
  __shared__ uint16_t s_list[GRID_WARPS][GRID_MAXWORDS];
  if (g >= N) return;
  const int e = (int)(g / p.A), ai = (int)(g - (long long)e * p.A);
  const int n = p.n_agents ? min(p.n_agents[e], p.A) : p.A;
  const int H = c.submap_h, W = c.submap_w; /* out[oy*H + ox]: W rows of H cells */
  const int wpr = H >> 2, nwords = W * wpr;
  uint8_t* out = s_out[wid];
  uint4* gout = reinterpret_cast<uint4*>(p.out + g * (long long)(H * W));
  const int type = __float_as_int(p.state[5 * N + g].w);
  const bool on = ai < n && (((type >> 16) & 0xff) & CASIM_SENSOR_OCCUPANCY_GRID) && p.n_maps > 0;
  if (!on) {
    for (int i = lane; i < (nwords >> 2); i += 32) gout[i] = make_uint4(0u, 0u, 0u, 0u);
    return;
  }
  const double2 q0 = reinterpret_cast<const double2*>(p.pose64 + 4 * g)[0];
  const double2 q1 = reinterpret_cast<const double2*>(p.pose64 + 4 * g)[1];
  const double px = q0.x, py = q0.y, ch = q1.x, sh = q1.y;
  int row0, col0;
  world_to_map_mul(c, px, py, row0, col0);
  /* Map.getSubmapByIndices (Map.py:81-105) */
  int sx = max(row0 - W / 2, 0), sy = max(col0 - H / 2, 0);
  if (sx + W > CASIM_MAP_DIM - 1) sx = CASIM_MAP_DIM - 1 - W;
  if (sy + H > CASIM_MAP_DIM - 1) sy = CASIM_MAP_DIM - 1 - H;
  /* cv2.getRotationMatrix2D(center=(col0,row0), angle=-heading*180/pi, 1), inverted as in warpAffine */
  double M0, M1, M2, M3, M4, M5;
  {
    double alpha = ch, beta = -sh;
    double cx = (double)(float)col0, cy = (double)(float)row0;
    M0 = alpha; M1 = beta; M2 = (1 - alpha) * cx - beta * cy;
    M3 = -beta; M4 = alpha; M5 = beta * cx + (1 - alpha) * cy;
    double D = M0 * M4 - M1 * M3;
    D = D != 0 ? 1. / D : 0;
    double A11 = M4 * D, A22 = M0 * D;
    M0 = A11; M1 *= -D; M3 *= -D; M4 = A22;
    double b1 = -M0 * M2 - M1 * M5, b2 = -M3 * M2 - M4 * M5;
    M2 = b1; M5 = b2;
  }
  const int BIAS = CASIM_GPAD << 10; /* 96 cells in 1/1024 units: shifts >> 5 >> 5 keep it exact */
  bool ok = true;                    /* every lookup stays inside the padded planes */
  for (int t = lane; t < 64; t += 32) {
    double rxd = 0, ryd = 0, cxd = 0, cyd = 0;
    if (t < W) {
      int y = sx + t;
      rxd = rint((M1 * y + M2) * 1024);
      ryd = rint((M4 * y + M5) * 1024);
    }
    if (t < H) {
      int x = sy + t;
      cxd = rint(M0 * x * 1024);
      cyd = rint(M3 * x * 1024);
    }
    ok = ok && fabs(rxd) < 5.0e8 && fabs(ryd) < 5.0e8 && fabs(cxd) < 5.0e8 && fabs(cyd) < 5.0e8;
    s_r[wid][t] = make_int2((int)rxd + 16 + BIAS, (int)ryd + 16 + BIAS);
    s_cx[wid][t] = (int)cxd;
    s_cy[wid][t] = (int)cyd;
  }
  __syncwarp();
  /* the four corner cells bound every source coordinate (monotone terms) */
  if (lane < 4) {
    int oy = (lane & 1) ? W - 1 : 0, ox = (lane & 2) ? H - 1 : 0;
    int xi = (s_r[wid][oy].x + s_cx[wid][ox]) >> 10, yi = (s_r[wid][oy].y + s_cy[wid][ox]) >> 10;
    ok = ok && xi >= 1 && yi >= 1 && xi < CASIM_GWORDS * 32 - CASIM_GBOX - 1 && yi < CASIM_GROWS - CASIM_GBOX - 1;
  }
  ok = __all_sync(FULL_MASK, ok);
  if (!ok) {
    /* far outside the map (agent off the grid): per-cell path with bounds checks on the unpadded plane 0 */
    const uint32_t* pl = p.map_planes + (size_t)(p.n_maps == 1 ? 0 : e) * CASIM_NUM_PLANES * CASIM_PLANE_WORDS;
    auto S = [&](int r, int cc) -> uint32_t {
      return (r >= 0 && r < CASIM_MAP_DIM && cc >= 0 && cc < CASIM_MAP_DIM) ? (uint32_t)map_bit(pl, r, cc) : 0u;
    };
    for (int wi = lane; wi < nwords; wi += 32) {
      int oy = wi / wpr, ox0 = (wi - oy * wpr) << 2;
      uint32_t word = 0;
      for (int k = 0; k < 4; k++) {
        long long Xl = ((long long)s_r[wid][oy].x - BIAS + s_cx[wid][ox0 + k]) >> 5;
        long long Yl = ((long long)s_r[wid][oy].y - BIAS + s_cy[wid][ox0 + k]) >> 5;
        int X = (int)Xl, Y = (int)Yl;
        int sxi = X >> 5, syi = Y >> 5, fx = X & 31, fy = Y & 31;
        uint32_t v = S(syi, sxi);
        if (fx > 0) v |= S(syi, sxi + 1);
        if (fy > 0) v |= S(syi + 1, sxi);
        if (fx > 0 && fy > 0) v |= S(syi + 1, sxi + 1);
        word |= v << k;
      }
      out[wi] = (uint8_t)word;
    }
  } else {
    const uint32_t* gp = p.grid_maps + (size_t)(p.n_maps == 1 ? 0 : e) * CASIM_GMAP_WORDS;
    const uint32_t* cls = gp + CASIM_NUM_PLANES * CASIM_GPLANE_WORDS;
    int nmixed = 0;
    /* ---- phase 1: classify every word.  A lane keeps its word column (its column terms stay in registers) and
     * walks down the rows: rpi rows of wpr words per iteration. ---- */
    const int rpi = 32 / wpr;
    const int lr = lane / wpr, wx = lane - lr * wpr;
    const bool act = lr < rpi;
    /* min over the word's end cells of (row term + column term) = row term + min of the two column terms */
    const int cmx = min(s_cx[wid][4 * wx], s_cx[wid][4 * wx + 3]), cmy = min(s_cy[wid][4 * wx], s_cy[wid][4 * wx + 3]);
    constexpr int U = GRID_U; /* rows in flight per lane: the classification loads of a batch are independent */
    for (int oy0 = 0; oy0 < W; oy0 += rpi * U) {
      uint32_t cw[U];
      int sh[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int oy = oy0 + u * rpi + lr;
        cw[u] = 0u;
        sh[u] = 0;
        if (act && oy < W) {
          const int2 r = s_r[wid][oy];
          const int xm = (r.x + cmx) >> 10, ym = (r.y + cmy) >> 10;
          cw[u] = __ldg(cls + ym * CASIM_GCLS_WORDS + (xm >> 4));
          sh[u] = (xm & 15) << 1;
        }
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int oy = oy0 + u * rpi + lr;
        const bool valid = act && oy < W;
        const uint32_t code = (cw[u] >> sh[u]) & 3u;
        const bool mixed = valid && code == 0u;
        if (valid && !mixed) out[oy * wpr + wx] = (code & 2u) ? (uint8_t)0xF : (uint8_t)0;
        unsigned bal = __ballot_sync(FULL_MASK, mixed);
        if (mixed) s_list[wid][nmixed + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)((oy << 4) | wx);
        nmixed += __popc(bal);
      }
    }
    __syncwarp();
    /* ---- phase 2: exact cells of the mixed words ---- */
    for (int m = lane; m < nmixed; m += 32) {
      const int en = s_list[wid][m];
      const int oy = en >> 4, w4 = en & 15;
      const int2 r = s_r[wid][oy];
      const int4 cx = *reinterpret_cast<const int4*>(&s_cx[wid][4 * w4]);
      const int4 cy = *reinterpret_cast<const int4*>(&s_cy[wid][4 * w4]);
      uint32_t word = grid_cell_exact(gp, (r.x + cx.x) >> 5, (r.y + cy.x) >> 5);
      word |= grid_cell_exact(gp, (r.x + cx.y) >> 5, (r.y + cy.y) >> 5) << 1;
      word |= grid_cell_exact(gp, (r.x + cx.z) >> 5, (r.y + cy.z) >> 5) << 2;
      word |= grid_cell_exact(gp, (r.x + cx.w) >> 5, (r.y + cy.w) >> 5) << 3;
      out[oy * wpr + w4] = (uint8_t)word;
    }
  }
  __syncwarp();
  /* ---- phase 3: coalesced copy-out ---- */
  /* nibble -> four 0/1 bytes: bit k of b lands on bit 8k of b * (1 + 2^7 + 2^14 + 2^21) */
  const uint32_t* so = reinterpret_cast<const uint32_t*>(out);
  for (int i = lane; i < (nwords >> 2); i += 32) {
    const uint32_t q = so[i];
    uint4 v;
    v.x = ((q & 0xFu) * 0x00204081u) & 0x01010101u;
    v.y = (((q >> 8) & 0xFu) * 0x00204081u) & 0x01010101u;
    v.z = (((q >> 16) & 0xFu) * 0x00204081u) & 0x01010101u;
    v.w = (((q >> 24) & 0xFu) * 0x00204081u) & 0x01010101u;
    gout[i] = v;
  }
}


/* ------------------------------------------------------------------------------------------------ */
/* k_grid_tma: OccupancyGridSensor for a map SHARED by all environments (n_maps == 1).
 *
 * Persistent kernel, one CTA per SM.  The CTA first stages the WHOLE padded map -- four footprint bit planes and one
 * classification byte per even anchor, 110 592 B (casim_internal.h CASIM_G2_*) -- into shared memory with nine TMA tile
 * loads (cp.async.bulk.tensor.2d on one mbarrier), issued before the programmatic-dependent-launch wait so that they
 * overlap the tail of the preceding kernel.  From then on no map lookup leaves the SM.  Warps then pull agents off a
 * global work counter (dynamic: agents near obstacle corners cost more) and produce each agent's grid in three passes:
 *   tables   the per-row / per-column fixed-point terms of OpenCV's warpAffine (fp64, once per agent);
 *   pass 1   one lane per 4x4 destination TILE.  Its source cells lie in the 8x8 box anchored at the even cell below the
 *            component-wise minimum of its corner cells (the per-axis terms are monotone and a tile spans < 5 cells per
 *            axis).  The box's classification byte says: all free / all occupied -> 4 words written at once; a single
 *            straight edge (every row, or every column, of the box identical) -> queued as an EDGE tile; else queued
 *            as OTHER;
 *   pass 2   edge tiles, one lane per 4-cell word: with an axis-aligned edge a cell's value is ONE comparison of its
 *            own fixed-point coordinate with the edge (footprint max column >= edge, or min column < edge, resp. rows) --
 *            exact, no lookup;
 *   pass 3   other tiles (corners, thin walls, noise): exact per-cell lookup in the footprint plane chosen by (fx>0, fy>0);
 *   copy-out the grid leaves shared memory as 16-byte coalesced stores (+ optional np.packbits form, 2 bytes per lane).
 * Maps are unions of rectangles (Map.get_occupancy_grid), so nearly every non-uniform tile is an edge tile; the result
 * is exact for ANY map.  Agents whose rotated window leaves the padded planes (within ~1 cell of the map border, or off
 * the map) take a bounds-checked per-cell path on the unpadded plane 0 in global memory. */
#ifndef G2_WARPS
#define G2_WARPS 32 /* 1024 threads, <= 64 registers: the kernel is latency-bound below ~8 warps per scheduler */
#endif
#define G2_THREADS (32 * G2_WARPS)
/* per warp: tiles[256] uint16 (bit 4r+k = cell (row r, column k) of the 4x4 tile), s_rx, s_ry, s_cx, s_cy [64] int,
 * s_rmin[16] int2, s_cmin[16] int2, queue[256] uint16 */
#define G2_WARP_BYTES (256 * 2 + 4 * 64 * 4 + 16 * 8 + 16 * 8 + 256 * 2)
/* + mbarrier (8 B) + ticket counter (4 B) + the laser tables: cos, sin of the beam angles, range samples (doubles),
 * reported values (floats) */
#define G2_LASER_OFF (CASIM_G2_BYTES + G2_WARPS * G2_WARP_BYTES + 16)
#define G2_SMEM_BYTES (G2_LASER_OFF + 3 * CASIM_MAX_BEAMS * 8 + (CASIM_MAX_RANGES + 1) * 4 + 12)

struct Grid2Args {
  DevCfg c;
  const float4* state;
  const double* pose64;
  const int32_t* n_agents;
  const uint32_t* map_planes; /* unpadded plane 0 (slow path) */
  uint8_t* out;
  uint8_t* packed;            /* optional np.packbits form, or NULL */
  float* laser;               /* LaserScanSensor output [E*A][nb] when the laser rides in this kernel, else NULL */
  int lg2P;                   /* laser: lanes per beam = 1 << lg2P */
  int E, A;
  unsigned inv_A;             /* floor(2^32 / A) + 1 when E * A < 2^32 / A (then umulhi(g, inv_A) == g / A for every agent), else 0 */
};

__device__ __forceinline__ uint32_t g2_cell(uint32_t a_map, int fx, int fy) {
  /* fx, fy: biased source coordinates in 1/1024 cell (OpenCV's X0 + adelta, before the >> 5); a_map: shared-space
   * address of the staged planes */
  const int xi = fx >> 10, yi = fy >> 10;
  const int plane = ((fx & 0x3E0) ? CASIM_G2_PLANE_WORDS : 0) + ((fy & 0x3E0) ? 2 * CASIM_G2_PLANE_WORDS : 0);
  uint32_t w;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(a_map + 4u * (uint32_t)(plane + yi * CASIM_G2_WORDS + (xi >> 5))));
  return (w >> (xi & 31)) & 1u;
}

__device__ __forceinline__ void g2_wait_map(uint32_t bar) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(0u)
        : "memory");
  }
}

/* 4 cells of one tile row against a threshold: bit k = (c_k < t) */
__device__ __forceinline__ uint32_t g2_row_lt(const int4 cc, int t) {
  return ((uint32_t)(cc.x - t) >> 31) | (((uint32_t)(cc.y - t) >> 31) << 1) | (((uint32_t)(cc.z - t) >> 31) << 2) |
         (((uint32_t)(cc.w - t) >> 31) << 3);
}

/* CTW, CTH: tile columns / rows as compile-time constants (15 x 15 = the reference's 60 x 60 submap), 0 = run time */
/* Shared-memory accessors on explicit 32-bit shared-space addresses.  With ordinary pointers into the dynamic shared
 * array the compiler rebuilt the CTA's shared-window base (S2UR SR_CgaCtaId / ULEA) in front of almost every access
 * (9.5 % of the kernel's instructions, profiles/r2g); the addresses below are formed once. */
/* OFF = byte offset folded into the instruction's immediate field (the per-warp arrays sit at fixed offsets from a_w) */
template <int OFF = 0> __device__ __forceinline__ int g2_ld32(uint32_t a) { int v; asm volatile("ld.shared.s32 %0, [%1+%2];" : "=r"(v) : "r"(a), "n"(OFF)); return v; }
template <int OFF = 0> __device__ __forceinline__ uint32_t g2_ldu16(uint32_t a) { uint32_t v; asm volatile("ld.shared.u16 %0, [%1+%2];" : "=r"(v) : "r"(a), "n"(OFF)); return v; }
template <int OFF = 0> __device__ __forceinline__ uint32_t g2_ldu8(uint32_t a) { uint32_t v; asm volatile("ld.shared.u8 %0, [%1+%2];" : "=r"(v) : "r"(a), "n"(OFF)); return v; }
template <int OFF = 0> __device__ __forceinline__ int2 g2_ld64(uint32_t a) { int2 v; asm volatile("ld.shared.v2.s32 {%0, %1}, [%2+%3];" : "=r"(v.x), "=r"(v.y) : "r"(a), "n"(OFF)); return v; }
template <int OFF = 0> __device__ __forceinline__ int4 g2_ld128(uint32_t a) {
  int4 v;
  asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4+%5];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a), "n"(OFF));
  return v;
}
template <int OFF = 0> __device__ __forceinline__ void g2_st32(uint32_t a, int v) { asm volatile("st.shared.s32 [%0+%2], %1;" ::"r"(a), "r"(v), "n"(OFF) : "memory"); }
template <int OFF = 0> __device__ __forceinline__ void g2_st64(uint32_t a, int2 v) { asm volatile("st.shared.v2.s32 [%0+%3], {%1, %2};" ::"r"(a), "r"(v.x), "r"(v.y), "n"(OFF) : "memory"); }
template <int OFF = 0> __device__ __forceinline__ void g2_stu16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0+%2], %1;" ::"r"(a), "h"((uint16_t)v), "n"(OFF) : "memory"); }
/* byte offsets of the per-warp arrays from the warp's staging base a_w */
enum { G2_TILE = 0, G2_RX = 512, G2_RY = 768, G2_CX = 1024, G2_CY = 1280, G2_RMIN = 1536, G2_CMIN = 1664, G2_Q = 1792 };

template <int CTW, int CTH>
__global__ void __launch_bounds__(G2_THREADS, 1) k_grid_tma(const __grid_constant__ CUtensorMap tmap, const Grid2Args p) {
  extern __shared__ __align__(128) uint8_t g2_smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t a_map = (uint32_t)__cvta_generic_to_shared(g2_smem); /* 4 footprint planes, then the classification bytes */
  const uint32_t a_cls = a_map + CASIM_NUM_PLANES * CASIM_G2_PLANE_WORDS * 4;
  const uint32_t a_w = a_map + CASIM_G2_BYTES + wid * G2_WARP_BYTES;    /* this warp's staging area */
  const uint32_t bar = a_map + CASIM_G2_BYTES + G2_WARPS * G2_WARP_BYTES;
  unsigned* s_ticket = reinterpret_cast<unsigned*>(g2_smem + CASIM_G2_BYTES + G2_WARPS * G2_WARP_BYTES + 8);

  /* ---- stage the map: the loads do not depend on the preceding kernels, so they go out before the PDL wait ---- */
  if (threadIdx.x == 0) {
    *s_ticket = 0u;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)CASIM_G2_BYTES) : "memory");
    const uint32_t dst = a_map;
#pragma unroll 1
    for (int k = 0; k < CASIM_G2_BOXES; k++)
      asm volatile(
          "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
              dst + k * CASIM_G2_BOX_ROWS * CASIM_G2_ROW_BYTES),
          "l"(&tmap), "r"(0), "r"(k * CASIM_G2_BOX_ROWS), "r"(bar)
          : "memory");
  }
  const DevCfg& c = p.c;
  /* laser tables (lane-dependent indices: shared memory, not the constant bank) */
  double* s_lcos = reinterpret_cast<double*>(g2_smem + G2_LASER_OFF);
  double* s_lsin = s_lcos + CASIM_MAX_BEAMS;
  double* s_lr = s_lsin + CASIM_MAX_BEAMS;
  float* s_lout = reinterpret_cast<float*>(s_lr + CASIM_MAX_RANGES);
  if (p.laser != nullptr) {
    static_assert(CASIM_MAX_BEAMS == CASIM_MAX_RANGES, "one fill loop for both tables");
    if (threadIdx.x < CASIM_MAX_BEAMS) {
      s_lcos[threadIdx.x] = c.laser_cos[threadIdx.x];
      s_lsin[threadIdx.x] = c.laser_sin[threadIdx.x];
      s_lr[threadIdx.x] = c.laser_r[threadIdx.x];
    }
    if (threadIdx.x <= CASIM_MAX_RANGES) s_lout[threadIdx.x] = c.laser_out[threadIdx.x];
  }
  casim_pdl_enter();
  __syncthreads(); /* the barrier word is initialised for everybody */

  const unsigned N = (unsigned)p.E * (unsigned)p.A;
  const int H = CTW ? 4 * CTW : c.submap_h, W = CTH ? 4 * CTH : c.submap_w; /* out[oy*H + ox]: W rows of H cells */
  const int TW = H >> 2, TH = W >> 2;                                       /* 4x4 tiles */
  const int NT = TW * TH;
  const int cells = H * W;
  const int BIAS = CASIM_G2_PAD << 10;
  /* Work distribution: the CTA owns a contiguous block of agents (all CTAs the same size +-1: the sum of ~277 agents'
   * costs varies by a few per cent only); inside the CTA the warps take agents off a SHARED-MEMORY ticket counter, so a
   * warp that drew expensive agents (obstacle corners) simply takes fewer.  (One global counter for all 4736 warps
   * serialises on its L2 atomic unit: measured 27 % of the stall samples.) */
  const unsigned lo = (unsigned)(((unsigned long long)N * blockIdx.x) / gridDim.x);
  const unsigned hi = (unsigned)(((unsigned long long)N * (blockIdx.x + 1)) / gridDim.x);
  auto ticket = [&]() -> unsigned {
    unsigned t = 0;
    if (lane == 0) t = lo + atomicAdd(s_ticket, 1u);
    t = __shfl_sync(FULL_MASK, t, 0);
    return t < hi ? t : N; /* N = no more work */
  };
  unsigned g = ticket();
  bool map_ready = false;
  /* an agent's inputs: loaded one agent ahead, so that their global-memory latency hides behind the previous agent */
  int in_type = 0, in_n = 0, in_ai = 0;
  float in_radius = 0.f;
  double2 in_q0 = make_double2(0., 0.), in_q1 = make_double2(0., 0.);
  auto fetch = [&](unsigned ag) {
    if (ag < N) {
      /* environment of the agent: one multiply-high instead of a division when the launcher could prove it exact */
      const int ee = p.inv_A ? (int)__umulhi(ag, p.inv_A) : (int)(ag / (unsigned)p.A);
      in_ai = (int)(ag - (unsigned)ee * (unsigned)p.A);
      in_n = p.n_agents ? min(p.n_agents[ee], p.A) : p.A;
      in_type = __float_as_int(p.state[5 * (size_t)N + ag].w);
      if (p.laser != nullptr) in_radius = p.state[2 * (size_t)N + ag].z;
      in_q0 = reinterpret_cast<const double2*>(p.pose64 + 4 * (size_t)ag)[0];
      in_q1 = reinterpret_cast<const double2*>(p.pose64 + 4 * (size_t)ag)[1];
    }
  };
  fetch(g);
  while (g < N) {
    const int ai = in_ai;
    const int n = in_n;
    uint32_t* gout = reinterpret_cast<uint32_t*>(p.out + (size_t)g * cells);
    const int type = in_type;
    const double2 q0 = in_q0, q1 = in_q1;
    const bool on = ai < n && (((type >> 16) & 0xff) & CASIM_SENSOR_OCCUPANCY_GRID);
    if (p.laser != nullptr) {
      /* ---- LaserScanSensor (LaserScanSensor.py:27-58) riding in this kernel: same arithmetic as k_laser above, the
       * 256 lattice samples read plane 0 of the map staged in shared memory ---- */
      const int nb = c.nb, nr = c.nr;
      float* lout = p.laser + (size_t)g * nb;
      if (!(ai < n && (((type >> 16) & 0xff) & CASIM_SENSOR_LASERSCAN))) {
        for (int b = lane; b < nb; b += 32) lout[b] = 0.f;
      } else {
        if (!map_ready) {
          g2_wait_map(bar);
          map_ready = true;
        }
        const double px = q0.x, py = q0.y, ch = q1.x, sh = q1.y;
        int r0, c0;
        const bool ego_in = world_to_map_mul(c, px, py, r0, c0);
        const double rc = (double)in_radius * c.inv_cell, rc2 = rc * rc;
        const int disc_thr = ego_in ? (int)fmin(ceil(rc2), 2.0e9) - 1 : -1; /* d^2 < rc2 <=> d^2 <= ceil(rc2) - 1 */
        const int P = 1 << p.lg2P;
        const int b = lane >> p.lg2P, part = lane & (P - 1);
        /* kd = samples that may fall into the ego disc: r_k / cell < rc + 1.5; a prefix when the r_k increase
         * (np.arange, LaserScanSensor.py:18-20), every sample otherwise */
        const int kd = !c.laser_sorted ? nr
                                       : __popc(__ballot_sync(FULL_MASK, lane < nr && disc_thr >= 0 &&
                                                                             s_lr[min(lane, CASIM_MAX_RANGES - 1)] * c.inv_cell < rc + 1.5));
        unsigned mask = 0;
        if (b < nb) {
          const double cs = s_lcos[b] * ch - s_lsin[b] * sh, sn = s_lsin[b] * ch + s_lcos[b] * sh;
          const double u0 = c.map_oy + px * c.inv_cell, du = cs * c.inv_cell;
          const double v0 = c.map_ox - py * c.inv_cell, dv = -(sn * c.inv_cell);
          /* Branch-free.  The staged plane is padded with free cells on every side, so an out-of-map sample needs no
           * bounds test: its (unsigned-)clamped address lands on a free padding cell.  The ego-disc exclusion (Map.py:73-79)
           * can only remove samples within rc + sqrt(2) cells of the agent (a cell index is within one cell of the point
           * it holds): only the first kd samples of a beam carry the test. */
          auto sample = [&](int k) -> uint32_t {
            const double rk = s_lr[k];
            const int col = __double2int_rd(fma(rk, du, u0)), row = __double2int_rd(fma(rk, dv, v0));
            const uint32_t pr = min((uint32_t)row + (uint32_t)CASIM_G2_PAD, (uint32_t)(CASIM_G2_DIM - 1));
            const uint32_t pc = min((uint32_t)col + (uint32_t)CASIM_G2_PAD, (uint32_t)(CASIM_G2_DIM - 1));
            uint32_t w;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(a_map + 4u * (pr * CASIM_G2_WORDS + (pc >> 5))));
            return (w >> (pc & 31u)) & 1u;
          };
          int k = part;
          for (; k < kd; k += P) { /* near the agent: with the disc test */
            const double rk = s_lr[k];
            const int col = __double2int_rd(fma(rk, du, u0)), row = __double2int_rd(fma(rk, dv, v0));
            const int dr = row - r0, dc = col - c0;
            mask |= ((dr * dr + dc * dc > disc_thr) ? sample(k) : 0u) << k;
          }
#pragma unroll 4
          for (; k < nr; k += P) mask |= sample(k) << k;
        }
        for (int o = 1; o < P; o <<= 1) mask |= __shfl_xor_sync(FULL_MASK, mask, o);
        if (b < nb && part == 0) {
          int idx = CASIM_MAX_RANGES; /* no hit: max range */
          if (mask) {
            const unsigned rest = mask & (mask - 1); /* drop the first hit */
            idx = rest ? (__ffs(rest) - 2) : (nr - 1);
          }
          lout[b] = s_lout[idx];
        }
      }
    }
    if (!on) {
      uint4* g4 = reinterpret_cast<uint4*>(gout);
      for (int i = lane; i < (cells >> 4); i += 32) g4[i] = make_uint4(0u, 0u, 0u, 0u);
      if (p.packed) {
        uint16_t* pk = reinterpret_cast<uint16_t*>(p.packed + (size_t)g * (cells >> 3));
        for (int i = lane; i < (cells >> 4); i += 32) pk[i] = 0;
      }
      g = ticket();
      fetch(g);
      continue;
    }
    {
      const double px = q0.x, py = q0.y, ch = q1.x, sh = q1.y;
      int row0, col0;
      world_to_map_mul(c, px, py, row0, col0);
      /* Map.getSubmapByIndices (Map.py:81-105) */
      int sx = max(row0 - W / 2, 0), sy = max(col0 - H / 2, 0);
      if (sx + W > CASIM_MAP_DIM - 1) sx = CASIM_MAP_DIM - 1 - W;
      if (sy + H > CASIM_MAP_DIM - 1) sy = CASIM_MAP_DIM - 1 - H;
      /* cv2.getRotationMatrix2D(center=(col0,row0), angle=-heading*180/pi, 1), inverted as in warpAffine */
      double M0, M1, M2, M3, M4, M5;
      {
        double alpha = ch, beta = -sh;
        double cx = (double)(float)col0, cy = (double)(float)row0;
        M0 = alpha; M1 = beta; M2 = (1 - alpha) * cx - beta * cy;
        M3 = -beta; M4 = alpha; M5 = beta * cx + (1 - alpha) * cy;
        double D = M0 * M4 - M1 * M3;
        D = D != 0 ? 1. / D : 0;
        double A11 = M4 * D, A22 = M0 * D;
        M0 = A11; M1 *= -D; M3 *= -D; M4 = A22;
        double b1 = -M0 * M2 - M1 * M5, b2 = -M3 * M2 - M4 * M5;
        M2 = b1; M5 = b2;
      }
      bool ok = true; /* every lookup stays inside the padded planes and every tile inside its 8x8 box */
#pragma unroll
      for (int t = lane; t < 64; t += 32) {
        double rxd = 0, ryd = 0, cxd = 0, cyd = 0;
        if (t < W) {
          int y = sx + t;
          rxd = rint((M1 * y + M2) * 1024);
          ryd = rint((M4 * y + M5) * 1024);
        }
        if (t < H) {
          int x = sy + t;
          cxd = rint(M0 * x * 1024);
          cyd = rint(M3 * x * 1024);
        }
        ok = ok && fabs(rxd) < 5.0e8 && fabs(ryd) < 5.0e8 && fabs(cxd) < 5.0e8 && fabs(cyd) < 5.0e8;
        g2_st32<G2_RX>(a_w + 4 * t, (int)rxd + 16 + BIAS);
        g2_st32<G2_RY>(a_w + 4 * t, (int)ryd + 16 + BIAS);
        g2_st32<G2_CX>(a_w + 4 * t, (int)cxd);
        g2_st32<G2_CY>(a_w + 4 * t, (int)cyd);
      }
      __syncwarp();
      int spx = 0, spy = 0, cpx = 0, cpy = 0;
      if (lane < 4) { /* the four corner cells bound every source coordinate (monotone terms) */
        const int oy = (lane & 1) ? W - 1 : 0, ox = (lane & 2) ? H - 1 : 0;
        const int xi = (g2_ld32<G2_RX>(a_w + 4 * oy) + g2_ld32<G2_CX>(a_w + 4 * ox)) >> 10, yi = (g2_ld32<G2_RY>(a_w + 4 * oy) + g2_ld32<G2_CY>(a_w + 4 * ox)) >> 10;
        ok = ok && xi >= 2 && yi >= 2 && xi <= CASIM_G2_DIM - 12 && yi <= CASIM_G2_DIM - 12;
      }
      if (lane < TH) { /* per tile row: minima of the row terms, and how far a tile reaches along each axis */
        const int ax = g2_ld32<G2_RX>(a_w + 16 * lane), bx = g2_ld32<G2_RX>(a_w + 16 * lane + 12), ay = g2_ld32<G2_RY>(a_w + 16 * lane), by = g2_ld32<G2_RY>(a_w + 16 * lane + 12);
        g2_st64<G2_RMIN>(a_w + 8 * lane, make_int2(min(ax, bx), min(ay, by)));
        spx = abs(ax - bx);
        spy = abs(ay - by);
      }
      if (lane < TW) {
        const int ax = g2_ld32<G2_CX>(a_w + 16 * lane), bx = g2_ld32<G2_CX>(a_w + 16 * lane + 12), ay = g2_ld32<G2_CY>(a_w + 16 * lane), by = g2_ld32<G2_CY>(a_w + 16 * lane + 12);
        g2_st64<G2_CMIN>(a_w + 8 * lane, make_int2(min(ax, bx), min(ay, by)));
        cpx = abs(ax - bx);
        cpy = abs(ay - by);
      }
      /* a tile's source cells span < 5 cells per axis (3 (|cos| + |sin|) <= 4.25 for a rotation): floor indices differ by
       * at most 5, +1 for the bilinear footprint, +1 for the even anchor = the 8x8 box */
      spx = __reduce_max_sync(FULL_MASK, spx) + __reduce_max_sync(FULL_MASK, cpx);
      spy = __reduce_max_sync(FULL_MASK, spy) + __reduce_max_sync(FULL_MASK, cpy);
      ok = __all_sync(FULL_MASK, ok) && max(spx, spy) < 5 * 1024;
      __syncwarp();
      const unsigned gcur = g;
      g = ticket(); /* the next agent's loads travel while this one is computed */
      fetch(g);
      if (!map_ready) { /* first agent of this warp: the staged map must have landed */
        g2_wait_map(bar);
        map_ready = true;
      }
      if (!ok) {
        /* rotated window partly outside the padded planes: per-cell path with bounds checks on the unpadded plane 0 */
        const uint32_t* pl = p.map_planes;
        auto S = [&](int r, int cc) -> uint32_t {
          return (r >= 0 && r < CASIM_MAP_DIM && cc >= 0 && cc < CASIM_MAP_DIM) ? (uint32_t)map_bit(pl, r, cc) : 0u;
        };
        for (int ti = lane; ti < NT; ti += 32) {
          const int ty = ti / TW, tx = ti - ty * TW;
          uint32_t bits = 0;
          for (int k = 0; k < 16; k++) {
            const int oy = 4 * ty + (k >> 2), ox = 4 * tx + (k & 3);
            long long Xl = ((long long)g2_ld32<G2_RX>(a_w + 4 * oy) - BIAS + g2_ld32<G2_CX>(a_w + 4 * ox)) >> 5;
            long long Yl = ((long long)g2_ld32<G2_RY>(a_w + 4 * oy) - BIAS + g2_ld32<G2_CY>(a_w + 4 * ox)) >> 5;
            int X = (int)Xl, Y = (int)Yl;
            int sxi = X >> 5, syi = Y >> 5, fx = X & 31, fy = Y & 31;
            uint32_t v = S(syi, sxi);
            if (fx > 0) v |= S(syi, sxi + 1);
            if (fy > 0) v |= S(syi + 1, sxi);
            if (fx > 0 && fy > 0) v |= S(syi + 1, sxi + 1);
            bits |= v << k;
          }
          g2_stu16<G2_TILE>(a_w + 2 * ti, bits);
        }
      } else {
        /* ---- pass 1: one lane per 4x4 tile: uniform tiles are done, the others are queued with their code ---- */
        int nq = 0;
        for (int t0 = 0; t0 < NT; t0 += 32) {
          const int ti = t0 + lane;
          const bool valid = ti < NT;
          const int ty = ti / TW, tx = ti - ty * TW;
          uint32_t code = 1u;
          if (valid) {
            const int2 rm = g2_ld64<G2_RMIN>(a_w + 8 * ty), cm = g2_ld64<G2_CMIN>(a_w + 8 * tx);
            code = g2_ldu8(a_cls + ((rm.y + cm.y) >> 11) * CASIM_G2_CLS_DIM + ((rm.x + cm.x) >> 11)); /* even anchor / 2 */
          }
          const uint32_t kind = code & 7u;
          const bool queued = kind != 1u && kind != 2u;
          if (valid && !queued) g2_stu16<G2_TILE>(a_w + 2 * ti, kind == 2u ? 0xFFFFu : 0u);
          const unsigned bq = __ballot_sync(FULL_MASK, queued);
          if (queued) g2_stu16<G2_Q>(a_w + 2 * (nq + __popc(bq & ((1u << lane) - 1u))), ti | (code << 8));
          nq += __popc(bq);
        }
        __syncwarp();
        /* ---- pass 2: the queued tiles, one lane per tile.  EDGE tile (kinds 3-6: every row, or every column, of the 8x8
         * box identical): one comparison per cell, no lookup.  OTHER tile: exact per-cell lookups. ---- */
        int no = 0; /* OTHER tiles: compacted in place to the front of the queue (slot no + rank <= the slot just read) */
        for (int m0 = 0; m0 < nq; m0 += 32) {
          const int m = m0 + lane;
          const uint32_t en = m < nq ? g2_ldu16<G2_Q>(a_w + 2 * m) : (1u << 8);
          const int ti = (int)(en & 255u);
          const int ty = ti / TW, tx = ti - ty * TW;
          const uint32_t kind = (en >> 8) & 7u, pp = (en >> 11) & 7u;
          const bool other = kind == 0u;
          const unsigned bo = __ballot_sync(FULL_MASK, other);
          if (other) g2_stu16<G2_Q>(a_w + 2 * (no + __popc(bo & ((1u << lane) - 1u))), ti); /* the ballot orders every lane's read first */
          no += __popc(bo);
          if (kind >= 3u) {
            uint32_t bits;
            const bool horiz = kind >= 5u, ge = (kind & 1u) != 0u; /* 3, 5: ">= p occupied"; 4, 6: "< p occupied" */
            const int sel = horiz ? 4 : 0; /* .y halves of the (x, y) pairs */
            const int anc = ((g2_ld32<G2_RMIN>(a_w + 8 * ty + sel) + g2_ld32<G2_CMIN>(a_w + 8 * tx + sel)) >> 11) << 1;
            /* ge: footprint max index = (v + 992) >> 10 >= anc + p  <=>  v >= (anc + p) * 1024 - 992
             * lt: footprint min index = v >> 10 < anc + p           <=>  v <  (anc + p) * 1024            */
            const int thr = ((anc + (int)pp) << 10) - (ge ? 992 : 0);
            const int4 rr = g2_ld128<G2_RX>(a_w + (horiz ? G2_RY - G2_RX : 0) + 16 * ty), cc = g2_ld128<G2_CX>(a_w + (horiz ? G2_CY - G2_CX : 0) + 16 * tx);
            bits = g2_row_lt(cc, thr - rr.x) | (g2_row_lt(cc, thr - rr.y) << 4) | (g2_row_lt(cc, thr - rr.z) << 8) |
                   (g2_row_lt(cc, thr - rr.w) << 12);
            if (ge) bits ^= 0xFFFFu;
            g2_stu16<G2_TILE>(a_w + 2 * ti, bits);
          }
        }
        __syncwarp();
        /* ---- pass 3: OTHER tiles, one lane per tile ROW (4 cells): exact lookups in the staged footprint planes; the
         * four lanes of a tile combine their nibbles with shuffles ---- */
        for (int m0 = 0; m0 < 4 * no; m0 += 32) {
          const int m = m0 + lane;
          const bool live = m < 4 * no;
          const int ti = live ? (int)g2_ldu16<G2_Q>(a_w + 2 * (m >> 2)) : 0;
          const int ty = ti / TW, tx = ti - ty * TW, r = m & 3;
          const int rxx = g2_ld32<G2_RX>(a_w + 4 * (4 * ty + r)), ryy = g2_ld32<G2_RY>(a_w + 4 * (4 * ty + r));
          const int4 cx = g2_ld128<G2_CX>(a_w + 16 * tx), cy = g2_ld128<G2_CY>(a_w + 16 * tx);
          uint32_t nib = g2_cell(a_map, rxx + cx.x, ryy + cy.x);
          nib |= g2_cell(a_map, rxx + cx.y, ryy + cy.y) << 1;
          nib |= g2_cell(a_map, rxx + cx.z, ryy + cy.z) << 2;
          nib |= g2_cell(a_map, rxx + cx.w, ryy + cy.w) << 3;
          nib <<= 4 * r;
          nib |= __shfl_xor_sync(FULL_MASK, nib, 1);
          nib |= __shfl_xor_sync(FULL_MASK, nib, 2);
          if (live && r == 0) g2_stu16<G2_TILE>(a_w + 2 * ti, nib);
        }
      }
      __syncwarp();
      /* ---- copy-out, one lane per tile: 4 words (one per row; the lanes of a tile row write 4 TW contiguous bytes).
       * nibble -> four 0/1 bytes: bit k of b lands on bit 8k of b * (1 + 2^7 + 2^14 + 2^21) ---- */
      for (int t0 = 0; t0 < NT; t0 += 32) {
        const int ti = t0 + lane;
        if (ti < NT) {
          const int ty = ti / TW, tx = ti - ty * TW;
          const uint32_t q = g2_ldu16<G2_TILE>(a_w + 2 * ti);
          uint32_t* o = gout + (4 * ty) * TW + tx;
          o[0] = ((q & 0xFu) * 0x00204081u) & 0x01010101u;
          o[TW] = (((q >> 4) & 0xFu) * 0x00204081u) & 0x01010101u;
          o[2 * TW] = (((q >> 8) & 0xFu) * 0x00204081u) & 0x01010101u;
          o[3 * TW] = ((q >> 12) * 0x00204081u) & 0x01010101u;
        }
      }
      if (p.packed) { /* np.packbits: 16 cells of a row -> 2 bytes, first cell in the most significant bit */
        uint16_t* pk = reinterpret_cast<uint16_t*>(p.packed + (size_t)gcur * (cells >> 3));
        for (int i = lane; i < (cells >> 4); i += 32) {
          uint32_t v = 0; /* nibbles of words 4i .. 4i+3 (row-major word index), word w = (oy, wx) -> tile (oy/4, wx), row oy%4 */
#pragma unroll
          for (int k = 0; k < 4; k++) {
            const int w = 4 * i + k, oy = w / TW, wx = w - oy * TW;
            const uint32_t nib = (g2_ldu16<G2_TILE>(a_w + 2 * ((oy >> 2) * TW + wx)) >> (4 * (oy & 3))) & 0xFu;
            v |= (__brev(nib) >> 28) << (4 * (3 - k)); /* cell 0 of the nibble to its top bit; first word in the top nibble */
          }
          pk[i] = (uint16_t)(((v >> 8) & 0xFFu) | ((v & 0xFFu) << 8)); /* byte 0 = cells 0-7 */
        }
      }
      __syncwarp(); /* the staging area is reused by the next agent */
    }
  }
  if (!map_ready) g2_wait_map(bar); /* nobody leaves while the bulk copies into this CTA's shared memory are in flight */
}

/* np.packbits of the dense grids, for the kernels that do not emit it themselves (per-env maps, odd submap sizes) */
__global__ void k_grid_pack(const uint8_t* __restrict__ dense, uint8_t* __restrict__ packed, size_t n_cells, int cells) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; /* one output byte */
  const int per = (cells + 7) >> 3;
  if (i >= (n_cells / cells) * (size_t)per) return;
  const size_t ag = i / per;
  const int b = (int)(i - ag * per);
  uint32_t v = 0;
  for (int k = 0; k < 8; k++) {
    const int cidx = 8 * b + k;
    v = (v << 1) | ((cidx < cells && dense[ag * cells + cidx]) ? 1u : 0u);
  }
  packed[i] = (uint8_t)v;
}

}  // namespace

cudaError_t casim_launch_laser(casim_handle* h, const float* state, float* laserscan, cudaStream_t s) {
  LaserArgs a;
  a.c = h->dc;
  a.state = reinterpret_cast<const float4*>(state);
  a.pose64 = h->pose64;
  a.n_agents = h->n_agents;
  a.map_planes = h->map_planes;
  a.grid_maps = h->grid_maps;
  a.n_maps = h->n_maps;
  a.out = laserscan;
  a.E = h->E;
  a.A = h->A;
  long long N = (long long)h->E * h->A;
  if (N >= (1ll << 31)) return cudaErrorInvalidValue;
  int nbp = 1, lg = 5;
  while (nbp < h->dc.nb) nbp <<= 1;
  for (int t = nbp; t > 1; t >>= 1) lg--;
  a.lg2P = lg;
  int wpc = LASER_WPC;
  dim3 grid((unsigned)((N + wpc - 1) / wpc)), block(32 * wpc);
  h->launches++;
  if (h->n_maps == 1 && h->grid_maps) return casim_launch_pdl<LaserArgs>(k_laser<true>, grid, block, 0, s, a);
  return casim_launch_pdl<LaserArgs>(k_laser<false>, grid, block, 0, s, a);
}

/* The tensor map of the shared-map layout: a 2-D uint32 tensor of CASIM_G2_ROWS rows x 12 words, loaded in boxes of 256
 * rows.  cuTensorMapEncodeTiled comes from the driver through the runtime's entry-point query (no -lcuda). */
cudaError_t casim_grid2_encode_tmap(casim_handle* h) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  if (!encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess) return e;
    if (qres != cudaDriverEntryPointSuccess || !fn) return cudaErrorNotSupported;
    encode = reinterpret_cast<EncodeFn>(fn);
  }
  static_assert(sizeof(CUtensorMap) <= sizeof(h->grid2_tmap), "tensor map does not fit its slot in the handle");
  const cuuint64_t dims[2] = {CASIM_G2_WORDS, CASIM_G2_ROWS};
  const cuuint64_t strides[1] = {CASIM_G2_ROW_BYTES};
  const cuuint32_t box[2] = {CASIM_G2_WORDS, CASIM_G2_BOX_ROWS};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = encode(reinterpret_cast<CUtensorMap*>(h->grid2_tmap), CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, h->grid2_map, dims,
                      strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                      CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    casim_set_error("cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
    return cudaErrorInvalidValue;
  }
  h->grid2_ready = 1;
  return cudaSuccess;
}

/* The shared-map occupancy-grid kernel can carry the LaserScanSensor too (both read the same staged map). */
bool casim_grid_can_fuse_laser(const casim_handle* h, const uint8_t* local_grid) {
  const int cells = h->dc.submap_w * h->dc.submap_h;
  return h->n_maps == 1 && h->grid2_map != nullptr && (h->dc.submap_h & 3) == 0 && (h->dc.submap_w & 3) == 0 &&
         (cells & 15) == 0 && (((uintptr_t)local_grid) & 15) == 0 && (long long)h->E * h->A < (1ll << 31);
}

cudaError_t casim_launch_grid(casim_handle* h, const float* state, uint8_t* local_grid, cudaStream_t s, float* fused_laser) {
  if (h->dc.submap_w > 64 || h->dc.submap_h > 64) return cudaErrorInvalidValue;
  GridArgs a;
  a.c = h->dc;
  a.state = reinterpret_cast<const float4*>(state);
  a.pose64 = h->pose64;
  a.n_agents = h->n_agents;
  a.map_planes = h->map_planes;
  a.grid_maps = h->grid_maps;
  a.n_maps = h->n_maps;
  a.out = local_grid;
  a.E = h->E;
  a.A = h->A;
  long long N = (long long)h->E * h->A;
  h->launches++;
  const int cells = h->dc.submap_w * h->dc.submap_h;
  /* fast path: rows are whole 4-cell words, every agent's grid is 16-byte aligned and a multiple of 16 bytes */
  const bool fast = (h->dc.submap_h & 3) == 0 && (cells & 15) == 0 && (((uintptr_t)local_grid) & 15) == 0;
  const bool tiles = fast && (h->dc.submap_w & 3) == 0 && N < (1ll << 31);
  if (h->grid_packed && (cells & 15) != 0) return cudaErrorInvalidValue; /* casim_set_grid_packed_output's size rule */
  cudaError_t err;
  if (tiles && h->n_maps == 1 && h->grid2_map) {
    /* shared map: persistent kernel, whole map staged into shared memory through TMA */
    if (!h->grid2_ready) {
      err = casim_grid2_encode_tmap(h);
      if (err != cudaSuccess) return err;
    }
    const bool ref_shape = h->dc.submap_h == 60 && h->dc.submap_w == 60; /* Config.SUBMAP_WIDTH / HEIGHT, config.py:83-84 */
    void (*kern)(const CUtensorMap, const Grid2Args) = ref_shape ? k_grid_tma<15, 15> : k_grid_tma<0, 0>;
    static bool attr_set[2] = {false, false};
    if (!attr_set[ref_shape]) {
      err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G2_SMEM_BYTES);
      if (err != cudaSuccess) return err;
      attr_set[ref_shape] = true;
    }
    Grid2Args b;
    b.c = h->dc;
    b.state = a.state;
    b.pose64 = h->pose64;
    b.n_agents = h->n_agents;
    b.map_planes = h->map_planes;
    b.out = local_grid;
    b.packed = h->grid_packed;
    b.laser = fused_laser;
    {
      int nbp = 1, lg = 5;
      while (nbp < h->dc.nb) nbp <<= 1;
      for (int t = nbp; t > 1; t >>= 1) lg--;
      b.lg2P = lg;
    }
    b.E = h->E;
    b.A = h->A;
    b.inv_A = (h->A >= 2 && (unsigned long long)N * (unsigned long long)h->A < (1ull << 32)) ? (unsigned)((1ull << 32) / (unsigned)h->A) + 1u : 0u;
    const long long want = (N + G2_WARPS - 1) / G2_WARPS;
    const int sms = h->num_sms > 0 ? h->num_sms : CASIM_NUM_SMS_FALLBACK;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(want < sms ? want : sms));
    cfg.blockDim = dim3(G2_THREADS);
    cfg.dynamicSmemBytes = G2_SMEM_BYTES;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = casim_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, *reinterpret_cast<const CUtensorMap*>(h->grid2_tmap), b);
  }
  if (fused_laser) return cudaErrorInvalidValue; /* the caller checks casim_grid_can_fuse_laser first */
  if (fast)
    err = casim_launch_pdl<GridArgs>(k_grid_gather, dim3((unsigned)((N + GRID_WARPS - 1) / GRID_WARPS)), dim3(32 * GRID_WARPS), 0, s, a);
  else
    err = casim_launch_pdl<GridArgs>(k_grid_generic, dim3((unsigned)N), dim3(256), 0, s, a);
  if (err != cudaSuccess || !h->grid_packed) return err;
  h->launches++;
  const size_t nbytes = (size_t)N * ((cells + 7) >> 3);
  k_grid_pack<<<(unsigned)((nbytes + 255) / 256), 256, 0, s>>>(local_grid, h->grid_packed, (size_t)N * cells, cells);
  return cudaGetLastError();
}
