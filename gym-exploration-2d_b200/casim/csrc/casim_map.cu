/*
 * casim_map.cu -- static map construction (once per reset, not on the per-step path).
 *
 *   k_raster      Map.get_occupancy_grid (Map.py:107-123) + world_coordinates_to_map_indices (Map.py:40-47):
 *                 axis-aligned rectangles (corner 1 = top-left, corner 3 = bottom-right) -> 300x300 bits.
 *   k_pack        a ready-made bool raster [n,300,300] -> bits.
 *   k_derive      the three OR-dilated planes used by the OccupancyGrid kernel (bilinear footprints
 *                 {S}, {S, right}, {S, down}, {S, right, down, diagonal}) and the optional uint8 copy.
 *   k_edt_cols / k_edt_rows   edfMap.update (information_models/edfMap.py:11-12): exact Euclidean distance
 *                 transform of the free space, kept as exact integer squared distances (the kernels take sqrt * cell size in
 *                 fp64 on the fly, which is what scipy.ndimage.distance_transform_edt * cellSize yields).
 */
#include "casim_device.cuh"

#include <mutex>

namespace {

__global__ void k_raster(DevCfg c, const float* rects, const int32_t* n_rects, int n_maps, int max_rects,
                         uint32_t* planes) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_maps * CASIM_PLANE_WORDS) return;
  int m = i / CASIM_PLANE_WORDS, w = i - m * CASIM_PLANE_WORDS;
  int row = w / CASIM_MAP_WORDS, wc = w - row * CASIM_MAP_WORDS;
  uint32_t bits = 0;
  int nrect = min(n_rects[m], max_rects);
  for (int k = 0; k < nrect; k++) {
    const float* r = rects + ((size_t)m * max_rects + k) * 4;
    int r0, c0, r1, c1;
    world_to_map_div(c, (double)r[0], (double)r[1], r0, c0);
    world_to_map_div(c, (double)r[2], (double)r[3], r1, c1);
    if (row < r0 || row > r1) continue;
    int lo = max(c0, wc * 32), hi = min(min(c1, wc * 32 + 31), CASIM_MAP_DIM - 1);
    for (int col = lo; col <= hi; col++) bits |= 1u << (col & 31);
  }
  planes[(size_t)m * CASIM_NUM_PLANES * CASIM_PLANE_WORDS + w] = bits;
}

__global__ void k_pack(const uint8_t* maps, int n_maps, uint32_t* planes) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_maps * CASIM_PLANE_WORDS) return;
  int m = i / CASIM_PLANE_WORDS, w = i - m * CASIM_PLANE_WORDS;
  int row = w / CASIM_MAP_WORDS, wc = w - row * CASIM_MAP_WORDS;
  const uint8_t* src = maps + ((size_t)m * CASIM_MAP_DIM + row) * CASIM_MAP_DIM;
  uint32_t bits = 0;
  for (int b = 0; b < 32; b++) {
    int col = wc * 32 + b;
    if (col < CASIM_MAP_DIM && src[col]) bits |= 1u << b;
  }
  planes[(size_t)m * CASIM_NUM_PLANES * CASIM_PLANE_WORDS + w] = bits;
}

__device__ __forceinline__ uint32_t plane_word(const uint32_t* p0, int row, int wc) {
  if (row < 0 || row >= CASIM_MAP_DIM || wc < 0 || wc >= CASIM_MAP_WORDS) return 0u;
  return p0[row * CASIM_MAP_WORDS + wc];
}

/* word of "S shifted so that bit c holds S(row, c+1)" */
__device__ __forceinline__ uint32_t right_word(const uint32_t* p0, int row, int wc) {
  return (plane_word(p0, row, wc) >> 1) | (plane_word(p0, row, wc + 1) << 31);
}

__global__ void k_derive(int n_maps, uint32_t* planes, uint8_t* map_out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_maps * CASIM_PLANE_WORDS) return;
  int m = i / CASIM_PLANE_WORDS, w = i - m * CASIM_PLANE_WORDS;
  int row = w / CASIM_MAP_WORDS, wc = w - row * CASIM_MAP_WORDS;
  uint32_t* base = planes + (size_t)m * CASIM_NUM_PLANES * CASIM_PLANE_WORDS;
  const uint32_t* p0 = base;
  uint32_t colmask = (wc == CASIM_MAP_WORDS - 1) ? ((1u << (CASIM_MAP_DIM - 32 * (CASIM_MAP_WORDS - 1))) - 1u) : 0xffffffffu;
  uint32_t s = plane_word(p0, row, wc);
  uint32_t r = right_word(p0, row, wc);
  uint32_t d = plane_word(p0, row + 1, wc);
  uint32_t dr = right_word(p0, row + 1, wc);
  base[1 * CASIM_PLANE_WORDS + w] = (s | r) & colmask;
  base[2 * CASIM_PLANE_WORDS + w] = (s | d) & colmask;
  base[3 * CASIM_PLANE_WORDS + w] = (s | r | d | dr) & colmask;
  if (map_out) {
    uint8_t* dst = map_out + ((size_t)m * CASIM_MAP_DIM + row) * CASIM_MAP_DIM;
    for (int b = 0; b < 32; b++) {
      int col = wc * 32 + b;
      if (col < CASIM_MAP_DIM) dst[col] = (s >> b) & 1u;
    }
  }
}

/* Word `pw` of padded row `pr` of plane 0 (S) of one map, for a padding of PAD cells on every side: padded column
 * 32 pw + b = map column 32 pw + b - PAD; cells outside the map are free. */
template <int PAD>
__device__ __forceinline__ uint32_t padded_s_word(const uint32_t* p0, int pr, int pw, int rows, int words) {
  const int row = pr - PAD;
  if (pr < 0 || pr >= rows || pw < 0 || pw >= words || row < 0 || row >= CASIM_MAP_DIM) return 0u;
  const int c0 = 32 * pw - PAD;
  const int wi = (c0 >= 0 ? c0 : c0 - 31) / 32, sh = c0 - 32 * wi; /* floor division */
  const uint32_t* src = p0 + row * CASIM_MAP_WORDS;
  const uint32_t lo = (wi >= 0 && wi < CASIM_MAP_WORDS) ? src[wi] : 0u;
  const uint32_t hi = (wi + 1 >= 0 && wi + 1 < CASIM_MAP_WORDS) ? src[wi + 1] : 0u;
  return sh ? __funnelshift_r(lo, hi, sh) : lo;
}

/* The four bilinear footprints {S}, {S, right}, {S, down}, {S, right, down, diagonal} of padded cell (pr, 32 pw + b),
 * derived IN PADDED SPACE: the cells just outside the map (column -1, row -1) see the map's first column / row through
 * their right / lower neighbour, which a padded copy of the unpadded footprint planes would lose. */
template <int PAD>
__device__ __forceinline__ uint32_t padded_footprint_word(const uint32_t* p0, int pl, int pr, int pw, int rows, int words) {
  const uint32_t s = padded_s_word<PAD>(p0, pr, pw, rows, words);
  if (pl == 0) return s;
  const uint32_t sn = padded_s_word<PAD>(p0, pr, pw + 1, rows, words);
  const uint32_t r = (s >> 1) | (sn << 31);
  if (pl == 1) return s | r;
  const uint32_t d = padded_s_word<PAD>(p0, pr + 1, pw, rows, words);
  if (pl == 2) return s | d;
  const uint32_t dn = padded_s_word<PAD>(p0, pr + 1, pw + 1, rows, words);
  return s | r | d | ((d >> 1) | (dn << 31));
}

/* OccupancyGrid gather path (per-env maps): the four footprint planes padded by 96 cells = 3 words */
__global__ void k_grid_pad(int n_maps, const uint32_t* planes, uint32_t* gmaps) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int per_map = CASIM_NUM_PLANES * CASIM_GPLANE_WORDS;
  if (i >= n_maps * per_map) return;
  int m = i / per_map, r = i - m * per_map;
  int pl = r / CASIM_GPLANE_WORDS, q = r - pl * CASIM_GPLANE_WORDS;
  int pr = q / CASIM_GWORDS, pw = q - pr * CASIM_GWORDS;
  const uint32_t* p0 = planes + (size_t)m * CASIM_NUM_PLANES * CASIM_PLANE_WORDS;
  gmaps[(size_t)m * CASIM_GMAP_WORDS + r] = padded_footprint_word<CASIM_GPAD>(p0, pl, pr, pw, CASIM_GROWS, CASIM_GWORDS);
}

/* 2-bit classification of every anchor (padded row pr, padded col pc): over the box rows [pr, pr+6) x cols
 * [pc, pc+6) of plane 0 (cells outside the padded array count as free): bit 0 = all free, bit 1 = all occupied.
 * One thread per output word (16 anchors). */
__global__ void k_grid_classify(int n_maps, uint32_t* gmaps) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int per_map = CASIM_GROWS * CASIM_GCLS_WORDS;
  if (i >= n_maps * per_map) return;
  int m = i / per_map, r = i - m * per_map;
  int pr = r / CASIM_GCLS_WORDS, cw = r - pr * CASIM_GCLS_WORDS;
  const uint32_t* p0 = gmaps + (size_t)m * CASIM_GMAP_WORDS; /* padded plane 0 */
  uint32_t out = 0;
  for (int a = 0; a < 16; a++) {
    int pc = cw * 16 + a;
    bool allz = true, allo = true;
    for (int dr = 0; dr < CASIM_GBOX; dr++) {
      int rr = pr + dr;
      uint32_t bits = 0;
      if (rr < CASIM_GROWS) {
        int w = pc >> 5, sh = pc & 31;
        uint32_t lo = p0[rr * CASIM_GWORDS + w];
        uint32_t hi = (w + 1 < CASIM_GWORDS) ? p0[rr * CASIM_GWORDS + w + 1] : 0u;
        bits = __funnelshift_r(lo, hi, sh) & ((1u << CASIM_GBOX) - 1u);
      }
      allz = allz && bits == 0u;
      allo = allo && bits == ((1u << CASIM_GBOX) - 1u);
    }
    out |= ((allz ? 1u : 0u) | (allo ? 2u : 0u)) << (2 * a);
  }
  gmaps[(size_t)m * CASIM_GMAP_WORDS + CASIM_NUM_PLANES * CASIM_GPLANE_WORDS + r] = out;
}

/* Shared-map layout of k_grid_tma: the four footprint planes of map 0, padded by CASIM_G2_PAD (= 32 + 10) cells */
__global__ void k_grid2_planes(const uint32_t* planes, uint32_t* g2) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= CASIM_NUM_PLANES * CASIM_G2_PLANE_WORDS) return;
  int pl = i / CASIM_G2_PLANE_WORDS, q = i - pl * CASIM_G2_PLANE_WORDS;
  int pr = q / CASIM_G2_WORDS, pw = q - pr * CASIM_G2_WORDS;
  g2[i] = padded_footprint_word<CASIM_G2_PAD>(planes, pl, pr, pw, CASIM_G2_DIM, CASIM_G2_WORDS);
}

/* One classification byte per even anchor (see casim_internal.h): the 8x8 box of padded plane 0 anchored at
 * (2 ay2, 2 ax2); cells beyond the padded array count as free. */
__global__ void k_grid2_classify(uint32_t* g2) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= CASIM_G2_CLS_BYTES) return;
  const int ay = 2 * (i / CASIM_G2_CLS_DIM), ax = 2 * (i % CASIM_G2_CLS_DIM);
  uint32_t rows[8];
  uint32_t all_or = 0, all_and = 0xffu, colpat = 0;
  bool rows_equal = true, rows_solid = true;
  for (int j = 0; j < 8; j++) {
    uint32_t bits = 0;
    const int rr = ay + j;
    if (rr < CASIM_G2_DIM) {
      const int w = ax >> 5, sh = ax & 31;
      uint32_t lo = g2[rr * CASIM_G2_WORDS + w];
      uint32_t hi = (w + 1 < CASIM_G2_WORDS) ? g2[rr * CASIM_G2_WORDS + w + 1] : 0u;
      bits = __funnelshift_r(lo, hi, sh) & 0xffu;
    }
    rows[j] = bits;
    all_or |= bits;
    all_and &= bits;
    rows_equal = rows_equal && bits == rows[0];
    rows_solid = rows_solid && (bits == 0u || bits == 0xffu);
    colpat |= (bits != 0u ? 1u : 0u) << j;
  }
  uint32_t code = 0;
  if (all_or == 0u)
    code = 1;
  else if (all_and == 0xffu)
    code = 2;
  else {
    const uint32_t pat = rows_equal ? rows[0] : (rows_solid ? colpat : 0x100u);
    const uint32_t base = rows_equal ? 3u : 5u;
    for (uint32_t pp = 1; pp < 8; pp++) {
      if (pat == ((0xffu << pp) & 0xffu)) code = base | (pp << 3);          /* columns / rows >= pp occupied */
      if (pat == ((1u << pp) - 1u)) code = (base + 1u) | (pp << 3);         /* columns / rows <  pp occupied */
    }
  }
  reinterpret_cast<uint8_t*>(g2 + CASIM_NUM_PLANES * CASIM_G2_PLANE_WORDS)[i] = (uint8_t)code;
}

/* pass 1: per column, distance (in cells) to the nearest occupied cell in that column (big = none) */
__global__ void k_edt_cols(int n_maps, const uint32_t* planes, int* gdist) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_maps * CASIM_MAP_DIM) return;
  int m = i / CASIM_MAP_DIM, col = i - m * CASIM_MAP_DIM;
  const uint32_t* p0 = planes + (size_t)m * CASIM_NUM_PLANES * CASIM_PLANE_WORDS;
  int* g = gdist + (size_t)m * CASIM_MAP_DIM * CASIM_MAP_DIM;
  const int BIG = 1 << 20;
  int d = BIG;
  for (int row = 0; row < CASIM_MAP_DIM; row++) {
    bool occ = (p0[row * CASIM_MAP_WORDS + (col >> 5)] >> (col & 31)) & 1u;
    d = occ ? 0 : min(d + 1, BIG);
    g[row * CASIM_MAP_DIM + col] = d;
  }
  d = BIG;
  for (int row = CASIM_MAP_DIM - 1; row >= 0; row--) {
    bool occ = (p0[row * CASIM_MAP_WORDS + (col >> 5)] >> (col & 31)) & 1u;
    d = occ ? 0 : min(d + 1, BIG);
    g[row * CASIM_MAP_DIM + col] = min(g[row * CASIM_MAP_DIM + col], d);
  }
}

/* pass 2: exact minimum over the row of dx^2 + g^2 (300 candidates per cell), sqrt, * cell */
__global__ void k_edt_rows(int n_maps, const int* gdist, int32_t* edf_d2, double* edf, double cell) {
  __shared__ int sg[CASIM_MAP_DIM];
  int m = blockIdx.x / CASIM_MAP_DIM, row = blockIdx.x - m * CASIM_MAP_DIM;
  const int* g = gdist + ((size_t)m * CASIM_MAP_DIM + row) * CASIM_MAP_DIM;
  for (int k = threadIdx.x; k < CASIM_MAP_DIM; k += blockDim.x) sg[k] = g[k];
  __syncthreads();
  const int BIG = 1 << 20;
  for (int col = threadIdx.x; col < CASIM_MAP_DIM; col += blockDim.x) {
    long long best = -1;
    for (int k = 0; k < CASIM_MAP_DIM; k++) {
      int gv = sg[k];
      if (gv >= BIG) continue;
      long long dx = col - k;
      long long v = dx * dx + (long long)gv * gv;
      if (best < 0 || v < best) best = v;
    }
    /* no obstacle anywhere: scipy returns the distance to a non-existent background; the exploration
     * kernels are only used with obstacles, keep a large finite value */
    size_t o = ((size_t)m * CASIM_MAP_DIM + row) * CASIM_MAP_DIM + col;
    const int32_t d2 = best < 0 ? (1 << 30) : (int32_t)best;
    edf_d2[o] = d2;
    edf[o] = sqrt((double)d2) * cell; /* edfMap.map (edfMap.py:11-12), the value the exploration kernels sample */
  }
}

}  // namespace

cudaError_t casim_launch_raster(casim_handle* h, const float* rects, const int32_t* n_rects, int n_maps,
                                int max_rects, cudaStream_t s) {
  int total = n_maps * CASIM_PLANE_WORDS;
  h->launches++;
  k_raster<<<(total + 255) / 256, 256, 0, s>>>(h->dc, rects, n_rects, n_maps, max_rects, h->map_planes);
  return cudaGetLastError();
}

cudaError_t casim_launch_pack_map(casim_handle* h, const uint8_t* static_map, int n_maps, cudaStream_t s) {
  int total = n_maps * CASIM_PLANE_WORDS;
  h->launches++;
  k_pack<<<(total + 255) / 256, 256, 0, s>>>(static_map, n_maps, h->map_planes);
  return cudaGetLastError();
}

cudaError_t casim_launch_map_derive(casim_handle* h, int n_maps, uint8_t* map_out, cudaStream_t s) {
  int total = n_maps * CASIM_PLANE_WORDS;
  h->launches += 3;
  k_derive<<<(total + 255) / 256, 256, 0, s>>>(n_maps, h->map_planes, map_out);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return err;
  {
    int tp = n_maps * CASIM_NUM_PLANES * CASIM_GPLANE_WORDS, tc = n_maps * CASIM_GROWS * CASIM_GCLS_WORDS;
    h->launches += 2;
    k_grid_pad<<<(tp + 255) / 256, 256, 0, s>>>(n_maps, h->map_planes, h->grid_maps);
    k_grid_classify<<<(tc + 127) / 128, 128, 0, s>>>(n_maps, h->grid_maps);
    err = cudaGetLastError();
    if (err != cudaSuccess) return err;
  }
  if (n_maps == 1 && h->grid2_map) { /* shared-map layout for the TMA-staged occupancy-grid kernel */
    h->launches += 2;
    k_grid2_planes<<<(CASIM_NUM_PLANES * CASIM_G2_PLANE_WORDS + 255) / 256, 256, 0, s>>>(h->map_planes, h->grid2_map);
    k_grid2_classify<<<(CASIM_G2_CLS_BYTES + 255) / 256, 256, 0, s>>>(h->grid2_map);
    err = cudaGetLastError();
    if (err != cudaSuccess) return err;
  }
  int* gdist = h->edf_scratch;
  k_edt_cols<<<(n_maps * CASIM_MAP_DIM + 127) / 128, 128, 0, s>>>(n_maps, h->map_planes, gdist);
  k_edt_rows<<<n_maps * CASIM_MAP_DIM, 128, 0, s>>>(n_maps, gdist, h->edf_d2, h->edf, h->dc.cell);
  return cudaGetLastError();
}

/* ------------------------------------------------------------------------------------------------------------------
 * casim_rasterize_noalloc -- the allocation-free variant of the map build (include/casim.h).  Same kernels as
 * casim_set_obstacles (k_raster -> k_derive -> k_edt_cols -> k_edt_rows), but every intermediate lives in ONE file-scope
 * device array instead of the handle's cudaMalloc'ed buffers:
 *
 *   per map (CASIM_NOALLOC_MAX_MAPS of them):  4 bit planes [4][300][10] u32 | column distances [300][300] i32 |
 *                                              squared distances [300][300] i32          = 768 000 B
 *
 * The array is held by every GPU that loads the library until the process exits.  Every word the kernels read is
 * written earlier in the same call (k_raster writes all of plane 0, k_derive all of planes 1-3, k_edt_cols every column
 * distance, k_edt_rows every squared distance), so nothing has to start at zero and the previous call's contents are
 * simply overwritten; two calls that overlap on the GPU (different streams) race on it. */
namespace {

constexpr size_t NA_PLANE_WORDS = (size_t)CASIM_NUM_PLANES * CASIM_PLANE_WORDS;        /* 12 000 */
constexpr size_t NA_CELLS = (size_t)CASIM_MAP_DIM * CASIM_MAP_DIM;                      /* 90 000 */
constexpr size_t NA_WORDS = CASIM_NOALLOC_MAX_MAPS * (NA_PLANE_WORDS + 2 * NA_CELLS);   /* 768 000 words = 3 072 000 B */

static __device__ __align__(256) uint32_t g_noalloc_scratch[NA_WORDS];

constexpr int NA_MAX_DEVICES = 64;
uint32_t* g_noalloc_ptr[NA_MAX_DEVICES]; /* device address of g_noalloc_scratch on each device (the symbol itself is a
                                            host-side stand-in and must never be handed to a kernel) */
std::mutex g_noalloc_mu;

/* once per device, with that device current */
cudaError_t noalloc_scratch(uint32_t** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= NA_MAX_DEVICES) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    uint32_t* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    if (!p) return cudaErrorInvalidSymbol;
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return cudaSuccess;
}

}  // namespace

extern "C" int casim_rasterize_noalloc(casim_handle* h, const float* rects, const int32_t* n_rects, int n_maps,
                                       int max_rects, uint8_t* map_out, double* edf_out, void* stream) {
  if (!h || !rects || !n_rects || (!map_out && !edf_out)) {
    casim_set_error("casim: casim_rasterize_noalloc: null argument (handle, rects, n_rects and one of map_out / edf_out)");
    return CASIM_EINVAL;
  }
  if (n_maps < 1 || n_maps > CASIM_NOALLOC_MAX_MAPS) {
    casim_set_error("casim: casim_rasterize_noalloc: n_maps = %d outside 1..CASIM_NOALLOC_MAX_MAPS (%d): the static "
                    "scratch holds no more; use casim_set_obstacles for larger batches", n_maps, CASIM_NOALLOC_MAX_MAPS);
    return CASIM_EINVAL;
  }
  if (max_rects < 0 || max_rects > CASIM_NOALLOC_MAX_RECTS) {
    casim_set_error("casim: casim_rasterize_noalloc: max_rects = %d outside 0..CASIM_NOALLOC_MAX_RECTS (%d)", max_rects,
                    CASIM_NOALLOC_MAX_RECTS);
    return CASIM_EINVAL;
  }
  uint32_t* scratch = nullptr;
  cudaError_t e = noalloc_scratch(&scratch);
  if (e != cudaSuccess) {
    casim_set_error("casim_rasterize_noalloc: cudaGetSymbolAddress(g_noalloc_scratch) failed: %s", cudaGetErrorString(e));
    return (int)e;
  }
  cudaStream_t s = (cudaStream_t)stream;
  uint32_t* planes = scratch;                                                    /* [n_maps][4][3000] */
  int* gdist = reinterpret_cast<int*>(scratch + CASIM_NOALLOC_MAX_MAPS * NA_PLANE_WORDS);   /* [n_maps][300][300] */
  int32_t* d2 = gdist + CASIM_NOALLOC_MAX_MAPS * NA_CELLS;                       /* [n_maps][300][300] */
  const int total = n_maps * CASIM_PLANE_WORDS;
  k_raster<<<(total + 255) / 256, 256, 0, s>>>(h->dc, rects, n_rects, n_maps, max_rects, planes);
  k_derive<<<(total + 255) / 256, 256, 0, s>>>(n_maps, planes, map_out);
  if (edf_out) {
    k_edt_cols<<<(n_maps * CASIM_MAP_DIM + 127) / 128, 128, 0, s>>>(n_maps, planes, gdist);
    k_edt_rows<<<n_maps * CASIM_MAP_DIM, 128, 0, s>>>(n_maps, gdist, d2, edf_out, h->dc.cell);
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    casim_set_error("casim_rasterize_noalloc: launch failed: %s", cudaGetErrorString(e));
    return (int)e;
  }
  h->launches += edf_out ? 4 : 2;
  return 0;
}
