/*
 * casim_flatten.cu -- the flattened observation wire format.
 *
 * Replaces MultiagentFlattenDictWrapper.observation (gym_collision_avoidance/envs/wrappers.py:37-45): for every
 * environment, for agent slot 0..A-1, for key in dict_keys: observation[agent][key].ravel(), concatenated into one
 * float32 row -- here for all E environments at once, out[E][A][S] with S = sum of the key widths, so that
 * out.view(E, A*S) is the array a VecEnv hands to the trainer (experiments/src/env_utils.py:7-39).  Absent agent slots
 * stay zero, like the zero-initialised self.observation of the reference (collision_avoidance_env.py:125-134).
 *
 * One thread per output float: the writes are fully coalesced, wide sources (other_agents_states, laserscan,
 * local_grid) are read coalesced, scalar sources are 4-byte gathers from the state planes (L2-resident).
 * Bound: HBM -- algorithmic bytes per agent = S x 4 written + the source bytes (4 per float, 1 per grid cell).
 */
#include "casim_device.cuh"

namespace {

#define FLAT_MAX_SEGS 16

struct FlatSeg {
  int begin, len; /* [begin, begin+len) inside an agent's row */
  int key;        /* enum casim_obs_key */
};

struct FlatArgs {
  const float4* state;
  casim_obs o;
  const int32_t* n_agents;
  float* out;
  FlatSeg seg[FLAT_MAX_SEGS];
  int n_segs, S, E, A, rows, nb, grid_cells;
};

__global__ void __launch_bounds__(256) k_flatten(const FlatArgs p) {
  casim_pdl_enter();
  const size_t N = (size_t)p.E * p.A;
  const size_t total = N * (size_t)p.S;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const size_t g = idx / (size_t)p.S;
    const int o = (int)(idx - g * (size_t)p.S);
    const int e = (int)(g / p.A), ai = (int)(g - (size_t)e * p.A);
    const int n = p.n_agents ? min(p.n_agents[e], p.A) : p.A;
    float v = 0.f;
    if (ai < n) {
      int k = 0;
#pragma unroll 1
      for (int s = 1; s < p.n_segs; s++)
        if (o >= p.seg[s].begin) k = s;
      const int j = o - p.seg[k].begin;
      const float* st = reinterpret_cast<const float*>(p.state);
#define WORD(w) st[((size_t)((w) >> 2) * N + g) * 4 + ((w)&3)]
      switch (p.seg[k].key) {
        case CASIM_KEY_DIST_TO_GOAL: v = WORD(CASIM_DIST_TO_GOAL); break;
        case CASIM_KEY_RADIUS: v = WORD(CASIM_RADIUS); break;
        case CASIM_KEY_HEADING_EGO_FRAME: v = WORD(CASIM_HEADING_EGO); break;
        case CASIM_KEY_HEADING_GLOBAL_FRAME: v = WORD(CASIM_HEADING); break;
        case CASIM_KEY_PREF_SPEED: v = WORD(CASIM_PREF_SPEED); break;
        case CASIM_KEY_POS_GLOBAL_FRAME: v = WORD(CASIM_PX + j); break;
        case CASIM_KEY_REL_GOAL: v = p.o.obs_misc[g * 4 + j]; break;
        case CASIM_KEY_NUM_OTHER_AGENTS: v = p.o.obs_misc[g * 4 + 2]; break;
        case CASIM_KEY_USE_PPO: v = p.o.obs_misc[g * 4 + 3]; break;
        case CASIM_KEY_OTHER_AGENT_STATES: /* config.py:171-177: row 0 of other_agents_states (sic) */
        case CASIM_KEY_OTHER_AGENTS_STATES: v = p.o.other_agents_states[g * (size_t)(p.rows * 10) + j]; break;
        case CASIM_KEY_LASERSCAN: v = p.o.laserscan[g * (size_t)p.nb + j]; break;
        case CASIM_KEY_LOCAL_GRID: v = p.o.local_grid[g * (size_t)p.grid_cells + j] ? 1.f : 0.f; break;
        default: break;
      }
#undef WORD
    }
    p.out[idx] = v;
  }
}

int key_width(const casim_handle* h, int key) {
  switch (key) {
    case CASIM_KEY_POS_GLOBAL_FRAME:
    case CASIM_KEY_REL_GOAL: return 2;
    case CASIM_KEY_OTHER_AGENT_STATES: return 10;
    case CASIM_KEY_OTHER_AGENTS_STATES: return h->dc.rows * 10;
    case CASIM_KEY_LASERSCAN: return h->dc.nb;
    case CASIM_KEY_LOCAL_GRID: return h->dc.submap_w * h->dc.submap_h;
    default: return (key >= 0 && key < CASIM_KEY_COUNT) ? 1 : -1;
  }
}

}  // namespace

int casim_flat_row_size_impl(const casim_handle* h, const int32_t* keys, int n_keys) {
  int S = 0;
  for (int i = 0; i < n_keys; i++) {
    int w = key_width(h, keys[i]);
    if (w < 0) return -1;
    S += w;
  }
  return S;
}

cudaError_t casim_launch_flatten(casim_handle* h, const float* state, const casim_obs* obs, const int32_t* keys,
                                 int n_keys, float* out, cudaStream_t s) {
  FlatArgs a;
  a.state = reinterpret_cast<const float4*>(state);
  a.o = *obs;
  a.n_agents = h->n_agents;
  a.out = out;
  a.n_segs = n_keys;
  int S = 0;
  for (int i = 0; i < n_keys; i++) {
    a.seg[i].begin = S;
    a.seg[i].len = key_width(h, keys[i]);
    a.seg[i].key = keys[i];
    S += a.seg[i].len;
  }
  a.S = S;
  a.E = h->E;
  a.A = h->A;
  a.rows = h->dc.rows;
  a.nb = h->dc.nb;
  a.grid_cells = h->dc.submap_w * h->dc.submap_h;
  const size_t total = (size_t)h->E * h->A * S;
  size_t blocks = (total + 255) / 256;
  const size_t cap = (size_t)h->num_sms * 16; /* grid-stride beyond 16 resident CTAs per SM */
  if (blocks > cap) blocks = cap;
  if (blocks == 0) return cudaSuccess;
  h->launches++;
  return casim_launch_pdl<FlatArgs>(k_flatten, dim3((unsigned)blocks), dim3(256), 0, s, a);
}
