/*
 * casim_ga3c.cu -- GA3C-CADRL policy: GA3CCADRLPolicy.find_next_action for every GA3C agent
 * (citations relative to gym_collision_avoidance/envs/policies/ of mlodel/gym-exploration-2d):
 *   GA3CCADRLPolicy.py:45-106    agents_to_ga3c_cadrl_state   -> k_ga3c_state
 *   GA3C_CADRL/network.py:65-98  normalise, LSTM(64) over <= 10 other agents, layer1, layer2
 *   GA3C_CADRL/network.py:45-51  fullyconnected1, logits_p, (softmax + 1e-4) / (1 + 11e-4)
 *   GA3CCADRLPolicy.py:38-42     argmax -> Actions table (network.py:8-17) -> [pref_speed * a0, a1]
 *
 * k_ga3c_state: one thread per agent.  The slot ORDER is an integer result that must match the reference
 * bit for bit, so the sort key (-round(dist, 2), p_orthog, list index) is evaluated in fp64 from the fp32 state and
 * Python's round(x, 2) (correctly rounded decimal, ties to even) is reproduced exactly with an error-free
 * product (fma) instead of a decimal conversion.  Values are then stored fp32 (the network's input dtype).
 *
 * k_ga3c_fwd<false>: fp32 SIMT forward, one CTA per tile of 64 agents, register-tiled (8 agents x 8 outputs per
 * thread), activations transposed in shared memory, weights streamed through L1/L2 (170 k parameters, shared by
 * every CTA).  This is the path compared against the fp32 oracle (1e-5).
 * k_ga3c_mma: bf16 tensor-core forward (weights converted once at upload), fp32 accumulate.
 */
#include <cuda_bf16.h>

#include "casim_device.cuh"

namespace {

#define G_IN CASIM_GA3C_INPUT /* 75 */
#define G_XS 76               /* row stride of the nn_input scratch */
#define G_SLOTS 10
#define G_OTHER 7
#define G_HOST 4
#define G_H 64
#define G_FC 256
#define G_ACT CASIM_GA3C_NUM_ACTIONS
#define G_TILE 64             /* agents per CTA */
#define G_THREADS 256

/* parameter blob offsets (include/casim.h) */
#define W_LSTM_K 0
#define W_LSTM_B (W_LSTM_K + (G_OTHER + G_H) * 4 * G_H)
#define W_L1_K (W_LSTM_B + 4 * G_H)
#define W_L1_B (W_L1_K + (G_HOST + G_H) * G_FC)
#define W_L2_K (W_L1_B + G_FC)
#define W_L2_B (W_L2_K + G_FC * G_FC)
#define W_FC_K (W_L2_B + G_FC)
#define W_FC_B (W_FC_K + G_FC * G_FC)
#define W_P_K (W_FC_B + G_FC)
#define W_P_B (W_P_K + G_FC * G_ACT)

struct Ga3cArgs {
  DevCfg c;
  const float4* state;
  const int32_t* n_agents;
  const float* w;          /* fp32 parameter blob */
  float* x;                /* [N][G_XS] scratch: the 75-float network input, fp32 */
  float* policy_actions;   /* [N][2] */
  int32_t* action_idx;     /* [N] or NULL */
  float* probs;            /* [N][11] or NULL */
  float* nn_input;         /* [N][75] or NULL */
  int E, A;
};

/* Python round(x, 2), returned as the integer number of hundredths.  float.__round__ rounds the EXACT binary
 * value to 2 decimals, ties to even.  200*|x| = p + e exactly (e from fma), so |x| is compared with the
 * midpoint (2k+1)/200 without any rounding. */
__device__ __forceinline__ long long py_round2_hundredths(double x) {
  double a = fabs(x);
  if (!(a < 1.0e13)) return 0; /* not reachable for distances in a 30 m map */
  double kf = floor(a * 100.0);
  double p = a * 200.0;
  double e = fma(a, 200.0, -p);
  double d = p - 2.0 * kf;      /* exact */
  double s = (d - 1.0) + e;     /* sign of 200a - (2k+1), exact in sign and zero */
  long long k = (long long)kf;
  if (s > 0.0 || (s == 0.0 && (k & 1))) k += 1;
  return x < 0.0 ? -k : k;
}

__global__ void __launch_bounds__(128) k_ga3c_state(const Ga3cArgs p) {
  const DevCfg& c = p.c;
  const size_t N = (size_t)p.E * p.A;
  const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= N) return;
  const int e = (int)(g / p.A), i = (int)(g - (size_t)e * p.A);
  const int n = p.n_agents ? min(p.n_agents[e], p.A) : p.A;
  float* x = p.x + g * G_XS;
  const int type = __float_as_int(p.state[5 * N + g].w);
  if (i >= n || (type & 0xff) != CASIM_POLICY_GA3C_CADRL) {
    x[0] = -1.f; /* marks "no GA3C agent in this slot" for the forward kernel */
    return;
  }
  const size_t base = (size_t)e * p.A;
  const float4 h0 = p.state[g], h2 = p.state[2 * N + g], h3 = p.state[3 * N + g], h5 = p.state[5 * N + g];
  const double px = h0.x, py = h0.y, rh = h2.z;
  /* agent.py:250-269 get_ref */
  double gx = (double)h2.x - px, gy = (double)h2.y - py;
  double dg = sqrt(gx * gx + gy * gy);
  double prx = gx, pry = gy;
  if (dg > 1e-8) {
    prx = gx / dg;
    pry = gy / dg;
  }
  const double orx = -pry, ory = prx;
  int kr[CASIM_MAX_AGENTS];
  double po[CASIM_MAX_AGENTS];
  int m = 0;
  for (int j = 0; j < n; j++) {
    kr[j] = INT_MIN; /* not a candidate */
    po[j] = 0.0;
    if (j == i) continue;
    const float4 o0 = p.state[base + j], o2 = p.state[2 * N + base + j];
    double rx = (double)o0.x - px, ry = (double)o0.y - py;
    double dc = sqrt(rx * rx + ry * ry);
    if (dc > c.sensing_horizon) continue;
    double d2o = dc - rh - (double)o2.z;
    kr[j] = (int)py_round2_hundredths(d2o);
    po[j] = rx * orx + ry * ory;
    m++;
  }
  int K = min(min(m, c.max_obs), G_SLOTS);
  for (int k = 1; k < G_IN; k++) x[k] = 0.f;
  x[0] = (float)K;
  x[1] = h3.z;  /* dist_to_goal */
  x[2] = h5.x;  /* heading_ego_frame */
  x[3] = h2.w;  /* pref_speed */
  x[4] = h2.z;  /* radius */
  /* sorted(key = (-round(dist, 2), p_orthog)) is stable: rank = number of candidates that sort before me */
  for (int j = 0; j < n; j++) {
    if (kr[j] == INT_MIN) continue;
    int rank = 0;
    for (int k = 0; k < n; k++) {
      if (k == j || kr[k] == INT_MIN) continue;
      bool before = (kr[k] > kr[j]) || (kr[k] == kr[j] && (po[k] < po[j] || (po[k] == po[j] && k < j)));
      rank += before;
    }
    int slot = rank - (m - K); /* keep the last K of the ascending order */
    if (slot < 0) continue;
    const float4 o0 = p.state[base + j], o1 = p.state[N + base + j], o2 = p.state[2 * N + base + j];
    double rx = (double)o0.x - px, ry = (double)o0.y - py;
    double vx = o1.x, vy = o1.y;
    float* w = x + 1 + G_HOST + G_OTHER * slot;
    w[0] = (float)(rx * prx + ry * pry);
    w[1] = (float)(rx * orx + ry * ory);
    w[2] = (float)(vx * prx + vy * pry);
    w[3] = (float)(vx * orx + vy * ory);
    w[4] = o2.z;
    w[5] = (float)(rh + (double)o2.z);
    w[6] = (float)(sqrt(rx * rx + ry * ry) - rh - (double)o2.z);
  }
  if (p.nn_input) {
    float* o = p.nn_input + g * G_IN;
    for (int k = 0; k < G_IN; k++) o[k] = x[k];
  }
}

/* network.py:125-128 normalisation vectors */
__constant__ float c_host_avg[4] = {0.0f, 0.0f, 1.0f, 0.5f};
__constant__ float c_host_std[4] = {5.0f, 3.14f, 1.0f, 1.0f};
__constant__ float c_oth_avg[7] = {0.0f, 0.0f, 0.0f, 0.0f, 0.5f, 0.0f, 1.0f};
__constant__ float c_oth_std[7] = {5.0f, 5.0f, 1.0f, 1.0f, 1.0f, 5.0f, 1.0f};
/* network.py:8-17 Actions, np.mgrid rows = start + k*step (filled by casim_ga3c_upload) */
__constant__ float c_actions[G_ACT][2];

__device__ __forceinline__ float sigmoid_f(float v) { return 1.0f / (1.0f + expf(-v)); }

/* acc[a][o] += sum_k act[k][ag0 + a] * W[k][col(o)] for this thread's 8 agents x 8 columns.
 * LSTM = true : columns {g*64 + tx*2 + {0,1}}: all four gates of units tx*2, tx*2+1 (row length 256);
 * LSTM = false: columns tx*8 .. tx*8+7. */
template <bool LSTM>
__device__ __forceinline__ void gemm_tile(const float* __restrict__ W, const float* __restrict__ bias, int K,
                                          const float* act /* smem [K][G_TILE] */, int tx, int ty, float acc[8][8]) {
  float b[8];
  if (LSTM) {
#pragma unroll
    for (int g = 0; g < 4; g++) {
      float2 v = __ldg(reinterpret_cast<const float2*>(bias + g * G_H + tx * 2));
      b[2 * g] = v.x;
      b[2 * g + 1] = v.y;
    }
  } else {
    float4 v0 = __ldg(reinterpret_cast<const float4*>(bias + tx * 8)), v1 = __ldg(reinterpret_cast<const float4*>(bias + tx * 8 + 4));
    b[0] = v0.x; b[1] = v0.y; b[2] = v0.z; b[3] = v0.w; b[4] = v1.x; b[5] = v1.y; b[6] = v1.z; b[7] = v1.w;
  }
#pragma unroll
  for (int a = 0; a < 8; a++)
#pragma unroll
    for (int o = 0; o < 8; o++) acc[a][o] = 0.f;
#pragma unroll 4
  for (int k = 0; k < K; k++) {
    float w[8];
    const float* wr = W + (size_t)k * G_FC;
    if (LSTM) {
#pragma unroll
      for (int g = 0; g < 4; g++) {
        float2 v = __ldg(reinterpret_cast<const float2*>(wr + g * G_H + tx * 2));
        w[2 * g] = v.x;
        w[2 * g + 1] = v.y;
      }
    } else {
      float4 v0 = __ldg(reinterpret_cast<const float4*>(wr + tx * 8)), v1 = __ldg(reinterpret_cast<const float4*>(wr + tx * 8 + 4));
      w[0] = v0.x; w[1] = v0.y; w[2] = v0.z; w[3] = v0.w; w[4] = v1.x; w[5] = v1.y; w[6] = v1.z; w[7] = v1.w;
    }
    const float4 a0 = *reinterpret_cast<const float4*>(act + k * G_TILE + ty * 8);
    const float4 a1 = *reinterpret_cast<const float4*>(act + k * G_TILE + ty * 8 + 4);
    const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
    for (int a = 0; a < 8; a++)
#pragma unroll
      for (int o = 0; o < 8; o++) acc[a][o] = fmaf(av[a], w[o], acc[a][o]);
  }
#pragma unroll
  for (int a = 0; a < 8; a++)
#pragma unroll
    for (int o = 0; o < 8; o++) acc[a][o] += b[o];
}

/* fp32 forward: CTA = 64 agents, 256 threads (tx = lane: output columns, ty = warp: 8 agents) */
__global__ void __launch_bounds__(G_THREADS) k_ga3c_fwd(const Ga3cArgs p) {
  extern __shared__ __align__(16) float s_act[]; /* [256][G_TILE] transposed activations */
  __shared__ float s_logit[G_TILE][12];
  __shared__ int s_len[G_TILE];
  const size_t N = (size_t)p.E * p.A;
  const size_t g0 = (size_t)blockIdx.x * G_TILE;
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  const float* w = p.w;

  /* sequence length per agent (x[0], un-normalised: network.py:85); -1 = slot without a GA3C agent */
  if (tid < G_TILE) {
    size_t g = g0 + tid;
    float x0 = g < N ? p.x[g * G_XS] : -1.f;
    s_len[tid] = (int)x0;
  }
  /* rows 0..6 = x_t (first step), rows 7..70 = h = 0 */
  for (int idx = tid; idx < (G_OTHER + G_H) * G_TILE; idx += G_THREADS) {
    int k = idx / G_TILE, a = idx - k * G_TILE;
    float v = 0.f;
    size_t g = g0 + a;
    if (k < G_OTHER && g < N) v = (p.x[g * G_XS + 1 + G_HOST + k] - c_oth_avg[k]) / c_oth_std[k];
    s_act[idx] = v;
  }
  __syncthreads();
  int maxlen = 0;
  for (int a = 0; a < G_TILE; a++) maxlen = max(maxlen, s_len[a]);
  maxlen = min(maxlen, G_SLOTS);

  float acc[8][8];
  float cst[8][2]; /* LSTM cell state of my 8 agents x 2 units */
#pragma unroll
  for (int a = 0; a < 8; a++) cst[a][0] = cst[a][1] = 0.f;
  for (int t = 0; t < maxlen; t++) {
    gemm_tile<true>(w + W_LSTM_K, w + W_LSTM_B, G_OTHER + G_H, s_act, tx, ty, acc);
    __syncthreads(); /* everyone has read x_t and h */
#pragma unroll
    for (int a = 0; a < 8; a++) {
      const int ag = ty * 8 + a;
      if (t < s_len[ag]) { /* dynamic_rnn(sequence_length): state copied through beyond the last valid step */
#pragma unroll
        for (int u = 0; u < 2; u++) { /* TF1 LSTMCell: i, j, f, o; forget_bias 1.0 */
          float gi = acc[a][u], gj = acc[a][2 + u], gf = acc[a][4 + u], go = acc[a][6 + u];
          float cn = sigmoid_f(gf + 1.0f) * cst[a][u] + sigmoid_f(gi) * tanhf(gj);
          cst[a][u] = cn;
          s_act[(G_OTHER + tx * 2 + u) * G_TILE + ag] = sigmoid_f(go) * tanhf(cn);
        }
      }
    }
    /* next step's inputs */
    if (t + 1 < maxlen)
      for (int idx = tid; idx < G_OTHER * G_TILE; idx += G_THREADS) {
        int k = idx / G_TILE, a = idx - k * G_TILE;
        size_t g = g0 + a;
        float v = 0.f;
        if (g < N && t + 1 < s_len[a]) v = (p.x[g * G_XS + 1 + G_HOST + G_OTHER * (t + 1) + k] - c_oth_avg[k]) / c_oth_std[k];
        s_act[idx] = v;
      }
    __syncthreads();
  }
  /* layer1 input = concat(host4, h64): host goes to rows 3..6, h already sits in rows 7..70 */
  if (tid < G_HOST * G_TILE) {
    int k = tid / G_TILE, a = tid - k * G_TILE;
    size_t g = g0 + a;
    s_act[(3 + k) * G_TILE + a] = g < N ? (p.x[g * G_XS + 1 + k] - c_host_avg[k]) / c_host_std[k] : 0.f;
  }
  __syncthreads();
  const float* Wk[3] = {w + W_L1_K, w + W_L2_K, w + W_FC_K};
  const float* Wb[3] = {w + W_L1_B, w + W_L2_B, w + W_FC_B};
#pragma unroll 1
  for (int layer = 0; layer < 3; layer++) {
    if (layer == 0)
      gemm_tile<false>(Wk[0], Wb[0], G_HOST + G_H, s_act + 3 * G_TILE, tx, ty, acc);
    else
      gemm_tile<false>(Wk[layer], Wb[layer], G_FC, s_act, tx, ty, acc);
    __syncthreads();
#pragma unroll
    for (int a = 0; a < 8; a++)
#pragma unroll
      for (int o = 0; o < 8; o++) s_act[(tx * 8 + o) * G_TILE + ty * 8 + a] = fmaxf(acc[a][o], 0.f);
    __syncthreads();
  }
  /* logits_p: 64 agents x 11 outputs, K = 256 */
  {
    const int a = tid >> 2, q = tid & 3;
    float l[3] = {0.f, 0.f, 0.f};
    for (int k = 0; k < G_FC; k++) {
      float v = s_act[k * G_TILE + a];
#pragma unroll
      for (int mm = 0; mm < 3; mm++) {
        int o = q + 4 * mm;
        if (o < G_ACT) l[mm] = fmaf(v, __ldg(w + W_P_K + k * G_ACT + o), l[mm]);
      }
    }
#pragma unroll
    for (int mm = 0; mm < 3; mm++) {
      int o = q + 4 * mm;
      if (o < G_ACT) s_logit[a][o] = l[mm] + __ldg(w + W_P_B + o);
    }
  }
  __syncthreads();
  if (tid < G_TILE) {
    const size_t g = g0 + tid;
    if (g < N && s_len[tid] >= 0) {
      const int flags = __float_as_int(p.state[5 * N + g].z);
      if (flags & CASIM_F_IS_DONE) { /* collision_avoidance_env.py:298-300 */
        p.policy_actions[2 * g] = 0.f;
        p.policy_actions[2 * g + 1] = 0.f;
        if (p.action_idx) p.action_idx[g] = -1;
      } else {
        float mx = s_logit[tid][0];
        for (int k = 1; k < G_ACT; k++) mx = fmaxf(mx, s_logit[tid][k]);
        float pr[G_ACT], sum = 0.f;
        for (int k = 0; k < G_ACT; k++) {
          pr[k] = expf(s_logit[tid][k] - mx);
          sum += pr[k];
        }
        int best = 0;
        for (int k = 0; k < G_ACT; k++) {
          pr[k] = (pr[k] / sum + 1e-4f) / (1.0f + 1e-4f * G_ACT);
          if (pr[k] > pr[best]) best = k;
        }
        if (p.probs)
          for (int k = 0; k < G_ACT; k++) p.probs[g * G_ACT + k] = pr[k];
        if (p.action_idx) p.action_idx[g] = best;
        const float pref = p.state[2 * N + g].w;
        p.policy_actions[2 * g] = (float)((double)pref * (double)c_actions[best][0]); /* GA3CCADRLPolicy.py:40-42 */
        p.policy_actions[2 * g + 1] = c_actions[best][1];
      }
    } else if (g < N && p.action_idx) {
      p.action_idx[g] = -1;
    }
  }
}

}  // namespace

cudaError_t casim_ga3c_upload(casim_handle* h, const float* weights_host, int n_params, cudaStream_t s) {
  cudaError_t e;
  if (!h->ga3c_w) {
    e = cudaMalloc((void**)&h->ga3c_w, (size_t)n_params * sizeof(float));
    if (e != cudaSuccess) return e;
  }
  e = cudaMemcpyAsync(h->ga3c_w, weights_host, (size_t)n_params * sizeof(float), cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess) return e;
  /* network.py:8-17: np.mgrid[start:stop:step] -> start + k*step in fp64, stored fp32 (all_actions is fp32) */
  float table[G_ACT][2];
  const double PI = 3.14159265358979323846;
  int r = 0;
  for (int k = 0; k < 5; k++, r++) { table[r][0] = 1.0f; table[r][1] = (float)(-PI / 6 + k * (PI / 12)); }
  for (int k = 0; k < 3; k++, r++) { table[r][0] = 0.5f; table[r][1] = (float)(-PI / 6 + k * (PI / 6)); }
  for (int k = 0; k < 3; k++, r++) { table[r][0] = 0.0f; table[r][1] = (float)(-PI / 6 + k * (PI / 6)); }
  e = cudaMemcpyToSymbolAsync(c_actions, table, sizeof(table), 0, cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(s); /* weights_host / table may go away after return */
}

cudaError_t casim_launch_ga3c(casim_handle* h, const float* state, float* policy_actions, int32_t* action_idx,
                              float* probs, float* nn_input, int use_tc, cudaStream_t s) {
  const size_t N = (size_t)h->E * h->A;
  if (!h->ga3c_x) {
    cudaError_t e = cudaMalloc((void**)&h->ga3c_x, N * G_XS * sizeof(float));
    if (e != cudaSuccess) return e;
  }
  Ga3cArgs a;
  a.c = h->dc;
  a.state = reinterpret_cast<const float4*>(state);
  a.n_agents = h->n_agents;
  a.w = h->ga3c_w;
  a.x = h->ga3c_x;
  a.policy_actions = policy_actions;
  a.action_idx = action_idx;
  a.probs = probs;
  a.nn_input = nn_input;
  a.E = h->E;
  a.A = h->A;
  h->launches += 2;
  k_ga3c_state<<<(unsigned)((N + 127) / 128), 128, 0, s>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  (void)use_tc;
  const size_t smem = (size_t)G_FC * G_TILE * sizeof(float);
  static bool attr_set = false;
  if (!attr_set) {
    e = cudaFuncSetAttribute(k_ga3c_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  k_ga3c_fwd<<<(unsigned)((N + G_TILE - 1) / G_TILE), G_THREADS, smem, s>>>(a);
  return cudaGetLastError();
}
