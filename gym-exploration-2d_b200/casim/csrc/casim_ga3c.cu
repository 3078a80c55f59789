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
 * k_ga3c_tc: bf16 tcgen05 forward, fp32 accumulate in TMEM (see the block comment above the kernel).
 */
#include <cuda_bf16.h>

#include <cstring>
#include <vector>

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
  uint4* xb;               /* [tiles][10 slots][128 rows] 16 B: 7 normalised bf16 inputs + bf16 1.0, or NULL */
  const uint8_t* wimg;     /* bf16 weight image in UMMA smem order (tensor-core path) */
  float* policy_actions;   /* [N][2] */
  int32_t* action_idx;     /* [N] or NULL */
  float* probs;            /* [N][11] or NULL */
  float* nn_input;         /* [N][75] or NULL */
  int E, A;
};

/* network.py:125-128 normalisation vectors */
__constant__ float c_host_avg[4] = {0.0f, 0.0f, 1.0f, 0.5f};
__constant__ float c_host_std[4] = {5.0f, 3.14f, 1.0f, 1.0f};
__constant__ float c_oth_avg[7] = {0.0f, 0.0f, 0.0f, 0.0f, 0.5f, 0.0f, 1.0f};
__constant__ float c_oth_std[7] = {5.0f, 5.0f, 1.0f, 1.0f, 1.0f, 5.0f, 1.0f};
/* network.py:8-17 Actions, np.mgrid rows = start + k*step (filled by casim_ga3c_upload) */
__constant__ float c_actions[G_ACT][2];

#define T_M 128 /* agents per CTA of the tensor-core path = UMMA M */

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi); /* .x = lo (low half) */
  return *reinterpret_cast<uint32_t*>(&v);
}

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

/* One CTA per group of 32 consecutive agent slots.  fp64 is scarce on B200, so every fp64 instruction runs with
 * (nearly) all lanes useful: phase A = one lane per agent (reference frame), phase B = one lane per (agent, other)
 * pair (distance, rounded key, p_orthog), phase C = stable rank of sorted(key = (-round(dist, 2), p_orthog)) per pair
 * (integer compares, fp64 only on key ties), then the 76-float rows are assembled in shared memory and leave the
 * CTA as contiguous stores (x, nn_input) and 16-byte bf16 chunks (xb). */
#define GS_AG 32
#define GS_THREADS 128
__global__ void __launch_bounds__(GS_THREADS) k_ga3c_state(const Ga3cArgs p) {
  __shared__ double s_fr[GS_AG][6];                 /* px, py, prx, pry, rh, - */
  __shared__ double s_po[GS_AG][CASIM_MAX_AGENTS];
  __shared__ int s_kr[GS_AG][CASIM_MAX_AGENTS];
  __shared__ int s_m[GS_AG], s_n[GS_AG], s_ok[GS_AG];
  __shared__ __align__(16) float s_row[GS_AG][G_XS];
  const DevCfg& c = p.c;
  const size_t N = (size_t)p.E * p.A;
  const int tid = threadIdx.x, A = p.A;
  const size_t gb = (size_t)blockIdx.x * GS_AG;
  const int nag = (int)min((size_t)GS_AG, N - gb);
  for (int k = tid; k < GS_AG * G_XS; k += GS_THREADS) (&s_row[0][0])[k] = 0.f;
  if (tid < GS_AG) { /* phase A */
    const int a = tid;
    int ok = 0, n = 0;
    if (a < nag) {
      const size_t g = gb + a;
      const int e = (int)(g / A), i = (int)(g - (size_t)e * A);
      n = p.n_agents ? min(p.n_agents[e], A) : A;
      const int type = __float_as_int(p.state[5 * N + g].w);
      ok = (i < n && (type & 0xff) == CASIM_POLICY_GA3C_CADRL) ? 1 : 0;
      if (ok) {
        const float4 h0 = p.state[g], h2 = p.state[2 * N + g];
        const double px = h0.x, py = h0.y;
        /* agent.py:250-269 get_ref */
        double gx = (double)h2.x - px, gy = (double)h2.y - py;
        const double dg = sqrt(gx * gx + gy * gy);
        if (dg > 1e-8) {
          gx = gx / dg;
          gy = gy / dg;
        }
        s_fr[a][0] = px;
        s_fr[a][1] = py;
        s_fr[a][2] = gx;
        s_fr[a][3] = gy;
        s_fr[a][4] = (double)h2.z;
      }
    }
    s_ok[a] = ok;
    s_n[a] = n;
    s_m[a] = 0;
  }
  __syncthreads();
  const int npairs = nag * A;
  for (int idx = tid; idx < npairs; idx += GS_THREADS) { /* phase B */
    const int a = idx / A, j = idx - a * A;
    int kr = INT_MIN;
    double po = 0.0;
    if (s_ok[a]) {
      const size_t g = gb + a;
      const int e = (int)(g / A), i = (int)(g - (size_t)e * A);
      if (j < s_n[a] && j != i) {
        const size_t o = (size_t)e * A + j;
        const float4 o0 = p.state[o], o2 = p.state[2 * N + o];
        const double rx = (double)o0.x - s_fr[a][0], ry = (double)o0.y - s_fr[a][1];
        const double dc = sqrt(rx * rx + ry * ry);
        if (!(dc > c.sensing_horizon)) {
          kr = (int)py_round2_hundredths(dc - s_fr[a][4] - (double)o2.z);
          po = rx * (-s_fr[a][3]) + ry * s_fr[a][2];
          atomicAdd(&s_m[a], 1);
        }
      }
    }
    s_kr[a][j] = kr;
    s_po[a][j] = po;
  }
  __syncthreads();
  for (int idx = tid; idx < npairs; idx += GS_THREADS) { /* phase C */
    const int a = idx / A, j = idx - a * A;
    const int kr = s_kr[a][j];
    if (kr == INT_MIN) continue;
    const double po = s_po[a][j];
    const int n = s_n[a], m = s_m[a];
    const int K = min(min(m, c.max_obs), G_SLOTS);
    int rank = 0; /* stable sort: number of candidates that sort before me */
    for (int k = 0; k < n; k++) {
      const int krk = s_kr[a][k];
      bool before = krk > kr;
      if (krk == kr && k != j) {
        const double pok = s_po[a][k];
        before = pok < po || (pok == po && k < j);
      }
      rank += (krk != INT_MIN) && before;
    }
    const int slot = rank - (m - K); /* keep the last K of the ascending order */
    if (slot < 0) continue;
    const size_t g = gb + a;
    const size_t o = (g / A) * A + j;
    const float4 o0 = p.state[o], o1 = p.state[N + o], o2 = p.state[2 * N + o];
    const double prx = s_fr[a][2], pry = s_fr[a][3], rh = s_fr[a][4];
    const double rx = (double)o0.x - s_fr[a][0], ry = (double)o0.y - s_fr[a][1];
    const double vx = o1.x, vy = o1.y;
    float* w = &s_row[a][1 + G_HOST + G_OTHER * slot];
    w[0] = (float)(rx * prx + ry * pry);
    w[1] = (float)po;
    w[2] = (float)(vx * prx + vy * pry);
    w[3] = (float)(vx * (-pry) + vy * prx);
    w[4] = o2.z;
    w[5] = (float)(rh + (double)o2.z);
    w[6] = (float)(sqrt(rx * rx + ry * ry) - rh - (double)o2.z);
  }
  if (tid < nag) {
    const int a = tid;
    if (s_ok[a]) {
      const size_t g = gb + a;
      const float4 h2 = p.state[2 * N + g], h3 = p.state[3 * N + g], h5 = p.state[5 * N + g];
      s_row[a][0] = (float)min(min(s_m[a], c.max_obs), G_SLOTS);
      s_row[a][1] = h3.z; /* dist_to_goal */
      s_row[a][2] = h5.x; /* heading_ego_frame */
      s_row[a][3] = h2.w; /* pref_speed */
      s_row[a][4] = h2.z; /* radius */
    } else {
      s_row[a][0] = -1.f; /* marks "no GA3C agent in this slot" for the forward kernels */
    }
  }
  __syncthreads();
  {
    float* x = p.x + gb * G_XS; /* rows are contiguous: G_XS floats per agent */
    const float* src = &s_row[0][0];
    for (int k = tid; k < nag * G_XS; k += GS_THREADS) x[k] = src[k];
    if (p.nn_input) {
      float* o = p.nn_input + gb * G_IN;
      for (int k = tid; k < nag * G_IN; k += GS_THREADS) {
        const int a = k / G_IN, q = k - a * G_IN;
        if (s_ok[a]) o[k] = s_row[a][q];
      }
    }
  }
  if (p.xb) { /* tensor-core path: per (slot, agent) the K-chunk [x_0..x_6 normalised, 1.0] of the LSTM A operand */
    for (int idx = tid; idx < G_SLOTS * GS_AG; idx += GS_THREADS) {
      const int sl = idx / GS_AG, a = idx - sl * GS_AG;
      if (a >= nag) continue;
      const size_t g = gb + a;
      uint4 q = make_uint4(0u, 0u, 0u, 0u);
      if (s_ok[a]) {
        const float* v = &s_row[a][1 + G_HOST + G_OTHER * sl];
        float nv[8];
#pragma unroll
        for (int k = 0; k < G_OTHER; k++) nv[k] = (v[k] - c_oth_avg[k]) / c_oth_std[k];
        nv[7] = 1.0f;
        q = make_uint4(pack_bf16(nv[0], nv[1]), pack_bf16(nv[2], nv[3]), pack_bf16(nv[4], nv[5]), pack_bf16(nv[6], nv[7]));
      }
      p.xb[(g / T_M) * (size_t)(G_SLOTS * T_M) + (size_t)sl * T_M + (g % T_M)] = q;
    }
  }
}

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

/* ------------------------------------------------------------------------------------------------------------
 * k_ga3c_tc: the same network on the 5th-generation tensor cores (tcgen05.mma, kind::f16, bf16 x bf16 -> fp32).
 *
 * CTA = 128 agents = one UMMA M=128 tile, two CTAs per SM (256 TMEM columns and ~107 KB of shared memory each), so
 * one CTA's MMAs overlap the other's epilogue.  Roles: warps 0-3 = epilogue (thread r owns agent row r = TMEM lane r),
 * warp 4 lane 0 = MMA issuer, warp 5 lane 0 = weight producer (cp.async.bulk of pre-formatted 8 KB slabs, 5-slot
 * ring).  Every GEMM is D[128 x N] = A[128 x K] * W[K x N] with D in TMEM columns 0..N-1:
 *   LSTM step : K = 80 = [h(64) | x_t(7) | 1 | 1 | 0..], N = 256 = gates i,j,f,o x 64; the two "1" columns carry the
 *               bias as a bf16 hi + lo pair; the 40 KB of LSTM weights stay in the ring for all <= 10 steps;
 *   layer1    : K = 80 = [h(64) | host(4) | 1 | 1 | 0..], N = 256;
 *   layer2/fc1: K = 256, N = 256, fp32 bias added in the epilogue;   logits_p: K = 256, N = 16 (11 used).
 * Operands sit in shared memory in the canonical K-major SWIZZLE_NONE layout: 8-row x 16-byte core matrices,
 * element (r, k) of a tile with R rows at (k/8)*R*16 + r*16 + (k%8)*2, i.e. LBO (next K chunk) = R*16 bytes,
 * SBO (next 8-row group) = 128 bytes.  The epilogue writes the next GEMM's A operand (h_t, ReLU activations) straight
 * into that layout as bf16: row r's 16-byte pieces are contiguous across a warp, so the stores are conflict-free.
 * Gate non-linearities use tanh.approx (sigmoid(x) = 0.5 tanh(x/2) + 0.5): one MUFU op each; the epilogue is
 * MUFU-bound (320 per agent per LSTM step), not tensor-bound.
 * ---------------------------------------------------------------------------------------------------------- */
#define T_THREADS 192
#define T_KL 80
#define T_SLAB 8192
#define T_RING 5
#define T_ACH 2048                 /* bytes of one 8-wide K chunk of the A tile (128 rows x 16 B) */
#define T_A_BYTES (32 * T_ACH)     /* K = 256 */
#define T_NSLAB 43                 /* 5 (lstm) + 5 (layer1) + 16 (layer2) + 16 (fc1) + 1 (logits_p, 32 chunks x 256 B) */
#define T_IMG_BYTES (T_NSLAB * T_SLAB)
#define T_NBIAS (G_FC + G_FC + 16) /* fp32 biases of layer2, fc1, logits_p appended to the image */
#define T_DYN_SMEM (T_A_BYTES + T_RING * T_SLAB)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
/* Bounded wait: a protocol bug must trap, not hang the device. */
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = 0;
  for (uint32_t it = 0;; it++) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (it == 64) t0 = clock64();
    if (it > 64 && clock64() - t0 > 4000000000ll) __trap();
  }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

/* shared-memory matrix descriptor, SWIZZLE_NONE, K-major (cute::UMMA::SmemDescriptor bit layout): start >> 4 in
 * [0,14), LBO >> 4 in [16,30), SBO >> 4 in [32,46), version 1 in [46,48), layout type 0 in [61,64) */
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         (1ull << 46);
}
/* instruction descriptor (cute::UMMA::InstrDescriptor): D fp32 [4,6) = 1, A bf16 [7,10) = 1, B bf16 [10,13) = 1,
 * both K-major (bits 15, 16 = 0), N >> 3 in [17,23), M >> 4 in [24,29) */
__device__ __forceinline__ uint32_t umma_idesc(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
/* thread t of warp w reads TMEM lane 32*(w%4) + t, 8 / 16 consecutive fp32 columns */
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float v[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
#pragma unroll
  for (int k = 0; k < 8; k++) v[k] = __uint_as_float(r[k]);
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float v[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, "
      "[%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
#pragma unroll
  for (int k = 0; k < 16; k++) v[k] = __uint_as_float(r[k]);
}
__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }

__global__ void __launch_bounds__(T_THREADS, 2) k_ga3c_tc(const Ga3cArgs p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(16) float s_bias[T_NBIAS];
  __shared__ __align__(8) uint64_t s_bar[2 * T_RING + 2];
  __shared__ uint32_t s_tmem;
  __shared__ int s_len[T_M];
  __shared__ int s_maxlen;

  const size_t N = (size_t)p.E * p.A;
  const size_t g0 = (size_t)blockIdx.x * T_M;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t sA = smem_u32(smem), sW = sA + T_A_BYTES;
  const uint32_t bar_full = smem_u32(&s_bar[0]), bar_empty = smem_u32(&s_bar[T_RING]);
  const uint32_t bar_a = smem_u32(&s_bar[2 * T_RING]), bar_d = smem_u32(&s_bar[2 * T_RING + 1]);

  if (tid == 0) {
    for (int k = 0; k < T_RING; k++) {
      mbar_init(bar_full + 8 * k, 1);
      mbar_init(bar_empty + 8 * k, 1);
    }
    mbar_init(bar_a, T_M);
    mbar_init(bar_d, 1);
    s_maxlen = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) { /* TMEM: 256 fp32 columns x 128 lanes; this warp also frees them */
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(256u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid < T_M) {
    const size_t g = g0 + tid;
    s_len[tid] = g < N ? min((int)p.x[g * G_XS], G_SLOTS) : -1; /* x[0] = sequence length, -1 = no GA3C agent here */
  }
  for (int k = tid; k < T_NBIAS; k += T_THREADS) s_bias[k] = __ldg(reinterpret_cast<const float*>(p.wimg + T_IMG_BYTES) + k);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid < T_M && s_len[tid] > 0) atomicMax(&s_maxlen, s_len[tid]);
  __syncthreads();
  const int maxlen = s_maxlen;
  const uint32_t tmem = s_tmem;

  if (warp == 5) {
    /* ===== weight producer: slab q of the image -> ring slot q % 5 ===== */
    if (lane == 0) {
      for (int q = 0; q < T_NSLAB; q++) {
        const int slot = q % T_RING, use = q / T_RING;
        if (use > 0) mbar_wait(bar_empty + 8 * slot, (use - 1) & 1);
        mbar_expect_tx(bar_full + 8 * slot, T_SLAB);
        bulk_g2s(sW + slot * T_SLAB, p.wimg + (size_t)q * T_SLAB, T_SLAB, bar_full + 8 * slot);
      }
    }
  } else if (warp == 4) {
    /* ===== MMA issuer ===== */
    if (lane == 0) {
      const uint32_t id256 = umma_idesc(T_M, 256), id16 = umma_idesc(T_M, 16);
      uint32_t a_phase = 0;
      for (int s = 0; s < T_RING; s++) mbar_wait(bar_full + 8 * s, 0); /* LSTM weights resident */
      for (int t = 0; t < maxlen; t++) {
        mbar_wait(bar_a, a_phase);
        a_phase ^= 1;
        tc_fence_after();
#pragma unroll
        for (int s = 0; s < T_KL / 16; s++)
          umma_bf16(tmem, umma_desc(sA + s * 2 * T_ACH, T_ACH, 128), umma_desc(sW + s * T_SLAB, 4096, 128), id256, s > 0);
        if (t == maxlen - 1)
          for (int s = 0; s < T_RING; s++) tc_commit(bar_empty + 8 * s);
        tc_commit(bar_d);
      }
      if (maxlen == 0)
        for (int s = 0; s < T_RING; s++) mbar_arrive(bar_empty + 8 * s);
      int q = T_RING;
      for (int layer = 0; layer < 3; layer++) { /* layer1 (5 slabs), layer2 (16), fullyconnected1 (16) */
        const int nslab = layer == 0 ? T_KL / 16 : G_FC / 16;
        mbar_wait(bar_a, a_phase);
        a_phase ^= 1;
        for (int s = 0; s < nslab; s++, q++) {
          const int slot = q % T_RING;
          mbar_wait(bar_full + 8 * slot, (q / T_RING) & 1);
          tc_fence_after();
          umma_bf16(tmem, umma_desc(sA + s * 2 * T_ACH, T_ACH, 128), umma_desc(sW + slot * T_SLAB, 4096, 128), id256, s > 0);
          tc_commit(bar_empty + 8 * slot);
        }
        tc_commit(bar_d);
      }
      { /* logits_p: one 8 KB slab = 32 K-chunks of 16 rows x 16 B */
        const int slot = q % T_RING;
        mbar_wait(bar_a, a_phase);
        a_phase ^= 1;
        mbar_wait(bar_full + 8 * slot, (q / T_RING) & 1);
        tc_fence_after();
        for (int s = 0; s < G_FC / 16; s++)
          umma_bf16(tmem, umma_desc(sA + s * 2 * T_ACH, T_ACH, 128), umma_desc(sW + slot * T_SLAB + s * 512, 256, 128), id16,
                    s > 0);
        tc_commit(bar_d);
      }
    }
  } else {
    /* ===== epilogue: thread r <-> agent row r <-> TMEM lane r ===== */
    const int r = tid;
    const size_t g = g0 + r;
    const int len = s_len[r];
    const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
    const uint32_t arow = sA + r * 16;
    const uint4* xb = p.xb + (size_t)blockIdx.x * (G_SLOTS * T_M) + r;
    const uint32_t one_bf = 0x3f80u; /* bf16 1.0 in the low half */
    auto sts16 = [](uint32_t addr, uint4 v) {
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    };
    /* layer1's K chunks 8, 9: [host0..3 normalised, 1, 1, 0, 0], zeros (network.py:83-84, 89-90) */
    auto write_host_chunks = [&]() {
      float hv[4] = {0.f, 0.f, 0.f, 0.f};
      if (g < N && len >= 0) {
#pragma unroll
        for (int k = 0; k < G_HOST; k++) hv[k] = (p.x[g * G_XS + 1 + k] - c_host_avg[k]) / c_host_std[k];
      }
      sts16(arow + 8 * T_ACH, make_uint4(pack_bf16(hv[0], hv[1]), pack_bf16(hv[2], hv[3]), 0x3f803f80u, 0u));
      sts16(arow + 9 * T_ACH, make_uint4(0u, 0u, 0u, 0u));
    };
    /* A for LSTM step 0: h = 0 | x_0, 1 | 1, 0.. */
#pragma unroll
    for (int c8 = 0; c8 < 8; c8++) sts16(arow + c8 * T_ACH, make_uint4(0u, 0u, 0u, 0u));
    if (maxlen > 0) {
      sts16(arow + 8 * T_ACH, xb[0]);
      sts16(arow + 9 * T_ACH, make_uint4(one_bf, 0u, 0u, 0u));
    } else {
      write_host_chunks();
    }
    fence_async_smem();
    mbar_arrive(bar_a);

    uint32_t d_phase = 0;
    float cst[G_H];
#pragma unroll
    for (int u = 0; u < G_H; u++) cst[u] = 0.f;
    for (int t = 0; t < maxlen; t++) {
      uint4 nx = make_uint4(0u, 0u, 0u, 0u);
      if (t + 1 < maxlen) nx = xb[(t + 1) * T_M];
      mbar_wait(bar_d, d_phase);
      d_phase ^= 1;
      tc_fence_after();
      const bool active = t < len; /* dynamic_rnn(sequence_length): state copied through beyond the last valid step */
#pragma unroll
      for (int uc = 0; uc < G_H / 8; uc++) {
        float gi[8], gj[8], gf[8], go[8];
        tmem_ld8(trow + uc * 8, gi);
        tmem_ld8(trow + 64 + uc * 8, gj);
        tmem_ld8(trow + 128 + uc * 8, gf);
        tmem_ld8(trow + 192 + uc * 8, go);
        tc_wait_ld();
        float hn[8];
#pragma unroll
        for (int u = 0; u < 8; u++) { /* TF1 LSTMCell: gates i, j, f, o; forget_bias 1.0 */
          const float cn = fmaf(sigmoid_fast(gf[u] + 1.0f), cst[uc * 8 + u], sigmoid_fast(gi[u]) * tanh_fast(gj[u]));
          hn[u] = sigmoid_fast(go[u]) * tanh_fast(cn);
          if (active) cst[uc * 8 + u] = cn;
        }
        if (active)
          sts16(arow + uc * T_ACH, make_uint4(pack_bf16(hn[0], hn[1]), pack_bf16(hn[2], hn[3]), pack_bf16(hn[4], hn[5]),
                                              pack_bf16(hn[6], hn[7])));
      }
      if (t + 1 < maxlen)
        sts16(arow + 8 * T_ACH, nx);
      else
        write_host_chunks();
      fence_async_smem();
      tc_fence_before();
      mbar_arrive(bar_a);
    }
    /* layer1, layer2, fullyconnected1: ReLU(D + bias) -> bf16 -> A of the next GEMM (in place: the MMAs that read
     * A have completed when bar_d fires) */
    for (int layer = 0; layer < 3; layer++) {
      mbar_wait(bar_d, d_phase);
      d_phase ^= 1;
      tc_fence_after();
      const float* bias = s_bias + (layer > 0 ? (layer - 1) * G_FC : 0); /* layer1's bias rides in the GEMM */
#pragma unroll 2
      for (int cc = 0; cc < G_FC / 16; cc++) {
        float v[16];
        tmem_ld16(trow + cc * 16, v);
        tc_wait_ld();
        if (layer > 0) {
#pragma unroll
          for (int k = 0; k < 16; k++) v[k] += bias[cc * 16 + k];
        }
#pragma unroll
        for (int k = 0; k < 16; k++) v[k] = fmaxf(v[k], 0.f);
        sts16(arow + (2 * cc) * T_ACH,
              make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7])));
        sts16(arow + (2 * cc + 1) * T_ACH, make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]),
                                                      pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15])));
      }
      fence_async_smem();
      tc_fence_before();
      mbar_arrive(bar_a);
    }
    /* logits_p -> softmax (+1e-4 floor) -> argmax -> action (network.py:45-51, GA3CCADRLPolicy.py:38-42) */
    mbar_wait(bar_d, d_phase);
    tc_fence_after();
    float lg[16];
    tmem_ld16(trow, lg);
    tc_wait_ld();
    tc_fence_before();
    if (g < N && len >= 0) {
      const int flags = __float_as_int(p.state[5 * N + g].z);
      if (flags & CASIM_F_IS_DONE) { /* collision_avoidance_env.py:298-300 */
        p.policy_actions[2 * g] = 0.f;
        p.policy_actions[2 * g + 1] = 0.f;
        if (p.action_idx) p.action_idx[g] = -1;
      } else {
        const float* pb = s_bias + 2 * G_FC;
        float mx = -INFINITY;
#pragma unroll
        for (int k = 0; k < G_ACT; k++) {
          lg[k] += pb[k];
          mx = fmaxf(mx, lg[k]);
        }
        float pr[G_ACT], sum = 0.f;
#pragma unroll
        for (int k = 0; k < G_ACT; k++) {
          pr[k] = expf(lg[k] - mx);
          sum += pr[k];
        }
        int best = 0;
        float pbest = -1.f;
#pragma unroll
        for (int k = 0; k < G_ACT; k++) {
          pr[k] = (pr[k] / sum + 1e-4f) / (1.0f + 1e-4f * G_ACT);
          if (pr[k] > pbest) {
            pbest = pr[k];
            best = k;
          }
        }
        if (p.probs) {
#pragma unroll
          for (int k = 0; k < G_ACT; k++) p.probs[g * G_ACT + k] = pr[k];
        }
        if (p.action_idx) p.action_idx[g] = best;
        const float pref = p.state[2 * N + g].w;
        p.policy_actions[2 * g] = (float)((double)pref * (double)c_actions[best][0]);
        p.policy_actions[2 * g + 1] = c_actions[best][1];
      }
    } else if (g < N && p.action_idx) {
      p.action_idx[g] = -1;
    }
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
  }
}

}  // namespace

static uint16_t f32_to_bf16(float f) { /* round to nearest even, like cvt.rn.bf16.f32 */
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static float bf16_to_f32(uint16_t b) {
  uint32_t u = (uint32_t)b << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

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
  /* tensor-core path: bf16 image of every layer in the shared-memory order k_ga3c_tc feeds to tcgen05.mma
   * (K-major, SWIZZLE_NONE: element (n, k) of a layer with R output rows at (k/8)*R*16 + n*16 + (k%8)*2) */
  std::vector<uint8_t> img(T_IMG_BYTES + T_NBIAS * sizeof(float), 0);
  auto put = [&](size_t layer_off, int rows, int k, int n, float v) {
    uint16_t b = f32_to_bf16(v);
    memcpy(&img[layer_off + (size_t)(k / 8) * rows * 16 + (size_t)n * 16 + (k % 8) * 2], &b, 2);
  };
  auto put_bias = [&](size_t layer_off, int k, int n, float v) { /* bf16 hi + lo pair against two "1" columns */
    uint16_t hi = f32_to_bf16(v);
    put(layer_off, G_FC, k, n, v);
    put(layer_off, G_FC, k + 1, n, v - bf16_to_f32(hi));
  };
  const float* w = weights_host;
  for (int n = 0; n < 4 * G_H; n++) { /* LSTM: A columns [h(64) | x(7) | 1 | 1]; TF kernel rows are [x(7); h(64)] */
    for (int k = 0; k < G_H; k++) put(0, G_FC, k, n, w[W_LSTM_K + (G_OTHER + k) * 4 * G_H + n]);
    for (int k = 0; k < G_OTHER; k++) put(0, G_FC, G_H + k, n, w[W_LSTM_K + k * 4 * G_H + n]);
    put_bias(0, G_H + G_OTHER, n, w[W_LSTM_B + n]);
  }
  const size_t o1 = 5 * T_SLAB, o2 = 10 * T_SLAB, o3 = 26 * T_SLAB, o4 = 42 * T_SLAB;
  for (int n = 0; n < G_FC; n++) { /* layer1: A columns [h(64) | host(4) | 1 | 1]; TF kernel rows are [host(4); h(64)] */
    for (int k = 0; k < G_H; k++) put(o1, G_FC, k, n, w[W_L1_K + (G_HOST + k) * G_FC + n]);
    for (int k = 0; k < G_HOST; k++) put(o1, G_FC, G_H + k, n, w[W_L1_K + k * G_FC + n]);
    put_bias(o1, G_H + G_HOST, n, w[W_L1_B + n]);
    for (int k = 0; k < G_FC; k++) {
      put(o2, G_FC, k, n, w[W_L2_K + k * G_FC + n]);
      put(o3, G_FC, k, n, w[W_FC_K + k * G_FC + n]);
    }
  }
  for (int n = 0; n < G_ACT; n++)
    for (int k = 0; k < G_FC; k++) put(o4, 16, k, n, w[W_P_K + k * G_ACT + n]);
  float* bias = reinterpret_cast<float*>(&img[T_IMG_BYTES]);
  memcpy(bias, w + W_L2_B, G_FC * sizeof(float));
  memcpy(bias + G_FC, w + W_FC_B, G_FC * sizeof(float));
  memcpy(bias + 2 * G_FC, w + W_P_B, G_ACT * sizeof(float));
  if (!h->ga3c_wb) {
    e = cudaMalloc(&h->ga3c_wb, img.size());
    if (e != cudaSuccess) return e;
  }
  if (!h->ga3c_xb) { /* allocated here, not in the launcher: the forward must be capturable into a CUDA graph */
    const size_t ntiles = ((size_t)h->E * h->A + T_M - 1) / T_M;
    const size_t bytes = ntiles * G_SLOTS * T_M * sizeof(uint4);
    e = cudaMalloc(&h->ga3c_xb, bytes);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(h->ga3c_xb, 0, bytes, s); /* rows past N in the last tile stay zero */
    if (e != cudaSuccess) return e;
  }
  if (!h->ga3c_x) {
    e = cudaMalloc((void**)&h->ga3c_x, (size_t)h->E * h->A * G_XS * sizeof(float));
    if (e != cudaSuccess) return e;
  }
  e = cudaFuncSetAttribute(k_ga3c_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((size_t)G_FC * G_TILE * sizeof(float)));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_ga3c_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_DYN_SMEM);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_ga3c_tc, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
  if (e != cudaSuccess) return e;
  e = cudaMemcpyAsync(h->ga3c_wb, img.data(), img.size(), cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(s); /* weights_host / table / img may go away after return */
}

cudaError_t casim_launch_ga3c(casim_handle* h, const float* state, float* policy_actions, int32_t* action_idx,
                              float* probs, float* nn_input, int use_tc, cudaStream_t s) {
  const size_t N = (size_t)h->E * h->A;
  if (!h->ga3c_x) {
    cudaError_t e = cudaMalloc((void**)&h->ga3c_x, N * G_XS * sizeof(float));
    if (e != cudaSuccess) return e;
  }
  const unsigned ntiles = (unsigned)((N + T_M - 1) / T_M);
  Ga3cArgs a;
  a.c = h->dc;
  a.state = reinterpret_cast<const float4*>(state);
  a.n_agents = h->n_agents;
  a.w = h->ga3c_w;
  a.x = h->ga3c_x;
  a.xb = use_tc ? reinterpret_cast<uint4*>(h->ga3c_xb) : nullptr;
  a.wimg = reinterpret_cast<const uint8_t*>(h->ga3c_wb);
  a.policy_actions = policy_actions;
  a.action_idx = action_idx;
  a.probs = probs;
  a.nn_input = nn_input;
  a.E = h->E;
  a.A = h->A;
  h->launches += 2;
  k_ga3c_state<<<(unsigned)((N + GS_AG - 1) / GS_AG), GS_THREADS, 0, s>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  if (use_tc)
    k_ga3c_tc<<<ntiles, T_THREADS, T_DYN_SMEM, s>>>(a);
  else
    k_ga3c_fwd<<<(unsigned)((N + G_TILE - 1) / G_TILE), G_THREADS, (size_t)G_FC * G_TILE * sizeof(float), s>>>(a);
  return cudaGetLastError();
}
