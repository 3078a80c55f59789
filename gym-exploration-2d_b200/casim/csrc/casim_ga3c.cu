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
 * k_ga3c_tc: tcgen05 forward on 16-bit operands (fp16; bf16 with -DU_FP16=0), fp32 accumulate in TMEM (see the block comment above the kernel).
 */
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <cstdlib>
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
  const float4* state0;          /* episode-start state kept by casim_reset, or NULL */
  const uint8_t* prev_game_over; /* [E] or NULL (no auto_reset) */
  const int32_t* n_agents;
  const float* w;          /* fp32 parameter blob */
  float* x;                /* [N][G_XS] scratch: the 75-float network input, fp32 */
  uint4* xb;               /* [tiles][10 slots][128 rows] 16 B: 7 normalised 16-bit inputs + 1.0, or NULL */
  const uint8_t* wimg;     /* 16-bit weight image in UMMA smem order (tensor-core path) */
  unsigned long long* trace; /* tuning aid (CASIM_GA3C_TRACE): [CTA][2 roles][64] globaltimer stamps, else NULL */
  float* policy_actions;   /* [N][2] */
  int32_t* action_idx;     /* [N] or NULL */
  int32_t* fix_list;       /* [N] agents whose tensor-core result is too close to call (top-2 margin < G_FIX_MARGIN), or NULL */
  int32_t* fix_count;      /* [1] length of fix_list: zeroed by k_ga3c_state, appended by k_ga3c_tc, read by the fp32 fix-up */
  int fix_mode;            /* k_ga3c_fwd: 1 = re-evaluate the agents of fix_list instead of all agents */
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

/* auto_reset: an environment whose episode ended last step restarts inside the coming casim_step from the episode-start
 * state kept by casim_reset (k_core, casim_step.cu); the policy is queried on that fresh state, flags cleared, exactly as
 * env.reset() followed by env.step() would do.  g = any element of the environment. */
__device__ __forceinline__ const float4* ga3c_src(const Ga3cArgs& p, size_t g) {
  if (p.prev_game_over != nullptr && p.prev_game_over[(unsigned)g / (unsigned)p.A]) return p.state0;
  return p.state;
}

/* "tensor cores, exact decisions": the tensor-core forward's probabilities were measured within 0.0055 of the fp32 forward's
 * with fp16 operands (0.039 with bf16; trained IROS18 weights, 163 044 agent-steps of the formation scenario,
 * profiles/r2z_ga3c_agreement_fp16.json; every disagreement of the argmax had an fp32 top-2 margin < 0.005), so an argmax
 * can only flip when that margin is below ~0.011.  Agents whose margin is below G_FIX_MARGIN (0.02 for fp16 operands, 0.10
 * for bf16: ~3.6 x / 2.5 x the measured error) are queued and re-evaluated by the fp32 kernel, which then overwrites their
 * action / probabilities: the action index equals the fp32 forward's for every agent. */
#ifndef G_FIX_AG
#define G_FIX_AG 2 /* agents per warp of the fix-up tile: 8 * G_FIX_AG agents per CTA (4: same kernel time, slower step) */
#endif
#ifndef U_FP16
#define U_FP16 1
#endif
#define G_FIX_MARGIN (U_FP16 ? 0.02f : 0.10f)

#define U_NA 144 /* agents per CTA of the tensor-core path */
#define U_XB_SLOTS 11

/* Operand format of the tensor-core forward.  tcgen05 kind::f16 takes fp16 or bf16 operands at the same rate; every operand
 * of this network is far inside fp16's range (normalised inputs, |h| < 1, ReLU activations of a few units, weights < 10), so
 * the three extra mantissa bits of fp16 are free accuracy: max |dp| against the fp32 forward 0.039 (bf16) -> see DESIGN.md 4.6.
 * U_FP16 = 0 restores bf16 operands ("op16" in the helper names = whichever 16-bit format is selected). */
#ifndef U_FP16
#define U_FP16 1
#endif
#define U_ONE16 (U_FP16 ? 0x3c00u : 0x3f80u) /* 1.0 in the operand format */
#define U_MAX16 65504.0f                       /* largest finite fp16: activations are clamped, never overflow to inf */
__device__ __forceinline__ uint32_t pack_op16(float lo, float hi) {
#if U_FP16
  __half2 v = __floats2half2_rn(fminf(fmaxf(lo, -U_MAX16), U_MAX16), fminf(fmaxf(hi, -U_MAX16), U_MAX16)); /* .x = lo (low half) */
#else
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi); /* .x = lo (low half) */
#endif
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
 * CTA as contiguous stores (x, nn_input) and 16-byte chunks of 16-bit operands (xb). */
#define GS_AG 32
#ifndef GS_THREADS
#define GS_THREADS 256 /* C5 sweep: 64: 108.6 us per step, 128: 98.5, 192: 97.7, 256: 96.8 (profiles/r1k_sweep_gs.txt) */
#endif
__global__ void __launch_bounds__(GS_THREADS) k_ga3c_state(const Ga3cArgs p) {
  __shared__ double s_fr[GS_AG][6];                 /* px, py, prx, pry, rh, - */
  __shared__ double s_po[GS_AG][CASIM_MAX_AGENTS];
  __shared__ int s_kr[GS_AG][CASIM_MAX_AGENTS];
  __shared__ int s_m[GS_AG], s_n[GS_AG], s_ok[GS_AG];
  __shared__ const float4* s_src[GS_AG];            /* the state this agent's ENVIRONMENT is read from (ga3c_src: state, or the
                                                       episode-start state when the environment restarts inside the coming step) */
  __shared__ __align__(16) float s_row[GS_AG][G_XS];
  const DevCfg& c = p.c;
  const size_t N = (size_t)p.E * p.A;
  const int tid = threadIdx.x, A = p.A;
  const size_t gb = (size_t)blockIdx.x * GS_AG;
  const int nag = (int)min((size_t)GS_AG, N - gb);
  for (int k = tid; k < GS_AG * G_XS; k += GS_THREADS) (&s_row[0][0])[k] = 0.f;
  if (blockIdx.x == 0 && tid == 0 && p.fix_count != nullptr) *p.fix_count = 0; /* the forward kernels follow in stream order */
  if (tid < GS_AG) { /* phase A */
    const int a = tid;
    int ok = 0, n = 0;
    if (a < nag) {
      const size_t g = gb + a;
      const int e = (int)(g / A), i = (int)(g - (size_t)e * A);
      n = p.n_agents ? min(p.n_agents[e], A) : A;
      const float4* src = ga3c_src(p, g);
      s_src[a] = src;
      const int type = __float_as_int(src[5 * N + g].w);
      ok = (i < n && (type & 0xff) == CASIM_POLICY_GA3C_CADRL) ? 1 : 0;
      if (ok) {
        const float4 h0 = src[g], h2 = src[2 * N + g];
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
        const float4* src = s_src[a]; /* o is an agent of the same environment */
        const float4 o0 = src[o], o2 = src[2 * N + o];
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
    const float4* src = s_src[a];
    const float4 o0 = src[o], o1 = src[N + o], o2 = src[2 * N + o];
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
      const float4* src = s_src[a];
      const float4 h2 = src[2 * N + g], h3 = src[3 * N + g], h5 = src[5 * N + g];
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
  if (p.xb) { /* tensor-core path: plane sl < 10 = per agent the K chunk [x_0..x_6 normalised, 1.0] of LSTM step sl;
                 plane 10 = header {host inputs normalised (network.py:83-84), 1, 1, sequence length or -1} */
    const size_t npad = (N + U_NA - 1) / U_NA * U_NA;
    for (int idx = tid; idx < U_XB_SLOTS * GS_AG; idx += GS_THREADS) {
      const int sl = idx / GS_AG, a = idx - sl * GS_AG;
      if (a >= nag) continue;
      uint4 q = make_uint4(0u, 0u, 0u, sl == G_SLOTS ? 0xffffffffu : 0u);
      if (s_ok[a]) {
        if (sl < G_SLOTS) {
          const float* v = &s_row[a][1 + G_HOST + G_OTHER * sl];
          float nv[8];
#pragma unroll
          for (int k = 0; k < G_OTHER; k++) nv[k] = (v[k] - c_oth_avg[k]) / c_oth_std[k];
          nv[7] = 1.0f;
          q = make_uint4(pack_op16(nv[0], nv[1]), pack_op16(nv[2], nv[3]), pack_op16(nv[4], nv[5]), pack_op16(nv[6], nv[7]));
        } else {
          float hv[4];
#pragma unroll
          for (int k = 0; k < G_HOST; k++) hv[k] = (s_row[a][1 + k] - c_host_avg[k]) / c_host_std[k];
          q = make_uint4(pack_op16(hv[0], hv[1]), pack_op16(hv[2], hv[3]), U_ONE16 | (U_ONE16 << 16), (uint32_t)(int)s_row[a][0]);
        }
      }
      p.xb[(size_t)sl * npad + gb + a] = q;
    }
  }
}

__device__ __forceinline__ float sigmoid_f(float v) { return 1.0f / (1.0f + expf(-v)); }

/* ---- weights staged through shared memory (cp.async), so that a tile's serial k loop never waits on L2 ---- */
#define G_WKC 16                               /* rows per streamed chunk of a dense layer */
#define G_WST 4                                /* chunks in the ring: three in flight behind the one in use */
#define G_WBUF ((G_OTHER + G_H) * G_FC)        /* floats: the whole LSTM kernel (71 x 256), or the dense ring (4 x 16 x 256) */

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
/* rows x 256 floats, global -> shared, all G_THREADS threads; the caller commits / waits */
__device__ __forceinline__ void stage_rows(float* dst, const float* src, int rows, int tid) {
  for (int i = tid; i < rows * (G_FC / 4); i += G_THREADS) cp_async16(dst + 4 * i, src + 4 * i);
}

/* acc[a][o] = bias + sum_k act[k][ty * AG + a] * W[k][col(o)] for this thread's AG agents x 8 columns, k ascending (an
 * agent's result does not depend on AG or on the staging).
 * LSTM step: columns {g*64 + tx*2 + {0,1}}: all four gates of units tx*2, tx*2+1; W = the kernel resident in shared memory. */
template <int AG>
__device__ __forceinline__ void lstm_tile(const float* sw /* smem [71][256] */, const float* __restrict__ bias,
                                          const float* act /* smem [K][8 * AG] */, int tx, int ty, float acc[AG][8]) {
  constexpr int TILE = 8 * AG;
#pragma unroll
  for (int a = 0; a < AG; a++)
#pragma unroll
    for (int o = 0; o < 8; o++) acc[a][o] = 0.f;
#pragma unroll(AG == 8 ? 4 : 8)
  for (int k = 0; k < G_OTHER + G_H; k++) {
    float w[8];
#pragma unroll
    for (int g = 0; g < 4; g++) {
      const float2 v = *reinterpret_cast<const float2*>(sw + k * G_FC + g * G_H + tx * 2);
      w[2 * g] = v.x;
      w[2 * g + 1] = v.y;
    }
    float av[AG];
    if (AG == 8) {
      const float4 a0 = *reinterpret_cast<const float4*>(act + k * TILE + ty * 8);
      const float4 a1 = *reinterpret_cast<const float4*>(act + k * TILE + ty * 8 + 4);
      av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w;
      av[AG - 4] = a1.x; av[AG - 3] = a1.y; av[AG - 2] = a1.z; av[AG - 1] = a1.w;
    } else if (AG == 4) {
      const float4 a0 = *reinterpret_cast<const float4*>(act + k * TILE + ty * 4);
      av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[AG - 1] = a0.w;
    } else {
#pragma unroll
      for (int a = 0; a < AG; a++) av[a] = act[k * TILE + ty * AG + a];
    }
#pragma unroll
    for (int a = 0; a < AG; a++)
#pragma unroll
      for (int o = 0; o < 8; o++) acc[a][o] = fmaf(av[a], w[o], acc[a][o]);
  }
#pragma unroll
  for (int g = 0; g < 4; g++) {
    const float2 v = __ldg(reinterpret_cast<const float2*>(bias + g * G_H + tx * 2));
#pragma unroll
    for (int a = 0; a < AG; a++) {
      acc[a][2 * g] += v.x;
      acc[a][2 * g + 1] += v.y;
    }
  }
}

/* Dense layer: column of output o = (o < 4 ? tx*4 + o : 128 + tx*4 + o - 4) (two conflict-free LDS.128 per k row);
 * W [K][256] streams from global memory through a two-chunk ring in shared memory.  Every thread of the CTA must call. */
__device__ __forceinline__ int dense_col(int tx, int o) { return o < 4 ? tx * 4 + o : G_FC / 2 + tx * 4 + o - 4; }

template <int AG>
__device__ __forceinline__ void dense_tile(const float* __restrict__ W, const float* __restrict__ bias, int K,
                                           const float* act /* smem [K][8 * AG] */, float* sw /* smem [G_WST][G_WKC][256] */,
                                           int tid, int tx, int ty, float acc[AG][8]) {
  constexpr int TILE = 8 * AG;
#pragma unroll
  for (int a = 0; a < AG; a++)
#pragma unroll
    for (int o = 0; o < 8; o++) acc[a][o] = 0.f;
  const int nchunks = (K + G_WKC - 1) / G_WKC;
  /* G_WST-deep ring: chunks ch+1 .. ch+G_WST-1 are in flight while chunk ch is used (one commit group per iteration, empty
   * past the end, so that "all but the newest G_WST-1 groups have landed" always means "chunk ch has landed") */
#pragma unroll
  for (int pre = 0; pre < G_WST - 1; pre++) {
    if (pre < nchunks) stage_rows(sw + pre * G_WKC * G_FC, W + (size_t)pre * G_WKC * G_FC, min(G_WKC, K - pre * G_WKC), tid);
    cp_async_commit();
  }
  for (int ch = 0; ch < nchunks; ch++) {
    const int k0 = ch * G_WKC, rows = min(G_WKC, K - k0);
    {
      const int nx = ch + G_WST - 1;
      if (nx < nchunks)
        stage_rows(sw + (nx % G_WST) * G_WKC * G_FC, W + (size_t)nx * G_WKC * G_FC, min(G_WKC, K - nx * G_WKC), tid);
      cp_async_commit();
      cp_async_wait<G_WST - 1>();
    }
    __syncthreads(); /* chunk ch has landed for everyone */
    const float* wc = sw + (ch % G_WST) * G_WKC * G_FC;
#pragma unroll(AG == 8 ? 4 : 8)
    for (int kk = 0; kk < rows; kk++) {
      const float4 v0 = *reinterpret_cast<const float4*>(wc + kk * G_FC + tx * 4);
      const float4 v1 = *reinterpret_cast<const float4*>(wc + kk * G_FC + G_FC / 2 + tx * 4);
      const float w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
      const int k = k0 + kk;
      float av[AG];
      if (AG == 8) {
        const float4 a0 = *reinterpret_cast<const float4*>(act + k * TILE + ty * 8);
        const float4 a1 = *reinterpret_cast<const float4*>(act + k * TILE + ty * 8 + 4);
        av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w;
        av[AG - 4] = a1.x; av[AG - 3] = a1.y; av[AG - 2] = a1.z; av[AG - 1] = a1.w;
      } else if (AG == 4) {
        const float4 a0 = *reinterpret_cast<const float4*>(act + k * TILE + ty * 4);
        av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[AG - 1] = a0.w;
      } else {
#pragma unroll
        for (int a = 0; a < AG; a++) av[a] = act[k * TILE + ty * AG + a];
      }
#pragma unroll
      for (int a = 0; a < AG; a++)
#pragma unroll
        for (int o = 0; o < 8; o++) acc[a][o] = fmaf(av[a], w[o], acc[a][o]);
    }
    __syncthreads(); /* everyone is done with this buffer before chunk ch + G_WST overwrites it */
  }
  const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + tx * 4));
  const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias + G_FC / 2 + tx * 4));
  const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
  for (int a = 0; a < AG; a++)
#pragma unroll
    for (int o = 0; o < 8; o++) acc[a][o] += bb[o];
}

/* fp32 forward: CTA = 8 * AG agents, 256 threads (tx = lane: output columns, ty = warp: AG agents).  AG = 8 for the full
 * batch; AG = G_FIX_AG = 2 for the fix-up of a short list (4x the CTAs, a quarter of the serial work per thread: the list is
 * short, so its cost is one CTA's latency).  The LSTM kernel is staged into shared memory once and stays for the ten steps;
 * the dense layers' weights stream through a two-chunk cp.async ring: 238.8 -> 141.0 us for the fix-up of the C5 bench
 * (20 % of the agents), 680 -> 550 us for the full-batch fp32 forward.  An agent's result does not depend on AG (same k order per output). */
template <int AG>
__global__ void __launch_bounds__(G_THREADS) k_ga3c_fwd(const Ga3cArgs p) {
  constexpr int TILE = 8 * AG;
  extern __shared__ __align__(16) float s_act[]; /* [256][TILE] transposed activations, then G_WBUF floats of weights */
  float* s_w = s_act + G_FC * TILE;
  __shared__ float s_logit[TILE][12];
  __shared__ int s_len[TILE];
  const size_t N = (size_t)p.E * p.A;
  const size_t g0 = (size_t)blockIdx.x * TILE;
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  const float* w = p.w;
  /* fix-up mode: slot a of this CTA is the agent fix_list[g0 + a] (N = empty slot); whole CTAs beyond the list leave */
  const int nfix = p.fix_mode ? *p.fix_count : 0;
  if (p.fix_mode && g0 >= (size_t)nfix) return;
  auto agent_of = [&](int a) -> size_t {
    if (!p.fix_mode) return g0 + a;
    return g0 + a < (size_t)nfix ? (size_t)p.fix_list[g0 + a] : N;
  };

  stage_rows(s_w, w + W_LSTM_K, G_OTHER + G_H, tid); /* the LSTM kernel stays in shared memory for all the steps */
  cp_async_commit();
  /* sequence length per agent (x[0], un-normalised: network.py:85); -1 = slot without a GA3C agent */
  if (tid < TILE) {
    size_t g = agent_of(tid);
    float x0 = g < N ? p.x[g * G_XS] : -1.f;
    s_len[tid] = (int)x0;
  }
  /* rows 0..6 = x_t (first step), rows 7..70 = h = 0 */
  for (int idx = tid; idx < (G_OTHER + G_H) * TILE; idx += G_THREADS) {
    int k = idx / TILE, a = idx - k * TILE;
    float v = 0.f;
    size_t g = agent_of(a);
    if (k < G_OTHER && g < N) v = (p.x[g * G_XS + 1 + G_HOST + k] - c_oth_avg[k]) / c_oth_std[k];
    s_act[idx] = v;
  }
  cp_async_wait<0>();
  __syncthreads();
  int maxlen = 0;
  for (int a = 0; a < TILE; a++) maxlen = max(maxlen, s_len[a]);
  maxlen = min(maxlen, G_SLOTS);

  float acc[AG][8];
  float cst[AG][2]; /* LSTM cell state of my 8 agents x 2 units */
#pragma unroll
  for (int a = 0; a < AG; a++) cst[a][0] = cst[a][1] = 0.f;
  for (int t = 0; t < maxlen; t++) {
    lstm_tile<AG>(s_w, w + W_LSTM_B, s_act, tx, ty, acc);
    __syncthreads(); /* everyone has read x_t and h */
#pragma unroll
    for (int a = 0; a < AG; a++) {
      const int ag = ty * AG + a;
      if (t < s_len[ag]) { /* dynamic_rnn(sequence_length): state copied through beyond the last valid step */
#pragma unroll
        for (int u = 0; u < 2; u++) { /* TF1 LSTMCell: i, j, f, o; forget_bias 1.0 */
          float gi = acc[a][u], gj = acc[a][2 + u], gf = acc[a][4 + u], go = acc[a][6 + u];
          float cn = sigmoid_f(gf + 1.0f) * cst[a][u] + sigmoid_f(gi) * tanhf(gj);
          cst[a][u] = cn;
          s_act[(G_OTHER + tx * 2 + u) * TILE + ag] = sigmoid_f(go) * tanhf(cn);
        }
      }
    }
    /* next step's inputs */
    if (t + 1 < maxlen)
      for (int idx = tid; idx < G_OTHER * TILE; idx += G_THREADS) {
        int k = idx / TILE, a = idx - k * TILE;
        size_t g = agent_of(a);
        float v = 0.f;
        if (g < N && t + 1 < s_len[a]) v = (p.x[g * G_XS + 1 + G_HOST + G_OTHER * (t + 1) + k] - c_oth_avg[k]) / c_oth_std[k];
        s_act[idx] = v;
      }
    __syncthreads();
  }
  /* layer1 input = concat(host4, h64): host goes to rows 3..6, h already sits in rows 7..70 */
  if (tid < G_HOST * TILE) {
    int k = tid / TILE, a = tid - k * TILE;
    size_t g = agent_of(a);
    s_act[(3 + k) * TILE + a] = g < N ? (p.x[g * G_XS + 1 + k] - c_host_avg[k]) / c_host_std[k] : 0.f;
  }
  __syncthreads();
  const float* Wk[3] = {w + W_L1_K, w + W_L2_K, w + W_FC_K};
  const float* Wb[3] = {w + W_L1_B, w + W_L2_B, w + W_FC_B};
#pragma unroll 1
  for (int layer = 0; layer < 3; layer++) {
    if (layer == 0)
      dense_tile<AG>(Wk[0], Wb[0], G_HOST + G_H, s_act + 3 * TILE, s_w, tid, tx, ty, acc);
    else
      dense_tile<AG>(Wk[layer], Wb[layer], G_FC, s_act, s_w, tid, tx, ty, acc);
    /* (dense_tile ends on a barrier: everyone has read the layer's input) */
#pragma unroll
    for (int a = 0; a < AG; a++)
#pragma unroll
      for (int o = 0; o < 8; o++) s_act[dense_col(tx, o) * TILE + ty * AG + a] = fmaxf(acc[a][o], 0.f);
    __syncthreads();
  }
  /* logits_p: TILE agents x 11 outputs, K = 256 */
  if (tid < 4 * TILE) {
    const int a = tid >> 2, q = tid & 3;
    float l[3] = {0.f, 0.f, 0.f};
    for (int k = 0; k < G_FC; k++) {
      float v = s_act[k * TILE + a];
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
  if (tid < TILE) {
    const size_t g = agent_of(tid);
    if (g < N && s_len[tid] >= 0) {
      const int flags = __float_as_int(ga3c_src(p, g)[5 * N + g].z);
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
        const float pref = ga3c_src(p, g)[2 * N + g].w;
        p.policy_actions[2 * g] = (float)((double)pref * (double)c_actions[best][0]); /* GA3CCADRLPolicy.py:40-42 */
        p.policy_actions[2 * g + 1] = c_actions[best][1];
      }
    } else if (g < N && p.action_idx) {
      p.action_idx[g] = -1;
    }
  }
}

/* ------------------------------------------------------------------------------------------------------------
 * k_ga3c_tc: the same network on the 5th-generation tensor cores (tcgen05.mma, kind::f16, 16-bit operands -> fp32),
 * TRANSPOSED: D^T[outputs x agents] = W^T[outputs x K] * X^T[K x agents].  Weights are the A operand (M = 128 output
 * rows <-> the 128 TMEM lanes), agents are the N dimension (TMEM columns).  Why: the epilogue is MUFU-bound (5 tanh
 * per LSTM unit per agent per step), MUFU cost is per warp instruction, and a thread can only read its own TMEM
 * lane -- with agents on lanes, 20480 agents = 640 row-warps cannot be spread evenly over 592 SM sub-partitions, and
 * partly filled warps cost as much as full ones.  With outputs on lanes every warp is full and the agents per CTA
 * are free: 144 per CTA, one CTA per SM.
 *
 * LSTM step: lane l <-> (unit u = l % 64, agent group l / 64); group 0 = agents 0..71 of the CTA, group 1 = 72..143.
 * Four accumulators (gates i, j, f, o; N = 80 columns each) so that a thread finds all four gates of its unit in
 * its own lane.  The group split rides on K: gate g's A tile is [W_g ; 0] for K-slabs 0..4 (which multiply agents
 * 0..79) and [0 ; W_g] for K-slabs 5..9 (agents 72..151), both cut out of one shared-memory image laid out
 * [0 | W_i | 0 | W_j | 0 | W_f | 0 | W_o | 0] (64-row blocks) by moving the descriptor start.  K = 80 per group =
 * [h(64) | x_t(7) | 1 | 1 | 0..]; the two "1" columns carry the bias as a 16-bit hi + lo pair.
 * layer1 / layer2 / fullyconnected1: two M = 128 blocks x N = 144; logits_p: one block (11 of 128 rows used).
 * Operands sit in shared memory in the canonical K-major SWIZZLE_NONE layout (8-row x 16-byte core matrices): element
 * (r, k) at (k/8)*LBO + r*16 + (k%8)*2, SBO = 128.  The activation tile X[agent][k] uses LBO = 153*16 bytes: the odd
 * pitch makes the epilogue's 2-byte column stores (32 lanes = 32 consecutive k of one agent) conflict-free.
 * Roles: warps 0-7 epilogue (warp w reads TMEM lanes 32*(w%4).., warps 4-7 take the other half of the columns),
 * warp 8 lane 0 = MMA issuer, warp 9 lane 0 = weight producer (cp.async.bulk: the 90 KB LSTM image once, then 8 KB
 * slabs of the dense layers through a ring that grows into the LSTM image's space when the recurrence is done).
 * Gate non-linearities use tanh.approx (sigmoid(x) = 0.5 tanh(x/2) + 0.5).
 * ---------------------------------------------------------------------------------------------------------- */
#define U_G 72                     /* agents per lane group */
#define U_XCH (153 * 16)           /* bytes per 8-wide K chunk of the activation tile (153 rows: 144 + overhang + odd) */
#define U_X_BYTES (32 * U_XCH)     /* K = 256 */
#define U_LBLK 1024                /* one 64-row block of a K chunk */
#define U_LCH (9 * U_LBLK)         /* bytes per K chunk of the LSTM block image */
#define U_L_BYTES (10 * U_LCH)     /* K = 80 */
#define U_SLAB 8192                /* K16 x 256 rows */
#define U_RING0 6                  /* ring slots available during the recurrence */
#define U_RING 17                  /* + 11 slots inside the LSTM image's space afterwards */
#define U_NSLAB 38                 /* 5 (layer1) + 16 (layer2) + 16 (fc1) + 1 (logits_p: 32 chunks x 256 B) */
#define U_IMG_BYTES (U_L_BYTES + U_NSLAB * U_SLAB)
#define U_NBIAS (G_FC + G_FC + 16) /* fp32 biases of layer2, fc1, logits_p appended to the image */
#define U_NREP 4                   /* replicas of the image in global memory */
#define U_IMG_STRIDE ((U_IMG_BYTES + U_NBIAS * 4 + 255) / 256 * 256)
#define U_DYN_SMEM (U_X_BYTES + U_L_BYTES + U_RING0 * U_SLAB)
#define U_EPI 256
#define U_THREADS 320

/* slab q of the stream (5 + 16 + 16 + 1) -> slab of the image this CTA loads at that point: each layer's K slabs
 * rotated by the CTA index */
__device__ __forceinline__ int rot_slab(int q, unsigned cta) {
  if (q < 5) return (int)((q + cta) % 5);
  if (q < 21) return 5 + (int)((q - 5 + cta) % 16);
  if (q < 37) return 21 + (int)((q - 21 + cta) % 16);
  return q;
}

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define TRACE(role, idx)                                                                          \
  do {                                                                                            \
    const int i_ = (idx);                                                                         \
    if (p.trace && i_ < 64) p.trace[((size_t)blockIdx.x * 2 + (role)) * 64 + i_] = gtime_ns();    \
  } while (0)

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
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); } /* epilogue warps only */

/* shared-memory matrix descriptor, SWIZZLE_NONE, K-major (cute::UMMA::SmemDescriptor bit layout): start >> 4 in
 * [0,14), LBO >> 4 in [16,30), SBO >> 4 in [32,46), version 1 in [46,48), layout type 0 in [61,64) */
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         (1ull << 46);
}
/* instruction descriptor (cute::UMMA::InstrDescriptor): D fp32 [4,6) = 1, A / B format [7,10) / [10,13): 0 = fp16, 1 = bf16,
 * both K-major (bits 15, 16 = 0), N >> 3 in [17,23), M >> 4 in [24,29) */
__device__ __forceinline__ uint32_t umma_idesc(int M, int N) {
  return (1u << 4) | (U_FP16 ? 0u : ((1u << 7) | (1u << 10))) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24); /* A / B format: 0 = fp16, 1 = bf16 */
}
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
/* thread t of warp w reads TMEM lane 32*(w%4) + t, 4 / 8 consecutive fp32 columns */
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float v[4]) {
  uint32_t r[4];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(taddr));
#pragma unroll
  for (int k = 0; k < 4; k++) v[k] = __uint_as_float(r[k]);
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float v[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
#pragma unroll
  for (int k = 0; k < 8; k++) v[k] = __uint_as_float(r[k]);
}
__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }
__device__ __forceinline__ void sts16(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts_op16(uint32_t addr, float v) {
#if U_FP16
  const __half b = __float2half_rn(fminf(fmaxf(v, -U_MAX16), U_MAX16));
#else
  const __nv_bfloat16 b = __float2bfloat16_rn(v);
#endif
  asm volatile("st.shared.b16 [%0], %1;" ::"r"(addr), "h"(*reinterpret_cast<const unsigned short*>(&b)) : "memory");
}

__global__ void __launch_bounds__(U_THREADS, 1) k_ga3c_tc(const Ga3cArgs p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_bar[2 * U_RING + 4];
  __shared__ uint32_t s_tmem;
  __shared__ int s_len[U_NA + 16];
  __shared__ int s_maxlen;

  const size_t N = (size_t)p.E * p.A;
  const size_t g0 = (size_t)blockIdx.x * U_NA;
  const size_t npad = (size_t)gridDim.x * U_NA; /* agents per xb slot plane */
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t sX = smem_u32(smem), sL = sX + U_X_BYTES, sR = sL + U_L_BYTES;
  const uint32_t bar_full = smem_u32(&s_bar[0]), bar_empty = smem_u32(&s_bar[U_RING]);
  const uint32_t bar_a = smem_u32(&s_bar[2 * U_RING]), bar_d = bar_a + 8, bar_l = bar_a + 16, bar_ldone = bar_a + 24;
  auto slot_addr = [&](int slot) { return slot < U_RING0 ? sR + slot * U_SLAB : sL + (slot - U_RING0) * U_SLAB; };

  if (tid == 0) {
    for (int k = 0; k < U_RING; k++) {
      mbar_init(bar_full + 8 * k, 1);
      mbar_init(bar_empty + 8 * k, 1);
    }
    mbar_init(bar_a, U_EPI);
    mbar_init(bar_d, 1);
    mbar_init(bar_l, 1);
    mbar_init(bar_ldone, 1);
    s_maxlen = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) { /* TMEM: all 512 fp32 columns x 128 lanes (one CTA per SM); this warp also frees them */
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  /* header chunk of every agent: {host0,host1 | host2,host3 | 1,1 | sequence length}; -1 = no GA3C agent here */
  uint4 hdr = make_uint4(0u, 0u, 0u, 0xffffffffu);
  if (tid < U_NA) {
    hdr = p.xb[(size_t)(U_XB_SLOTS - 1) * npad + g0 + tid];
    s_len[tid] = min((int)hdr.w, G_SLOTS);
  } else if (tid < U_NA + 16) {
    s_len[tid] = -1;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid < U_NA && s_len[tid] > 0) atomicMax(&s_maxlen, s_len[tid]);
  __syncthreads();
  const int maxlen = s_maxlen;
  const uint32_t tmem = s_tmem;

  if (warp == 9) {
    /* ===== weight producer ===== */
    if (lane == 0) {
      /* All CTAs want the same bytes at the same time: to keep them from queueing on the same L2 lines, each CTA
       * reads its own replica (blockIdx % U_NREP) and walks every layer's K slabs from its own starting point
       * (accumulation order over K is free); the MMA issuer uses the same rotation. */
      const uint8_t* img = p.wimg + (size_t)(blockIdx.x % U_NREP) * U_IMG_STRIDE;
      mbar_expect_tx(bar_l, U_L_BYTES);
      for (int k = 0; k < 10; k++) {
        const int kk = (k + blockIdx.x) % 10;
        bulk_g2s(sL + kk * U_LCH, img + (size_t)kk * U_LCH, U_LCH, bar_l);
      }
      const uint8_t* slabs = img + U_L_BYTES;
      for (int q = 0; q < U_NSLAB; q++) {
        const int slot = q % U_RING;
        if (q == U_RING0) mbar_wait(bar_ldone, 0); /* the recurrence is over: its weights' space joins the ring */
        if (q >= U_RING) mbar_wait(bar_empty + 8 * slot, ((q / U_RING) - 1) & 1);
        mbar_expect_tx(bar_full + 8 * slot, U_SLAB);
        bulk_g2s(slot_addr(slot), slabs + (size_t)rot_slab(q, blockIdx.x) * U_SLAB, U_SLAB, bar_full + 8 * slot);
      }
    }
  } else if (warp == 8) {
    /* ===== MMA issuer ===== */
    if (lane == 0) {
      const uint32_t id80 = umma_idesc(128, 80), id144 = umma_idesc(128, U_NA);
      uint32_t a_phase = 0;
      int ti = 0;
      TRACE(1, ti++);
      mbar_wait(bar_l, 0);
      TRACE(1, ti++);
      for (int t = 0; t < maxlen; t++) {
        mbar_wait(bar_a, a_phase);
        a_phase ^= 1;
        tc_fence_after();
        TRACE(1, ti++);
#pragma unroll 1
        for (int g = 0; g < 4; g++)
#pragma unroll
          for (int s = 0; s < 10; s++) {
            const int grp = s / 5, sc = s % 5;
            umma_f16(tmem + g * 80, umma_desc(sL + (2 * sc) * U_LCH + (2 * g + 1 - grp) * U_LBLK, U_LCH, 128),
                      umma_desc(sX + (2 * sc) * U_XCH + grp * U_G * 16, U_XCH, 128), id80, s > 0);
          }
        if (t == maxlen - 1) tc_commit(bar_ldone);
        tc_commit(bar_d);
      }
      if (maxlen == 0) mbar_arrive(bar_ldone);
      int q = 0;
      for (int layer = 0; layer < 3; layer++) { /* layer1 (5 slabs), layer2 (16), fullyconnected1 (16) */
        const int nslab = layer == 0 ? 5 : G_FC / 16;
        mbar_wait(bar_a, a_phase);
        a_phase ^= 1;
        TRACE(1, ti++);
        for (int s = 0; s < nslab; s++, q++) {
          const int slot = q % U_RING;
          mbar_wait(bar_full + 8 * slot, (q / U_RING) & 1);
          tc_fence_after();
          const int ks = rot_slab(q, blockIdx.x) - (q - s); /* which K slab of this layer sits in the slot */
          const uint64_t bd = umma_desc(sX + (2 * ks) * U_XCH, U_XCH, 128);
          umma_f16(tmem, umma_desc(slot_addr(slot), 4096, 128), bd, id144, s > 0);
          umma_f16(tmem + U_NA, umma_desc(slot_addr(slot) + 2048, 4096, 128), bd, id144, s > 0);
          tc_commit(bar_empty + 8 * slot);
        }
        tc_commit(bar_d);
        TRACE(1, ti++);
      }
      { /* logits_p: one 8 KB slab = 32 K-chunks of 16 rows x 16 B; rows 16..127 of the A tile read whatever follows */
        const int slot = q % U_RING;
        mbar_wait(bar_a, a_phase);
        a_phase ^= 1;
        mbar_wait(bar_full + 8 * slot, (q / U_RING) & 1);
        tc_fence_after();
        for (int s = 0; s < G_FC / 16; s++)
          umma_f16(tmem, umma_desc(slot_addr(slot) + s * 512, 256, 128), umma_desc(sX + (2 * s) * U_XCH, U_XCH, 128), id144,
                    s > 0);
        tc_commit(bar_d);
      }
    }
  } else {
    /* ===== epilogue: TMEM lane l <-> output row; warps 4-7 take the second half of the columns ===== */
    const int l = (warp & 3) * 32 + lane, hw = warp >> 2;
    const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    /* X: h = 0 (chunks 0..7), chunk 8 = first LSTM input (or layer1's host chunk), chunk 9 = [1, 0..] / zeros;
     * rows 144..152 are only read as overhang of the second agent group and stay zero */
    for (int k = tid; k < 10 * 153; k += U_EPI) sts16(sX + (k / 153) * U_XCH + (k % 153) * 16, make_uint4(0u, 0u, 0u, 0u));
    epi_bar();
    const uint32_t xrow = sX + tid * 16; /* tid < U_NA: my agent's row */
    const uint4 host_chunk = make_uint4(hdr.x, hdr.y, hdr.z, 0u);
    if (tid < U_NA) {
      if (maxlen > 0) {
        sts16(xrow + 8 * U_XCH, p.xb[g0 + tid]);
        sts16(xrow + 9 * U_XCH, make_uint4(U_ONE16, 0u, 0u, 0u));
      } else {
        sts16(xrow + 8 * U_XCH, host_chunk);
      }
    }
    fence_async_smem();
    mbar_arrive(bar_a);

    uint32_t d_phase = 0;
    int ti = 0;
    if (tid == 0) TRACE(0, ti++);
    {
      /* recurrence: unit u of agent group grp; my 36 columns c0..c0+35 of the group's 72 */
      const int u = l & 63, grp = l >> 6, c0 = hw * (U_G / 2);
      const uint32_t hdst = sX + (u >> 3) * U_XCH + (u & 7) * 2 + (grp * U_G + c0) * 16;
      const int* lenp = s_len + grp * U_G + c0;
      float cst[U_G / 2];
#pragma unroll
      for (int c = 0; c < U_G / 2; c++) cst[c] = 0.f;
      for (int t = 0; t < maxlen; t++) {
        uint4 nx = make_uint4(0u, 0u, 0u, 0u);
        if (tid < U_NA && t + 1 < maxlen) nx = p.xb[(size_t)(t + 1) * npad + g0 + tid];
        mbar_wait(bar_d, d_phase);
        d_phase ^= 1;
        tc_fence_after();
        if (tid == 0) TRACE(0, ti++);
#pragma unroll
        for (int cc = 0; cc < U_G / 2; cc += 4) {
          float gi[4], gj[4], gf[4], go[4];
          tmem_ld4(trow + c0 + cc, gi);
          tmem_ld4(trow + 80 + c0 + cc, gj);
          tmem_ld4(trow + 160 + c0 + cc, gf);
          tmem_ld4(trow + 240 + c0 + cc, go);
          tc_wait_ld();
#pragma unroll
          for (int k = 0; k < 4; k++) { /* TF1 LSTMCell: gates i, j, f, o; forget_bias 1.0 */
            const float cn = fmaf(sigmoid_fast(gf[k] + 1.0f), cst[cc + k], sigmoid_fast(gi[k]) * tanh_fast(gj[k]));
            const float hn = sigmoid_fast(go[k]) * tanh_fast(cn);
            if (t < lenp[cc + k]) { /* dynamic_rnn(sequence_length): state copied through beyond the last valid step */
              cst[cc + k] = cn;
              sts_op16(hdst + (cc + k) * 16, hn);
            }
          }
        }
        if (tid < U_NA) {
          if (t + 1 < maxlen) {
            sts16(xrow + 8 * U_XCH, nx);
          } else { /* layer1's K chunks 8, 9: [host0..3 normalised, 1, 1, 0, 0], zeros (network.py:83-84, 89-90) */
            sts16(xrow + 8 * U_XCH, host_chunk);
            sts16(xrow + 9 * U_XCH, make_uint4(0u, 0u, 0u, 0u));
          }
        }
        fence_async_smem();
        tc_fence_before();
        mbar_arrive(bar_a);
        if (tid == 0) TRACE(0, ti++);
      }
    }
    /* layer1, layer2, fullyconnected1: ReLU(D + bias) -> 16-bit -> X (in place: the MMAs that read X have completed
     * when bar_d fires); output row o = mb*128 + l is K index o of the next GEMM */
    const float* biasg = reinterpret_cast<const float*>(p.wimg + U_IMG_BYTES);
    for (int layer = 0; layer < 3; layer++) {
      float b0 = 0.f, b1 = 0.f; /* layer1's bias rides in the GEMM */
      if (layer > 0) {
        b0 = __ldg(biasg + (layer - 1) * G_FC + l);
        b1 = __ldg(biasg + (layer - 1) * G_FC + 128 + l);
      }
      mbar_wait(bar_d, d_phase);
      d_phase ^= 1;
      tc_fence_after();
      if (tid == 0) TRACE(0, ti++);
#pragma unroll
      for (int mb = 0; mb < 2; mb++) {
        const float b = mb ? b1 : b0;
        const int o = mb * 128 + l;
        const uint32_t dst = sX + (o >> 3) * U_XCH + (o & 7) * 2 + hw * U_G * 16;
#pragma unroll 3
        for (int cc = 0; cc < U_G; cc += 8) {
          float v[8];
          tmem_ld8(trow + mb * U_NA + hw * U_G + cc, v);
          tc_wait_ld();
#pragma unroll
          for (int k = 0; k < 8; k++) sts_op16(dst + (cc + k) * 16, fmaxf(v[k] + b, 0.f));
        }
      }
      fence_async_smem();
      tc_fence_before();
      mbar_arrive(bar_a);
      if (tid == 0) TRACE(0, ti++);
    }
    /* logits_p: lanes 0..10 hold the 11 logits of every agent (column); transpose through shared memory (X is dead) */
    mbar_wait(bar_d, d_phase);
    tc_fence_after();
    if (tid == 0) TRACE(0, ti++);
    float* s_logit = reinterpret_cast<float*>(smem); /* [U_NA][12] */
    {
      const float pb = l < G_ACT ? __ldg(biasg + 2 * G_FC + l) : 0.f;
#pragma unroll 3
      for (int cc = 0; cc < U_G; cc += 8) {
        float v[8];
        tmem_ld8(trow + hw * U_G + cc, v);
        tc_wait_ld();
        if (l < G_ACT) {
#pragma unroll
          for (int k = 0; k < 8; k++) s_logit[(hw * U_G + cc + k) * 12 + l] = v[k] + pb;
        }
      }
    }
    tc_fence_before();
    epi_bar();
    const size_t g = g0 + tid;
    if (tid < U_NA && g < N) {
      if (s_len[tid] < 0) {
        if (p.action_idx) p.action_idx[g] = -1;
      } else {
        /* softmax (+1e-4 floor) -> argmax -> action (network.py:45-51, GA3CCADRLPolicy.py:38-42) */
        const int flags = __float_as_int(ga3c_src(p, g)[5 * N + g].z);
        if (flags & CASIM_F_IS_DONE) { /* collision_avoidance_env.py:298-300 */
          p.policy_actions[2 * g] = 0.f;
          p.policy_actions[2 * g + 1] = 0.f;
          if (p.action_idx) p.action_idx[g] = -1;
        } else {
          float lg[G_ACT], mx = -INFINITY;
#pragma unroll
          for (int k = 0; k < G_ACT; k++) {
            lg[k] = s_logit[tid * 12 + k];
            mx = fmaxf(mx, lg[k]);
          }
          float pr[G_ACT], sum = 0.f;
#pragma unroll
          for (int k = 0; k < G_ACT; k++) {
            pr[k] = expf(lg[k] - mx);
            sum += pr[k];
          }
          int best = 0;
          float pbest = -1.f, psecond = -1.f;
#pragma unroll
          for (int k = 0; k < G_ACT; k++) {
            pr[k] = (pr[k] / sum + 1e-4f) / (1.0f + 1e-4f * G_ACT);
            if (pr[k] > pbest) {
              psecond = pbest;
              pbest = pr[k];
              best = k;
            } else {
              psecond = fmaxf(psecond, pr[k]);
            }
          }
          if (p.fix_list != nullptr && pbest - psecond < G_FIX_MARGIN) /* too close to call on the tensor cores: fp32 decides */
            p.fix_list[atomicAdd(p.fix_count, 1)] = (int32_t)g;
          if (p.probs) {
#pragma unroll
            for (int k = 0; k < G_ACT; k++) p.probs[g * G_ACT + k] = pr[k];
          }
          if (p.action_idx) p.action_idx[g] = best;
          const float pref = ga3c_src(p, g)[2 * N + g].w;
          p.policy_actions[2 * g] = (float)((double)pref * (double)c_actions[best][0]);
          p.policy_actions[2 * g + 1] = c_actions[best][1];
        }
      }
    }
  }
  if (tid == 0) TRACE(0, 63);
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

}  // namespace

static uint16_t f32_to_op16(float f) { /* round to nearest even, like cvt.rn.bf16.f32 / cvt.rn.f16.f32 */
#if U_FP16
  const __half hv = __float2half_rn(f);
  uint16_t r;
  memcpy(&r, &hv, 2);
  return r;
#endif
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static float op16_to_f32(uint16_t b) {
#if U_FP16
  __half hv;
  memcpy(&hv, &b, 2);
  return __half2float(hv);
#endif
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
  /* tensor-core path: 16-bit image of every layer in the shared-memory order k_ga3c_tc feeds to tcgen05.mma as the
   * A operand (K-major, SWIZZLE_NONE: element (row, k) of a layer at (k/8)*LBO + row*16 + (k%8)*2) */
  std::vector<uint8_t> img(U_IMG_BYTES + U_NBIAS * sizeof(float), 0);
  auto put = [&](size_t layer_off, size_t lbo, int k, int row, float v) {
    uint16_t b = f32_to_op16(v);
    memcpy(&img[layer_off + (size_t)(k / 8) * lbo + (size_t)row * 16 + (k % 8) * 2], &b, 2);
  };
  auto put_bias = [&](size_t layer_off, size_t lbo, int k, int row, float v) { /* 16-bit hi + lo pair against two "1" columns */
    uint16_t hi = f32_to_op16(v);
    put(layer_off, lbo, k, row, v);
    put(layer_off, lbo, k + 1, row, v - op16_to_f32(hi));
  };
  const float* w = weights_host;
  /* LSTM: per K chunk nine 64-row blocks [0 | W_i | 0 | W_j | 0 | W_f | 0 | W_o | 0]; K order [h(64) | x(7) | 1 | 1];
   * TF kernel rows are [x(7); h(64)], columns gate*64 + unit */
  for (int g = 0; g < 4; g++)
    for (int u = 0; u < G_H; u++) {
      const int n = g * G_H + u, row = (2 * g + 1) * 64 + u;
      for (int k = 0; k < G_H; k++) put(0, U_LCH, k, row, w[W_LSTM_K + (G_OTHER + k) * 4 * G_H + n]);
      for (int k = 0; k < G_OTHER; k++) put(0, U_LCH, G_H + k, row, w[W_LSTM_K + k * 4 * G_H + n]);
      put_bias(0, U_LCH, G_H + G_OTHER, row, w[W_LSTM_B + n]);
    }
  const size_t o1 = U_L_BYTES, o2 = o1 + 5 * U_SLAB, o3 = o2 + 16 * U_SLAB, o4 = o3 + 16 * U_SLAB;
  for (int n = 0; n < G_FC; n++) { /* layer1: K order [h(64) | host(4) | 1 | 1]; TF kernel rows are [host(4); h(64)] */
    for (int k = 0; k < G_H; k++) put(o1, 4096, k, n, w[W_L1_K + (G_HOST + k) * G_FC + n]);
    for (int k = 0; k < G_HOST; k++) put(o1, 4096, G_H + k, n, w[W_L1_K + k * G_FC + n]);
    put_bias(o1, 4096, G_H + G_HOST, n, w[W_L1_B + n]);
    for (int k = 0; k < G_FC; k++) {
      put(o2, 4096, k, n, w[W_L2_K + k * G_FC + n]);
      put(o3, 4096, k, n, w[W_FC_K + k * G_FC + n]);
    }
  }
  for (int n = 0; n < G_ACT; n++)
    for (int k = 0; k < G_FC; k++) put(o4, 256, k, n, w[W_P_K + k * G_ACT + n]);
  float* bias = reinterpret_cast<float*>(&img[U_IMG_BYTES]);
  memcpy(bias, w + W_L2_B, G_FC * sizeof(float));
  memcpy(bias + G_FC, w + W_FC_B, G_FC * sizeof(float));
  memcpy(bias + 2 * G_FC, w + W_P_B, G_ACT * sizeof(float));
  if (!h->ga3c_wb) {
    e = cudaMalloc(&h->ga3c_wb, (size_t)U_NREP * U_IMG_STRIDE);
    if (e != cudaSuccess) return e;
  }
  for (int r = 1; r < U_NREP; r++) {
    e = cudaMemcpyAsync((uint8_t*)h->ga3c_wb + (size_t)r * U_IMG_STRIDE, img.data(), img.size(), cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return e;
  }
  if (!h->ga3c_xb) { /* allocated here, not in the launcher: the forward must be capturable into a CUDA graph */
    const size_t ntiles = ((size_t)h->E * h->A + U_NA - 1) / U_NA;
    const size_t bytes = ntiles * U_XB_SLOTS * U_NA * sizeof(uint4);
    e = cudaMalloc(&h->ga3c_xb, bytes);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(h->ga3c_xb, 0, bytes, s); /* agents past N in the last tile stay zero (sequence length 0) */
    if (e != cudaSuccess) return e;
  }
  if (!h->ga3c_x) {
    e = cudaMalloc((void**)&h->ga3c_x, (size_t)h->E * h->A * G_XS * sizeof(float));
    if (e != cudaSuccess) return e;
  }
  e = cudaFuncSetAttribute(k_ga3c_fwd<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(((size_t)G_FC * G_TILE + G_WBUF) * sizeof(float)));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_ga3c_fwd<G_FIX_AG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(((size_t)G_FC * 8 * G_FIX_AG + G_WBUF) * sizeof(float)));
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_ga3c_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)U_DYN_SMEM);
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
  const unsigned ntiles = (unsigned)((N + U_NA - 1) / U_NA);
  Ga3cArgs a;
  a.c = h->dc;
  a.state = reinterpret_cast<const float4*>(state);
  a.state0 = reinterpret_cast<const float4*>(h->state0);
  a.prev_game_over = (h->dc.auto_reset && h->init) ? h->prev_game_over : nullptr;
  a.n_agents = h->n_agents;
  a.w = h->ga3c_w;
  a.x = h->ga3c_x;
  a.xb = (use_tc & 1) ? reinterpret_cast<uint4*>(h->ga3c_xb) : nullptr;
  a.wimg = reinterpret_cast<const uint8_t*>(h->ga3c_wb);
  a.trace = nullptr;
  a.fix_list = nullptr;
  a.fix_count = nullptr;
  a.fix_mode = 0;
  const bool exact = (use_tc & 1) && (use_tc & 4); /* bit 2: fp32 re-evaluation of the agents the tensor-core forward cannot call */
  if (exact) {
    if (!h->ga3c_fix) {
      cudaError_t e = cudaMalloc((void**)&h->ga3c_fix, (N + 1) * sizeof(int32_t));
      if (e != cudaSuccess) return e;
      e = cudaMemsetAsync(h->ga3c_fix, 0, (N + 1) * sizeof(int32_t), s);
      if (e != cudaSuccess) return e;
    }
    a.fix_count = h->ga3c_fix;
    a.fix_list = h->ga3c_fix + 1;
  }
  static const char* trace_path = casim_tuning_str("CASIM_GA3C_TRACE"); /* tuning aid: dumps phase time stamps, synchronises */
  if (use_tc && trace_path) {
    cudaMalloc((void**)&a.trace, (size_t)ntiles * 128 * sizeof(unsigned long long));
    cudaMemsetAsync(a.trace, 0, (size_t)ntiles * 128 * sizeof(unsigned long long), s);
  }
  a.policy_actions = policy_actions;
  a.action_idx = action_idx;
  a.probs = probs;
  a.nn_input = nn_input;
  a.E = h->E;
  a.A = h->A;
  const bool skip_state = (use_tc & 2) != 0; /* profiling aid: forward only, on the state vectors of the last call */
  use_tc &= 1;
  h->launches += skip_state ? 1 : 2;
  if (!skip_state) {
    k_ga3c_state<<<(unsigned)((N + GS_AG - 1) / GS_AG), GS_THREADS, 0, s>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  if (use_tc)
    k_ga3c_tc<<<ntiles, U_THREADS, U_DYN_SMEM, s>>>(a);
  if (use_tc && a.trace) {
    std::vector<unsigned long long> t((size_t)ntiles * 128);
    cudaStreamSynchronize(s);
    cudaMemcpy(t.data(), a.trace, t.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    cudaFree(a.trace);
    if (FILE* f = fopen(trace_path, "wb")) {
      fwrite(t.data(), sizeof(unsigned long long), t.size(), f);
      fclose(f);
    }
    return cudaGetLastError();
  }
  if (use_tc) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess || !exact) return e;
    if (skip_state) return cudaErrorInvalidValue; /* the state kernel zeroes the fix-up counter */
    h->launches++;
    a.fix_mode = 1;
    k_ga3c_fwd<G_FIX_AG><<<(unsigned)((N + 8 * G_FIX_AG - 1) / (8 * G_FIX_AG)), G_THREADS,
                           ((size_t)G_FC * 8 * G_FIX_AG + G_WBUF) * sizeof(float), s>>>(a);
    return cudaGetLastError();
  } else
    k_ga3c_fwd<8><<<(unsigned)((N + G_TILE - 1) / G_TILE), G_THREADS, ((size_t)G_FC * G_TILE + G_WBUF) * sizeof(float), s>>>(a);
  return cudaGetLastError();
}
