/*
 * casim_oracle_ga3c.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see casim_oracle.c).
 *
 * Plain-C restatement of the GA3C-CADRL policy of mlodel/gym-exploration-2d
 * (gym_collision_avoidance/envs/policies/...):
 *   GA3CCADRLPolicy.py:45-106  agents_to_ga3c_cadrl_state  (fp64, Python round(x, 2), stable sort, clip)
 *   GA3CCADRLPolicy.py:34-43   find_next_action            (argmax -> action table)
 *   GA3C_CADRL/network.py:8-17   Actions
 *   GA3C_CADRL/network.py:65-98  NetworkVP_rnn._create_graph  (normalise, LSTM(64), layer1, layer2)
 *   GA3C_CADRL/network.py:45-51  _create_graph_outputs        (fullyconnected1, logits_p, softmax + floor)
 *   GA3C_CADRL/network.py:101-151 Config (sizes, normalisation vectors)
 *
 * Parity status: the STATE VECTOR (a23) is PINNED against the reference function executed in the build
 * container (oracle/gen_golden_ga3c.py -> tests/golden/ga3c_state.npz).  The NETWORK NUMERICS (a24) cannot be pinned to
 * TensorFlow: tensorflow==1.15.0 (requirements.txt:1) is not installable here and the reference holds no test vectors.
 * They are pinned instead (round 2) to an independent numpy evaluation of the TF graph, float64 and float32, on the
 * reference's shipped IROS18 weights and 1574 reference-generated state vectors (oracle/gen_golden_ga3c_probs.py ->
 * tests/golden/ga3c_probs.npz; test_ga3c_network_vs_tf_graph_golden: |dp| <= 5e-6, argmax identical).  What stays
 * unpinned is TensorFlow's own float32 rounding.  This file restates the graph and TF1 tf.contrib.rnn.LSTMCell semantics (one kernel [7+64, 256],
 * gate order i, j, f, o, forget_bias 1.0, zero initial state, dynamic_rnn(sequence_length) returning the state
 * at the last valid step) in fp32 and is cross-checked against torch.nn.LSTMCell / torch.nn.functional in
 * tests/test_oracle_golden.py.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/casim.h"

#define NF CASIM_STATE_WORDS
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

#define NN_IN CASIM_GA3C_INPUT /* 75 */
#define NN_SLOTS 10            /* network.py:137 MAX_NUM_OTHER_AGENTS_OBSERVED of the network */
#define NN_OTHER 7
#define NN_HOST 4
#define NN_H 64
#define NN_FC 256
#define NN_ACT CASIM_GA3C_NUM_ACTIONS

/* Python round(x, 2) as an integer number of hundredths: correctly rounded decimal, ties to even
 * (float.__round__ -> dtoa mode 3).  glibc printf is correctly rounded on the exact binary value. */
static long long py_round2_hundredths(double x) {
  char buf[64];
  snprintf(buf, sizeof buf, "%.2f", x);
  double r = strtod(buf, NULL);
  return llround(r * 100.0);
}

/* network.py:8-17 Actions: np.mgrid[start:stop:step] rows = start + k*step */
void casim_oracle_ga3c_actions(double* table /* [11][2] */) {
  int r = 0;
  for (int k = 0; k < 5; k++) { table[2 * r] = 1.0; table[2 * r + 1] = -M_PI / 6 + k * (M_PI / 12); r++; }
  for (int k = 0; k < 3; k++) { table[2 * r] = 0.5; table[2 * r + 1] = -M_PI / 6 + k * (M_PI / 6); r++; }
  for (int k = 0; k < 3; k++) { table[2 * r] = 0.0; table[2 * r + 1] = -M_PI / 6 + k * (M_PI / 6); r++; }
}

/* agent.py:250-269 get_ref (the axes left by the last update_ego_frame) */
static void ref_axes(const double* s, double* prll, double* orth) {
  double gx = s[CASIM_GX] - s[CASIM_PX], gy = s[CASIM_GY] - s[CASIM_PY];
  double d = sqrt(gx * gx + gy * gy);
  if (d > 1e-8) { prll[0] = gx / d; prll[1] = gy / d; } else { prll[0] = gx; prll[1] = gy; }
  orth[0] = -prll[1];
  orth[1] = prll[0];
}

/* GA3CCADRLPolicy.py:45-106 with the id dropped (:38): x[75] doubles. */
static void ga3c_state(const casim_config* c, const double* env, int n, int i, double* x) {
  const double* h = env + i * NF;
  memset(x, 0, sizeof(double) * NN_IN);
  x[1] = h[CASIM_DIST_TO_GOAL];
  x[2] = h[CASIM_HEADING_EGO];
  x[3] = h[CASIM_PREF_SPEED];
  x[4] = h[CASIM_RADIUS];
  double prll[2], orth[2];
  ref_axes(h, prll, orth);
  long long kr[CASIM_MAX_AGENTS];
  double po[CASIM_MAX_AGENTS];
  int idx[CASIM_MAX_AGENTS], m = 0;
  for (int j = 0; j < n; j++) {
    if (j == i) continue;
    const double* o = env + j * NF;
    double rx = o[CASIM_PX] - h[CASIM_PX], ry = o[CASIM_PY] - h[CASIM_PY];
    double p_orth = rx * orth[0] + ry * orth[1];
    double dc = sqrt(rx * rx + ry * ry);
    double d2o = dc - h[CASIM_RADIUS] - o[CASIM_RADIUS];
    if (dc > c->sensing_horizon) continue;
    kr[m] = py_round2_hundredths(d2o);
    po[m] = p_orth;
    idx[m] = j;
    m++;
  }
  /* sorted(key = (-round(dist, 2), p_orthog)), stable (:66) */
  for (int a = 1; a < m; a++) {
    long long k = kr[a];
    double p = po[a];
    int id = idx[a], b = a - 1;
    while (b >= 0 && (-kr[b] > -k || (-kr[b] == -k && po[b] > p))) {
      kr[b + 1] = kr[b]; po[b + 1] = po[b]; idx[b + 1] = idx[b];
      b--;
    }
    kr[b + 1] = k; po[b + 1] = p; idx[b + 1] = id;
  }
  /* keep the LAST Config.MAX_NUM_OTHER_AGENTS_OBSERVED (:68); the network has 10 slots */
  int K = m < c->max_other_agents_observed ? m : c->max_other_agents_observed;
  if (K > NN_SLOTS) K = NN_SLOTS;
  for (int r = 0; r < K; r++) {
    const double* o = env + idx[m - K + r] * NF;
    double rx = o[CASIM_PX] - h[CASIM_PX], ry = o[CASIM_PY] - h[CASIM_PY];
    double* w = x + 1 + NN_HOST + NN_OTHER * r;
    w[0] = rx * prll[0] + ry * prll[1];
    w[1] = rx * orth[0] + ry * orth[1];
    w[2] = o[CASIM_VX] * prll[0] + o[CASIM_VY] * prll[1];
    w[3] = o[CASIM_VX] * orth[0] + o[CASIM_VY] * orth[1];
    w[4] = o[CASIM_RADIUS];
    w[5] = h[CASIM_RADIUS] + o[CASIM_RADIUS];
    w[6] = sqrt(rx * rx + ry * ry) - h[CASIM_RADIUS] - o[CASIM_RADIUS];
  }
  x[0] = (double)K;
}

/* network.py:125-128,147-148 normalisation vectors */
static const float HOST_AVG[4] = {0.0f, 0.0f, 1.0f, 0.5f}, HOST_STD[4] = {5.0f, 3.14f, 1.0f, 1.0f};
static const float OTH_AVG[7] = {0.0f, 0.0f, 0.0f, 0.0f, 0.5f, 0.0f, 1.0f}, OTH_STD[7] = {5.0f, 5.0f, 1.0f, 1.0f, 1.0f, 5.0f, 1.0f};

static float sigmoidf_(float v) { return 1.0f / (1.0f + expf(-v)); }

/* offsets into the parameter blob (include/casim.h: casim_ga3c_set_weights) */
#define W_LSTM_K 0
#define W_LSTM_B (W_LSTM_K + (NN_OTHER + NN_H) * 4 * NN_H)
#define W_L1_K (W_LSTM_B + 4 * NN_H)
#define W_L1_B (W_L1_K + (NN_HOST + NN_H) * NN_FC)
#define W_L2_K (W_L1_B + NN_FC)
#define W_L2_B (W_L2_K + NN_FC * NN_FC)
#define W_FC_K (W_L2_B + NN_FC)
#define W_FC_B (W_FC_K + NN_FC * NN_FC)
#define W_P_K (W_FC_B + NN_FC)
#define W_P_B (W_P_K + NN_FC * NN_ACT)
#define W_V_K (W_P_B + NN_ACT)
#define W_END (W_V_K + NN_FC + 1)

static void dense(const float* in, int nin, const float* k, const float* b, int nout, int relu, float* out) {
  for (int o = 0; o < nout; o++) {
    float acc = 0.0f;
    for (int i = 0; i < nin; i++) acc += in[i] * k[i * nout + o];
    acc += b[o];
    out[o] = relu && acc < 0.0f ? 0.0f : acc;
  }
}

/* network.py:65-98 + :45-51: x [75] fp32 -> softmax_p [11] */
static void ga3c_network(const float* w, const float* x, float* probs) {
  float host[NN_HOST], hc[NN_H], cc[NN_H], in[NN_OTHER + NN_H], gates[4 * NN_H];
  for (int k = 0; k < NN_HOST; k++) host[k] = (x[1 + k] - HOST_AVG[k]) / HOST_STD[k];
  memset(hc, 0, sizeof hc);
  memset(cc, 0, sizeof cc);
  int len = (int)x[0]; /* sequence_length = x[:, 0], un-normalised (network.py:85) */
  if (len > NN_SLOTS) len = NN_SLOTS;
  for (int t = 0; t < len; t++) {
    for (int k = 0; k < NN_OTHER; k++) in[k] = (x[1 + NN_HOST + NN_OTHER * t + k] - OTH_AVG[k]) / OTH_STD[k];
    memcpy(in + NN_OTHER, hc, sizeof hc);
    dense(in, NN_OTHER + NN_H, w + W_LSTM_K, w + W_LSTM_B, 4 * NN_H, 0, gates);
    for (int u = 0; u < NN_H; u++) { /* TF1 LSTMCell: i, j, f, o; forget_bias = 1.0 */
      float gi = gates[u], gj = gates[NN_H + u], gf = gates[2 * NN_H + u], go = gates[3 * NN_H + u];
      float cn = sigmoidf_(gf + 1.0f) * cc[u] + sigmoidf_(gi) * tanhf(gj);
      cc[u] = cn;
      hc[u] = sigmoidf_(go) * tanhf(cn);
    }
  }
  float l1in[NN_HOST + NN_H], a1[NN_FC], a2[NN_FC], a3[NN_FC], logits[NN_ACT];
  memcpy(l1in, host, sizeof host);
  memcpy(l1in + NN_HOST, hc, sizeof hc);
  dense(l1in, NN_HOST + NN_H, w + W_L1_K, w + W_L1_B, NN_FC, 1, a1);
  dense(a1, NN_FC, w + W_L2_K, w + W_L2_B, NN_FC, 1, a2);
  dense(a2, NN_FC, w + W_FC_K, w + W_FC_B, NN_FC, 1, a3);
  dense(a3, NN_FC, w + W_P_K, w + W_P_B, NN_ACT, 0, logits);
  float mx = logits[0], sum = 0.0f;
  for (int k = 1; k < NN_ACT; k++) if (logits[k] > mx) mx = logits[k];
  for (int k = 0; k < NN_ACT; k++) { probs[k] = expf(logits[k] - mx); sum += probs[k]; }
  for (int k = 0; k < NN_ACT; k++) probs[k] = (probs[k] / sum + 1e-4f) / (1.0f + 1e-4f * NN_ACT);
}

typedef struct {
  const casim_config* c; int e0, e1, A; const int32_t* n_agents; const double* state; const float* w;
  double* policy_actions; int32_t* action_idx; float* probs; double* nn_input;
} ga3c_job;

static void* ga3c_job_main(void* pj) {
  ga3c_job* j = (ga3c_job*)pj;
  double table[NN_ACT * 2];
  casim_oracle_ga3c_actions(table);
  for (int e = j->e0; e < j->e1; e++) {
    int n = j->n_agents ? j->n_agents[e] : j->A;
    const double* env = j->state + (size_t)e * j->A * NF;
    for (int i = 0; i < j->A; i++) {
      size_t g = (size_t)e * j->A + i;
      if (j->action_idx) j->action_idx[g] = -1;
      if (i >= n) continue;
      const double* s = env + i * NF;
      if (((int)s[CASIM_TYPE] & 0xff) != CASIM_POLICY_GA3C_CADRL) continue;
      double x[NN_IN];
      ga3c_state(j->c, env, n, i, x);
      if (j->nn_input) memcpy(j->nn_input + g * NN_IN, x, sizeof x);
      if ((int)s[CASIM_FLAGS] & CASIM_F_IS_DONE) { /* collision_avoidance_env.py:298-300 */
        j->policy_actions[2 * g] = 0; j->policy_actions[2 * g + 1] = 0;
        continue;
      }
      float xf[NN_IN], pr[NN_ACT];
      for (int k = 0; k < NN_IN; k++) xf[k] = (float)x[k]; /* feed_dict -> float32 placeholder */
      ga3c_network(j->w, xf, pr);
      int best = 0;
      for (int k = 1; k < NN_ACT; k++) if (pr[k] > pr[best]) best = k;
      if (j->probs) memcpy(j->probs + g * NN_ACT, pr, sizeof pr);
      if (j->action_idx) j->action_idx[g] = best;
      j->policy_actions[2 * g] = s[CASIM_PREF_SPEED] * table[2 * best]; /* GA3CCADRLPolicy.py:40-42 */
      j->policy_actions[2 * g + 1] = table[2 * best + 1];
    }
  }
  return NULL;
}

int casim_oracle_ga3c_forward(const casim_config* c, int E, int A, const int32_t* n_agents, const double* state,
                              const float* weights, int n_params, double* policy_actions, int32_t* action_idx,
                              float* probs, double* nn_input, int n_threads) {
  if (n_params != CASIM_GA3C_NUM_PARAMS || W_END != CASIM_GA3C_NUM_PARAMS) return CASIM_EINVAL;
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 64) n_threads = 64;
  ga3c_job jobs[64];
  pthread_t th[64];
  for (int t = 0; t < n_threads; t++) {
    ga3c_job* j = &jobs[t];
    j->c = c; j->A = A; j->n_agents = n_agents; j->state = state; j->w = weights;
    j->policy_actions = policy_actions; j->action_idx = action_idx; j->probs = probs; j->nn_input = nn_input;
    j->e0 = (int)((long long)E * t / n_threads);
    j->e1 = (int)((long long)E * (t + 1) / n_threads);
    if (n_threads == 1) ga3c_job_main(j); else pthread_create(&th[t], NULL, ga3c_job_main, j);
  }
  if (n_threads > 1) for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
  return 0;
}

/* network forward alone on ready-made inputs (cross-check against torch in the tests) */
int casim_oracle_ga3c_network(const float* weights, const float* x, int B, float* probs) {
  for (int b = 0; b < B; b++) ga3c_network(weights, x + (size_t)b * NN_IN, probs + (size_t)b * NN_ACT);
  return 0;
}
