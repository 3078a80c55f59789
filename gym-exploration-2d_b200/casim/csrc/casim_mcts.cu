/*
 * casim_mcts.cu -- Dec-MCTS tree bookkeeping on the device: the reference's
 * `Tree.grow` (policies/pydecmcts/DecMCTS.py:273-360) for T trees in lock-step, one node per tree per iteration,
 * with the node arrays resident in HBM.  One iteration = three launches on one stream, no host round trip:
 *
 *   k_mcts_select   _get_system_state (:182-194) -> the cells the other robots claim;  _select (:140-153) with _UCT
 *                   (:14-18) from the root down to a childless node;  the action sequences of that node's 9 one-step
 *                   extensions (expansion feasibility, _expansion :201-231) and of its `nsims` random completions
 *                   (_simulate :233-269), written straight into casim_explore_rollouts' [E,A,K,H] action buffer;
 *   k_rollouts      (casim_explore.cu) evaluates them;
 *   k_mcts_update   _expansion: a child per feasible action, in action order;  the node's statistics (:323-339);
 *                   discounted back-propagation to the root (:341-357).
 *
 * One warp per tree.  Node numbering follows the reference (creation order), the arithmetic is fp64 without FMA
 * contraction in the reference's evaluation order, so a tree grown here equals the tree casim/mcts.py's host planner
 * grows from the same random draws, node for node (tests/test_gpu_mcts.py).  The draws themselves (np.random in the
 * reference, ig_mcts.py:257 and DecMCTS.py:190) come from the caller.
 */
#include "casim_internal.h"

#include <math_constants.h>

namespace {

constexpr int MC_ACT = CASIM_MCTS_ACTIONS; /* 9: vel_list x dphi_list, ig_mcts.py:259-265 */
constexpr int MC_END = 255;
constexpr int MC_WARPS = 4; /* trees per CTA */

/* Counter-based stand-in for np.random.choice(9) (ig_mcts.py:257) when the caller passes no draws: splitmix64 of
 * (seed, global draw index), top 32 bits scaled to 0..8.  casim/mcts.py:hash_draws is the same function in numpy. */
__device__ __forceinline__ int mcts_draw(unsigned long long seed, unsigned long long index) {
  unsigned long long x = seed + 0x9E3779B97F4A7C15ull * (index + 1ull);
  x ^= x >> 30;
  x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27;
  x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return (int)(((x >> 32) * 9ull) >> 32);
}

__device__ __forceinline__ size_t nidx(const casim_mcts_trees& t, int tree, int node) { return (size_t)tree * t.M + node; }

__global__ void __launch_bounds__(32 * MC_WARPS)
    k_mcts_select(const casim_mcts_trees t, const uint8_t* __restrict__ draws, unsigned long long draw_seed, int iteration,
                  const double* __restrict__ comm_u, uint8_t* __restrict__ actions, uint32_t* __restrict__ exclude,
                  int root_only) {
  const int tree = blockIdx.x * MC_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (tree >= t.T) return;
  const int H = t.H, K = MC_ACT + t.nsims;
  const size_t slot = (size_t)t.slot[tree];

  /* ---- the other robots' claimed cells: one weighted-random entry of every valid distribution, OR-ed ---- */
  if (exclude) {
    uint32_t acc[4] = {0u, 0u, 0u, 0u};
    for (int cm = 0; cm < t.n_comms; cm++) {
      const size_t ct = (size_t)cm * t.T + tree;
      if (!t.comm_valid[ct]) continue;
      const double u = comm_u[ct];
      const double* q = t.comm_q + ct * t.comm_n;
      double cum = 0.0;
      int pick = 0;
      for (int i = 0; i < t.comm_n; i++) { /* (u > cumsum(q)).sum(), clipped: np.random.choice(n, p=q) */
        cum = cum + q[i];
        pick += (u > cum) ? 1 : 0;
      }
      pick = min(pick, t.comm_n - 1);
      const uint32_t* m = t.comm_masks + (ct * t.comm_n + pick) * 128;
#pragma unroll
      for (int w = 0; w < 4; w++) acc[w] |= m[lane + 32 * w];
    }
#pragma unroll
    for (int w = 0; w < 4; w++) exclude[slot * 128 + lane + 32 * w] = acc[w];
  }

  /* ---- selection: descend by UCT while the node has children ---- */
  int node = 0;
  if (!root_only) {
    for (;;) {
      const size_t ni = nidx(t, tree, node);
      if (t.nchild[ni] == 0) break;
      const int ch = lane < MC_ACT ? t.child[ni * MC_ACT + lane] : -1;
      double val = -CUDART_INF;
      if (ch >= 0) {
        const size_t ci = nidx(t, tree, ch);
        const double n_j = t.N[ci], mu_j = t.mu[ci], n_p = t.N[ni];
        /* mu_j + 2 * c_p * (2 * log(n_p) / n_j) ** 0.5 ; unvisited children first (:147-150) */
        val = n_j == 0.0 ? CUDART_INF : mu_j + (2.0 * t.c_p) * sqrt((2.0 * log(n_p)) / n_j);
      }
      int best = lane;
#pragma unroll
      for (int d = 16; d >= 1; d >>= 1) { /* first maximum, like np.argmax */
        const double ov = __shfl_xor_sync(0xffffffffu, val, d);
        const int ob = __shfl_xor_sync(0xffffffffu, best, d);
        if (ov > val || (ov == val && ob < best)) {
          val = ov;
          best = ob;
        }
      }
      node = __shfl_sync(0xffffffffu, ch, best);
    }
  }
  if (lane == 0) t.node[tree] = node;

  /* ---- the K sequences of this node ---- */
  const size_t ni = nidx(t, tree, node);
  const int stage = t.stage[ni];
  const uint8_t* path = t.path + ni * H;
  uint8_t* out = actions + slot * (size_t)K * H;
  for (int idx = lane; idx < K * H; idx += 32) {
    const int k = idx / H, d = idx - k * H;
    int v = path[d];
    if (k < MC_ACT) { /* expansion probe: path + action k, feasibility only; nothing to probe at the horizon */
      if (stage >= H) v = MC_END;
      else if (d == stage) v = k;
      else if (d == stage + 1) v = CASIM_ROLLOUT_PROBE;
    } else if (root_only) {
      v = MC_END;
    } else if (d >= stage) {
      const size_t di = ((size_t)tree * t.nsims + (k - MC_ACT)) * H + d;
      v = draws ? draws[di] : mcts_draw(draw_seed, (unsigned long long)iteration * t.T * t.nsims * H + di);
    }
    out[idx] = (uint8_t)v;
  }
}

__global__ void __launch_bounds__(32 * MC_WARPS)
    k_mcts_update(const casim_mcts_trees t, const uint8_t* __restrict__ actions, const float* __restrict__ rewards,
                  const int32_t* __restrict__ n_feasible, double* __restrict__ avg_out, int expand_only,
                  int* __restrict__ overflow) {
  const int tree = blockIdx.x * MC_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (tree >= t.T) return;
  const int H = t.H, K = MC_ACT + t.nsims;
  const size_t slot = (size_t)t.slot[tree];
  const int node = t.node[tree];
  const size_t ni = nidx(t, tree, node);
  const int stage = t.stage[ni];

  /* ---- expansion: a child per feasible action, in action order; nothing at stage == horizon ---- */
  const bool feas = lane < MC_ACT && stage < H && n_feasible[slot * K + lane] == stage + 1;
  const unsigned fm = __ballot_sync(0xffffffffu, feas);
  const int cnt = __popc(fm), base = t.n_nodes[tree];
  if (cnt > 0) {
    if (base + cnt > t.M) { /* the caller under-sized the node arrays: refuse rather than overwrite */
      if (lane == 0) atomicExch(overflow, 1);
      return;
    }
    if (feas) {
      const int k = __popc(fm & ((1u << lane) - 1u)), nw = base + k;
      const size_t wi = nidx(t, tree, nw);
      t.parent[wi] = node;
      t.stage[wi] = (uint8_t)(stage + 1);
      t.child[ni * MC_ACT + k] = nw;
    }
    for (int k = 0; k < cnt; k++) { /* path of child k = the node's path + its action */
      const int a = __fns(fm, 0, k + 1);
      uint8_t* dst = t.path + nidx(t, tree, base + k) * H;
      if (lane < H) dst[lane] = lane == stage ? (uint8_t)a : t.path[ni * H + lane];
    }
    if (lane == 0) {
      t.nchild[ni] = (uint8_t)(t.nchild[ni] + cnt);
      t.n_nodes[tree] = base + cnt;
    }
  }
  if (expand_only) return;

  /* ---- the node's statistics: average accumulated left to right, first strict maximum (:323-339) ---- */
  const float* rw = rewards + slot * K + MC_ACT;
  double acc = 0.0, best = (double)rw[0];
  int kb = 0;
  for (int i = 0; i < t.nsims; i++) {
    const double r = (double)rw[i];
    acc = acc + r;
    if (r > best) {
      best = r;
      kb = i;
    }
  }
  const double avg = acc / (double)t.nsims;
  const uint8_t seq = lane < H ? actions[(slot * K + MC_ACT + kb) * (size_t)H + lane] : (uint8_t)MC_END;
  if (lane == 0) {
    t.mu[ni] = avg;
    t.best_reward[ni] = best;
    t.N[ni] = 1.0;
    t.has_rollout[ni] = 1;
    if (avg_out) avg_out[tree] = avg;
  }
  if (lane < H) t.best_seq[ni * H + lane] = seq;

  /* ---- discounted back-propagation (:341-357) ---- */
  int cur = node;
  for (;;) {
    const int up = t.parent[nidx(t, tree, cur)];
    if (up < 0) break;
    const size_t ui = nidx(t, tree, up);
    const double n_old = t.N[ui], mu_old = t.mu[ui];
    const bool better = best > t.best_reward[ui];
    __syncwarp();
    if (lane == 0) {
      t.mu[ui] = ((t.gamma * mu_old) * n_old + avg) / (n_old + 1.0);
      t.N[ui] = t.gamma * n_old + 1.0;
      if (better) {
        t.best_reward[ui] = best;
        t.has_rollout[ui] = 1;
      }
    }
    if (better && lane < H) t.best_seq[ui * H + lane] = seq;
    cur = up;
  }
}

}  // namespace

cudaError_t casim_launch_mcts_select(casim_handle* h, const casim_mcts_trees* t, const uint8_t* draws,
                                     unsigned long long draw_seed, int iteration, const double* comm_u, uint8_t* actions,
                                     uint32_t* exclude, int root_only, cudaStream_t s) {
  h->launches++;
  k_mcts_select<<<(t->T + MC_WARPS - 1) / MC_WARPS, 32 * MC_WARPS, 0, s>>>(*t, draws, draw_seed, iteration, comm_u, actions,
                                                                         exclude, root_only);
  return cudaGetLastError();
}

cudaError_t casim_launch_mcts_update(casim_handle* h, const casim_mcts_trees* t, const uint8_t* actions, const float* rewards,
                                     const int32_t* n_feasible, double* avg_out, int expand_only, int* overflow,
                                     cudaStream_t s) {
  h->launches++;
  k_mcts_update<<<(t->T + MC_WARPS - 1) / MC_WARPS, 32 * MC_WARPS, 0, s>>>(*t, actions, rewards, n_feasible, avg_out,
                                                                         expand_only, overflow);
  return cudaGetLastError();
}
