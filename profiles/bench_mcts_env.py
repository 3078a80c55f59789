"""Side benchmark: one ENVIRONMENT step with ig_mcts explorers (collision_avoidance_env.py:341-379 _take_action_dmcts + the
step itself) for a batch of C4-like environments -- team belief update, Ncycles x explorers-per-env planning rounds of
Ntree Dec-MCTS iterations each with the action distributions exchanged in between, then the fused step.
  python profiles/bench_mcts_env.py [--E 1024] [--ntree 100] [--nsims 10] [--H 10] [--ncycles 5] [--steps 3] [--host]
Wall clock per env.step(); the reference's defaults are Ntree 100, Nsims 10, horizon 10, Ncycles 5 (ig_mcts.py:30-60)."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gym-exploration-2d_b200"))
import numpy as np
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--E", type=int, default=1024)
    ap.add_argument("--ntree", type=int, default=100)
    ap.add_argument("--nsims", type=int, default=10)
    ap.add_argument("--H", type=int, default=10)
    ap.add_argument("--ncycles", type=int, default=5)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--host", action="store_true", help="host tree bookkeeping (BatchedDecMCTS) instead of the device trees")
    a = ap.parse_args()
    from casim import plugins as P
    from casim.env import BatchedCollisionAvoidanceEnv
    from casim.config import Config

    class Cfg(Config):
        pass

    rng = np.random.default_rng(1)
    envs_agents = []
    for e in range(a.E):
        x = float(rng.uniform(-9, -5))
        y = float(rng.uniform(-9, -5))
        ags = [P.Agent(x, 0.4, 9, 0, 0.5, 1.0, 0.0, P.ig_mcts, P.FirstOrderDynamics, [P.OtherAgentsStatesSensor], 0),
               P.Agent(0.3, y, 0, 9, 0.5, 1.0, 1.5, P.ig_mcts, P.FirstOrderDynamics, [P.OtherAgentsStatesSensor], 1),
               P.Agent(x + 3.0, -0.3, 0, 0, 0.2, 1.0, 0.0, P.StaticPolicy, P.UnicycleDynamics, [], 2),
               P.Agent(0.0, -3.0, 0, 0, 0.2, 1.0, 0.0, P.StaticPolicy, P.UnicycleDynamics, [], 3)]
        for ag in ags[:2]:
            ag.policy.set_param(Ntree=a.ntree, Nsims=a.nsims, mcts_horizon=a.H, Ncycles=a.ncycles, mcts_gamma=1)
            if a.host:
                ag.policy.device_trees = False
        envs_agents.append(ags)
    crossing = [[(10, 10), (2, 10), (2, 2), (10, 2)], [(-2, 10), (-10, 10), (-10, 2), (-2, 2)],
                [(10, -2), (2, -2), (2, -10), (10, -10)], [(-2, -2), (-10, -2), (-10, -10), (-2, -10)]]
    env = BatchedCollisionAvoidanceEnv(num_envs=a.E, max_agents=4, config=Cfg)
    env.set_agents(envs_agents)
    env.set_static_map(crossing)
    env.reset()
    env.step()                                   # warm-up: allocator blocks, sort kernels, lazy module loads
    torch.cuda.synchronize()
    times = []
    for _ in range(a.steps):
        t0 = time.perf_counter()
        env.step()
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
    iters = 2 * a.E * a.ncycles * a.ntree
    print(json.dumps({"metric": "seconds per env.step with ig_mcts explorers", "envs": a.E, "explorers": 2 * a.E,
                      "ntree": a.ntree, "nsims": a.nsims, "horizon": a.H, "ncycles": a.ncycles,
                      "planner": "host" if a.host else "device", "seconds_per_step": float(np.median(times)),
                      "all": times, "tree_iterations_per_s": iters / float(np.median(times)),
                      "simulations_per_s": iters * a.nsims / float(np.median(times))}))
    env.close()


if __name__ == "__main__":
    main()
