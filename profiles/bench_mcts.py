"""Side benchmark: one Dec-MCTS planning step (ig_mcts.find_next_action's tree growth, ig_mcts.py:79-109) for every
explorer of the C4 scene with casim/mcts.py + casim_explore_rollouts: Ntree iterations x Nsims simulations x horizon.
  python profiles/bench_mcts.py [--ntree 100] [--nsims 10] [--H 10]
Wall clock of the whole planning step (draws + launches + the final D2H of the node arrays + the action distribution).
Default: the device-resident trees (DeviceDecMCTS: casim_mcts_select -> casim_explore_rollouts -> casim_mcts_update per
iteration); --host times the host bookkeeping planner (BatchedDecMCTS) the same way."""
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
    ap.add_argument("--host", action="store_true")
    a = ap.parse_args()
    from casim import lib, scenarios as S
    from casim.sim import BatchedSim
    from casim.mcts import BatchedDecMCTS, DeviceDecMCTS, sim_evaluator
    init, types, n, rects = S.exploration(a.E, seed=99)
    cfg = lib.default_config()
    cfg.game_over_mode = 0
    sim = BatchedSim(a.E, 4, cfg=cfg, sensors=("other_agents_states",))
    sim.enable_exploration()
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    for _ in range(3):
        sim.explore_step()
        sim.step(policy_actions=sim.buf["policy_actions"])
    T, ev, masks_of, _ = sim_evaluator(sim)                # _ = (env index, agent index) of the trees
    t_eval = [0.0]

    def timed(actions, exclude):
        t = time.perf_counter()
        out = ev(actions, exclude)
        t_eval[0] += time.perf_counter() - t
        return out

    sel = np.zeros((a.E, 4), bool)
    sel[_[0], _[1]] = True
    if a.host:
        BatchedDecMCTS(T, ev, horizon=a.H).grow(nsims=a.nsims, gamma=1.0)     # warm-up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        tree = BatchedDecMCTS(T, timed, horizon=a.H, c_p=1.0, seed=1)
        for _i in range(a.ntree):
            tree.grow(nsims=a.nsims, gamma=1.0)
        act = tree.best_action_index()
        dt = time.perf_counter() - t0
        dev_s = None
    else:
        w = DeviceDecMCTS(sim, sel, horizon=a.H, nsims=a.nsims, n_grows=a.ntree)     # warm-up: allocator blocks, sort kernels
        w.grow(nsims=a.nsims, gamma=1.0)
        w.best_action_index()
        del w
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        tree = DeviceDecMCTS(sim, sel, horizon=a.H, c_p=1.0, seed=1, nsims=a.nsims, n_grows=a.ntree)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        tree.grow_n(a.ntree, gamma=1.0)
        ev1.record()
        act = tree.best_action_index()
        dt = time.perf_counter() - t0
        dev_s = ev0.elapsed_time(ev1) * 1e-3
        t_eval[0] = float("nan")
    sims = T * a.ntree * a.nsims
    print(json.dumps({"metric": "planning steps/s (trees x grow iterations)", "trees": T, "ntree": a.ntree, "nsims": a.nsims,
                      "horizon": a.H, "seconds_per_planning_step_all_trees": dt, "planner": "host" if a.host else "device",
                      "evaluator_seconds": t_eval[0] if a.host else None,
                      "host_bookkeeping_seconds": dt - t_eval[0] if a.host else None,
                      "device_seconds_grow_loop": dev_s, "simulations_per_s": sims / dt,
                      "tree_iterations_per_s": T * a.ntree / dt, "gpu_launches": tree.launches,
                      "nodes_per_tree_mean": float(tree.d["n_nodes"].double().mean().item()) if not a.host else float(tree.n_nodes.mean()), "actions_hist": np.bincount(act, minlength=9).tolist()}))


if __name__ == "__main__":
    main()
