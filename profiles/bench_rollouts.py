"""Side benchmark (not the bench.py contract line): the Dec-MCTS rollout evaluator casim_explore_rollouts on the
C4 scene -- 1024 envs x (2 explorers + 2 targets) x K rollouts x H steps -- with the CPU oracle
(oracle/casim_oracle_explore.c, one thread) timed on a bounded sample of the same agents beside it.
  python profiles/bench_rollouts.py [--K 10] [--H 10] [--iters 20] [--no-cpu]
Prints one JSON line.  Inputs resident in HBM; CUDA events on the launching stream; the action tensor
(E*A*K*H bytes) and the belief (E*A*14.4 KB) exceed nothing: the working set is L2 resident by nature of the
planner (it re-reads one belief many times), which is the configuration the planner runs in."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gym-exploration-2d_b200"))
sys.path.insert(0, ROOT)
import numpy as np
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--E", type=int, default=1024)
    ap.add_argument("--K", type=int, default=10)
    ap.add_argument("--H", type=int, default=10)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    from casim import lib, layout as L, scenarios as S
    from casim.sim import BatchedSim
    E, K, Hn = a.E, a.K, a.H
    init, types, n, rects = S.exploration(E, seed=99)
    cfg = lib.default_config()
    cfg.game_over_mode = 0
    sim = BatchedSim(E, 4, cfg=cfg, sensors=("other_agents_states",))
    sim.enable_exploration()
    sim.set_obstacles(rects)
    sim.reset(init, types, n)
    for _ in range(3):
        sim.explore_step()
        sim.step(policy_actions=sim.buf["policy_actions"])
    rng = np.random.default_rng(3)
    acts_np = rng.integers(0, 9, (E, 4, K, Hn)).astype(np.uint8)
    actions = torch.from_numpy(acts_np).cuda()
    n_expl = int(((types & 0xff) == L.POLICY_IG_GREEDY).sum())
    res = {}
    for want in (True, False):
        for _ in range(5):
            out = sim.explore_rollouts(actions, want_end_pose=want, want_obsvd=want)
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(a.iters):
            out = sim.explore_rollouts(actions, want_end_pose=want, want_obsvd=want)
        t1.record()
        torch.cuda.synchronize()
        res[want] = t0.elapsed_time(t1) / a.iters * 1e3
    line = {"metric": "rollout-steps/s", "unit": "rollout-steps/s", "value": n_expl * K * Hn / (res[False] * 1e-6),
            "us_per_launch": res[False], "us_per_launch_with_masks": res[True],
            "config": {"workload": "C4 scene, %d envs x 4 agents (%d explorers) x K=%d rollouts x H=%d steps" % (E, n_expl, K, Hn)},
            "gpu_launches": a.iters, "dtype": "f64 decisions / fp32 belief", "data": "synthetic"}
    if not a.no_cpu:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from oracle import oracle
        ocfg = oracle.default_config()
        ocfg.game_over_mode = 0
        smap = oracle.rasterize(ocfg, rects.astype(np.float64))
        state = sim.state_numpy()
        belief = sim.belief.cpu().numpy().astype(np.float64)
        rew = out["rewards"].cpu().numpy()
        ig = np.argwhere((types & 0xff) == L.POLICY_IG_GREEDY)
        sample = ig[:64]
        t = time.perf_counter()
        worst = 0.0
        for e, ag in sample:
            pose = state[e, ag, [L.PX, L.PY, L.HEADING]]
            o = oracle.explore_rollouts(ocfg, smap, pose, state[e, ag, L.RADIUS], belief[e, ag], acts_np[e, ag])
            worst = max(worst, float(np.abs(o["rewards"] - rew[e, ag]).max()))
        dt = time.perf_counter() - t
        t = time.perf_counter()
        for _ in range(len(sample)):
            oracle.edt(smap)
        t_edt = time.perf_counter() - t          # every oracle call re-derives the map's EDT: not charged to the baseline
        dt = max(dt - t_edt, 1e-9)
        line["cpu_baseline"] = {"value": len(sample) * K * Hn / dt, "unit": "rollout-steps/s", "cores": 1, "kind": "port",
                                "sample": "%d explorer agents x %d x %d (EDT rebuild time subtracted)" % (len(sample), K, Hn)}
        line["max_abs_reward_diff_vs_oracle"] = worst
    print(json.dumps(line))


if __name__ == "__main__":
    main()
