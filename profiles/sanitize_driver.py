#!/usr/bin/env python
"""One small invocation of every per-step kernel (the workloads of bench.py at 48 envs) for compute-sanitizer:
    compute-sanitizer --tool memcheck  python profiles/sanitize_driver.py
    compute-sanitizer --tool racecheck python profiles/sanitize_driver.py
(racecheck / memcheck on the shuffle + shared-memory kernels)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gym-exploration-2d_b200"))

import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from casim import lib  # noqa: E402
from casim.sim import BatchedSim  # noqa: E402

which = sys.argv[1:] or ["c2", "full", "full_permap", "c4", "c5", "c5tc", "rvo"]
for name in which:
    wl = {"full_permap": "full", "c5tc": "c5"}.get(name, name)
    W = bench.WORKLOADS[wl]
    E, A = 48, W["A"]
    cfg = lib.default_config()
    cfg.auto_reset = 1
    sim = BatchedSim(E, A, cfg=cfg, sensors=W["sensors"])
    init, types, n, rects = bench.scenario(wl, E, A, seed=7)
    if rects is not None:
        if name == "full_permap":
            from casim.scenarios import per_env_crossing_rects
            sim.set_obstacles(per_env_crossing_rects(E, seed=3))
        else:
            sim.set_obstacles(rects)
    if wl == "c4":
        sim.enable_exploration()
    if wl == "c5":
        sim.ga3c_set_weights(bench.ga3c_weights())
    if wl == "full":
        sim.enable_packed_grid()
    sim.reset(init, types, n)
    for t in range(3):
        pol = None
        if wl == "c4":
            pol = sim.explore_step()
        elif wl == "c5":
            pol = sim.ga3c_forward(use_tensor_cores=(name == "c5tc"))
        elif wl == "rvo":
            pol = sim.rvo_step()
        sim.step(policy_actions=pol)
    if wl == "c4":
        acts = np.random.default_rng(0).integers(0, 9, (E, A, 6, 10)).astype(np.uint8)
        sim.explore_rollouts(acts)
        sim.explore_team_update()
    if wl == "full":
        sim.flatten_obs(["dist_to_goal", "other_agents_states", "laserscan", "local_grid"])
        sim.reset_envs(sim.buf["game_over"])
    torch.cuda.synchronize()
    print(name, "ok, rewards sum %.5f" % float(sim.buf["rewards"].sum()))
    sim.close()
