"""Tuning aid: phase timeline of k_ga3c_tc (CASIM_GA3C_TRACE) at the C5 shape."""
import os, sys
import numpy as np
os.environ["CASIM_GA3C_TRACE"] = "/tmp/ga3c_trace.bin"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gym-exploration-2d_b200")]
import torch
import bench
from casim import lib
from casim.sim import BatchedSim
E, A = 1024, 20
cfg = lib.default_config(); cfg.near_goal_threshold = 0.2
init, types, n, rects = bench.scenario("c5", E, A, seed=1234)
sim = BatchedSim(E, A, cfg=cfg, sensors=("other_agents_states",), want_debug=False)
sim.ga3c_set_weights(bench.ga3c_weights(), want_debug=False)
sim.reset(init, types, n)
for _ in range(4):
    sim.ga3c_forward(use_tensor_cores=True)
torch.cuda.synchronize()
t = np.fromfile("/tmp/ga3c_trace.bin", dtype=np.uint64).reshape(-1, 2, 64).astype(np.int64)
t0 = t[:, 1, 0].min()
for cta in (0, 71, 142):
    ep, mm = t[cta, 0], t[cta, 1]
    print("CTA", cta, "mma start", mm[0] - t0, "lstm weights at", mm[1] - t0)
    print("  mma :", [int(x - t0) for x in mm[2:] if x])
    print("  epi :", [int(x - t0) for x in ep[:62] if x], "end", int(ep[63] - t0))
ends = t[:, 0, 63] - t0
print("end min/mean/max", ends.min(), ends.mean(), ends.max())
