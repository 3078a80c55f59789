#!/usr/bin/env python
"""How often does the bf16 tensor-core forward pick the action the fp32 forward (== the fp64 evaluation of the reference's
TF graph on every golden vector) picks?  Trained IROS18 weights; formation scenario states over a few steps, and the
reference-generated golden state vectors.  Usage (GPU box): python profiles/ga3c_agreement.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gym-exploration-2d_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import helpers as H  # noqa: E402
from casim import lib, scenarios as S  # noqa: E402
from casim.sim import BatchedSim  # noqa: E402

w = H.load("ga3c_iros18.npz")["weights"]
cfg = lib.default_config()
cfg.game_over_mode = 1
cfg.near_goal_threshold = 0.2
E, A = 1024, 20
init, types, n = S.formations(E, A, seed=11)
sim = BatchedSim(E, A, cfg=cfg, sensors=("other_agents_states",))
sim.reset(init, types, n)
sim.ga3c_set_weights(w)
tot = same = 0
maxdp = 0.0
hist = []
low = {0.02: 0, 0.05: 0, 0.10: 0}
for t in range(8):
    sim.ga3c_forward(use_tensor_cores=False)
    torch.cuda.synchronize()
    p32, i32 = sim.buf["probs"].clone(), sim.buf["action_idx"].clone()
    a32 = sim.buf["policy_actions"].clone()
    sim.ga3c_forward(use_tensor_cores=True)
    torch.cuda.synchronize()
    ptc, itc = sim.buf["probs"], sim.buf["action_idx"]
    live = i32 >= 0
    tot += int(live.sum())
    same += int((itc[live] == i32[live]).sum())
    maxdp = max(maxdp, float((ptc[live] - p32[live]).abs().max()))
    srt = torch.sort(p32[live], dim=-1).values
    margin = srt[:, -1] - srt[:, -2]
    srt_tc = torch.sort(ptc[live], dim=-1).values
    for thr in low:                                   # what the exact mode would re-evaluate at this margin threshold
        low[thr] += int(((srt_tc[:, -1] - srt_tc[:, -2]) < thr).sum())
    bad = itc[live] != i32[live]
    hist.append(float(margin[bad].max()) if bool(bad.any()) else 0.0)
    sim.step(policy_actions=a32)
print(json.dumps({"set": "formations 1024x20, 8 steps", "agents": tot, "agreement": same / tot, "max_abs_dprob": maxdp,
                  "largest_margin_of_a_disagreement": max(hist),
                  "fraction_with_bf16_margin_below": {str(k): v / tot for k, v in low.items()}}))
sim.close()
