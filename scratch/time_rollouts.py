import sys, time
sys.path.insert(0, "gym-exploration-2d_b200"); sys.path.insert(0, "tests")
import numpy as np, torch
from casim import lib, scenarios as S
from casim.sim import BatchedSim
E, K, Hn = 1024, 10, 10
init, types, n, rects = S.exploration(E, seed=99)
cfg = lib.default_config(); cfg.game_over_mode = 0
sim = BatchedSim(E, 4, cfg=cfg, sensors=("other_agents_states",)); sim.enable_exploration()
sim.set_obstacles(rects); sim.reset(init, types, n)
for _ in range(3):
    sim.explore_step(); sim.step(policy_actions=sim.buf["policy_actions"])
rng = np.random.default_rng(3)
actions = torch.from_numpy(rng.integers(0, 9, (E, 4, K, Hn)).astype(np.uint8)).cuda()
for want in (True, False):
    for _ in range(5): sim.explore_rollouts(actions, want_end_pose=want, want_obsvd=want)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): sim.explore_rollouts(actions, want_end_pose=want, want_obsvd=want)
    b.record(); torch.cuda.synchronize()
    us = a.elapsed_time(b) / 20 * 1e3
    print("rollouts E=%d x 2 explorers x K=%d x H=%d outputs=%s: %.1f us/launch, %.2f M rollout-steps/s" % (E, K, Hn, want, us, E * 2 * K * Hn / us))
