"""Where does a C2 step_host call spend its time?  (diagnostic, GPU box)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gym-exploration-2d_b200")]
import torch
from casim import lib, scenarios as S
from casim.sim import BatchedSim
from casim.config import NUM_STATS
E, A = 4096, 10
cfg = lib.default_config(); cfg.auto_reset = 1
sim = BatchedSim(E, A, cfg=cfg, sensors=(), want_debug=False, want_misc=False, want_stats=True)
init, types, n = S.open_space(E, A, seed=1, sensors=0)
sim.reset(init, types, n)
host = {"rewards": torch.zeros((E, A)).pin_memory(), "done": torch.zeros((E, A), dtype=torch.uint8).pin_memory(),
        "game_over": torch.zeros((E,), dtype=torch.uint8).pin_memory(), "stats": torch.zeros((NUM_STATS,)).pin_memory()}
act_host = torch.zeros((E, A, 2)).pin_memory(); act_dev = torch.zeros((E, A, 2), device="cuda")
def t(fn, n=500):
    for _ in range(20): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
hv = BatchedSim.HostView(host); empty = BatchedSim.HostView({})
print("step async (no sync)        %.1f us" % t(lambda: sim.step()))
print("step + stream sync          %.1f us" % t(lambda: (sim.step(), torch.cuda.current_stream().synchronize())))
print("step_host no io             %.1f us" % t(lambda: sim.step_host(empty)))
print("step_host outputs (dict)    %.1f us" % t(lambda: sim.step_host(host)))
print("step_host outputs (view)    %.1f us" % t(lambda: sim.step_host(hv)))
print("step_host actions only      %.1f us" % t(lambda: sim.step_host(empty, ext_actions_host=act_host, ext_actions_dev=act_dev)))
print("step_host actions + outputs %.1f us" % t(lambda: sim.step_host(hv, ext_actions_host=act_host, ext_actions_dev=act_dev)))
