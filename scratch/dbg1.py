import sys; sys.path[:0]=['/root/repo','/root/repo/gym-exploration-2d_b200','/root/repo/tests']
import numpy as np, torch, helpers as H
from casim import layout as L, lib
from casim.sim import BatchedSim
from oracle import oracle
rng = np.random.default_rng(21)
E, A = 64, 6
b = H.random_batch(rng, E, A, sensor_choices=(1,), policies=[L.POLICY_NONCOOPERATIVE, L.POLICY_STATIC, L.POLICY_NONCOOPERATIVE], dynamics=[L.DYN_UNICYCLE], spread=3.0)
cfg = lib.default_config(); cfg.auto_reset=1
sim = BatchedSim(E, A, cfg=cfg, sensors=("other_agents_states",))
sim.reset(b["init"], b["types"], b["n_agents"])
prev_go = np.zeros((E,), np.uint8)
for t in range(30):
    ostate = sim.state_numpy()
    pre = ostate.copy()
    oobs = oracle.step(cfg, ostate, n_agents=b["n_agents"], init=b["init"], types0=b["types"], prev_game_over=prev_go)
    sim.step(); torch.cuda.synchronize()
    g = sim.buf["other_agents_states"].cpu().numpy()
    d = np.abs(g - oobs.other_agents_states)
    bad = d > 1e-5 + 1e-5*np.abs(oobs.other_agents_states)
    if bad.any():
        idx = np.argwhere(bad)
        print("t", t, idx[:5])
        e,a,r,c = idx[0]
        print("gpu row", g[e,a,r]); print("ora row", oobs.other_agents_states[e,a,r])
        print("host pre", pre[e,a]); print("host post", ostate[e,a]); print("gpu post", sim.state_numpy()[e,a])
        print("prev_go", prev_go[e])
        break
    prev_go = oobs.game_over.copy()
