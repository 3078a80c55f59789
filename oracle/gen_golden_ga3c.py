"""TEST INFRASTRUCTURE -- generates tests/golden/ga3c_state.npz by running the UNMODIFIED reference function
GA3CCADRLPolicy.agents_to_ga3c_cadrl_state (gym_collision_avoidance/envs/policies/GA3CCADRLPolicy.py:45-106)
in the build container.  Run through ``python oracle/gen_golden.py ga3c``.

TensorFlow 1.15 is absent: the policy object is created with object.__new__ (its __init__ would build the TF
graph), only the state-vector builder is executed -- it needs numpy alone.  The network numerics therefore stay
"parity unpinned" (oracle/casim_oracle_ga3c.c header).
"""
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(_HERE), "tests", "golden")


def main(ns, ncases=160, seed=4242):
    from oracle import gen_golden as G
    from casim import layout as L
    from gym_collision_avoidance.envs.policies.GA3CCADRLPolicy import GA3CCADRLPolicy
    rng = np.random.default_rng(seed)
    pol = object.__new__(GA3CCADRLPolicy)
    A = 20
    states = np.zeros((ncases, A, L.STATE_WORDS))
    n_agents = np.zeros((ncases,), np.int32)
    max_obs = np.zeros((ncases,), np.int32)
    nn_input = np.zeros((ncases, A, 75))
    t = L.encode_type(L.POLICY_GA3C_CADRL, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS)
    for k in range(ncases):
        n = int(rng.choice([2, 3, 6, 10, 11, 14, 20]))
        n_agents[k] = n
        # Config.MAX_NUM_OTHER_AGENTS_OBSERVED: 9 by default (config.py:72); 19 when a script raises
        # MAX_NUM_AGENTS_IN_ENVIRONMENT to 20 would overflow the 10 network slots, so stay <= 10
        max_obs[k] = int(rng.choice([9, 9, 10, 3]))
        ns.Config.MAX_NUM_OTHER_AGENTS_OBSERVED = int(max_obs[k])
        init = np.zeros((n, 7))
        init[:, 0:2] = rng.uniform(-6, 6, (n, 2))
        init[:, 2:4] = rng.uniform(-6, 6, (n, 2))
        init[:, 4] = rng.choice([0.2, 0.3, 0.5, 0.5, 0.8], n)
        init[:, 5] = rng.uniform(0.5, 1.5, n)
        init[:, 6] = rng.uniform(-np.pi, np.pi, n)
        if k % 4 == 1:
            # lattice positions and equal radii: many ties of round(dist, 2), decided by p_orthog / list order
            init[:, 0:2] = rng.integers(-4, 5, (n, 2)) * 0.5
            init[:, 4] = 0.25
            for i in range(n):          # avoid coincident agents
                init[i, 0] += 0.0 if i % 2 else 0.25
        if k % 4 == 2:
            # distances engineered onto x.xx5 decimal boundaries (Python round-half-even on the binary value)
            init[0, 0:2] = 0.0
            for i in range(1, n):
                ang = rng.uniform(-np.pi, np.pi)
                d = rng.integers(100, 900) / 100.0 + 0.005 + init[0, 4] + init[i, 4]
                init[i, 0:2] = d * np.cos(ang), d * np.sin(ang)
        init = np.float32(init).astype(np.float64)
        # any policy class will do for the Agent objects: only their kinematic attributes are read (:45-106)
        tb = L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS)
        agents = G.build_agents(ns, init, np.full((n,), tb, np.int32))
        for i, ag in enumerate(agents):
            s = G.extract_state(ag, t)
            s[L.VX], s[L.VY] = np.float32(rng.uniform(-1, 1, 2))
            s[L.DIST_TO_GOAL] = np.float32(np.hypot(s[L.GX] - s[L.PX], s[L.GY] - s[L.PY]))
            s[L.HEADING_EGO] = np.float32(s[L.HEADING_EGO])
            s = np.float32(s).astype(np.float64)
            s[L.TYPE] = t
            G.inject_state(ag, s)
            states[k, i] = s
        for i in range(n):
            obs = pol.agents_to_ga3c_cadrl_state(agents[i], agents[:i] + agents[i + 1:])
            assert obs[0] == agents[i].id
            nn_input[k, i] = obs[1:]
    ns.Config.MAX_NUM_OTHER_AGENTS_OBSERVED = 9
    np.savez_compressed(os.path.join(GOLDEN, "ga3c_state.npz"), state=states, n_agents=n_agents, max_obs=max_obs,
                        nn_input=nn_input)
    print("wrote ga3c_state.npz: %d cases, up to %d agents" % (ncases, A))


def gen_checkpoint_fixture():
    """tests/golden/ga3c_iros18.npz: the 170 764 trained parameters of the reference's default checkpoint
    (GA3C_CADRL/checkpoints/IROS18/network_01900000, GA3CCADRLPolicy.py:24-30) read by casim.ga3c_checkpoint, plus
    per-variable float64 sums as known answers for the reader."""
    from oracle import ref_harness as rh
    from casim.ga3c_checkpoint import GA3C_VARIABLES, ga3c_weights_from_checkpoint, read_checkpoint
    prefix = os.path.join(rh.REFERENCE_ROOT, "gym_collision_avoidance", "envs", "policies", "GA3C_CADRL", "checkpoints",
                          "IROS18", "network_01900000")
    w = ga3c_weights_from_checkpoint(prefix)
    t = read_checkpoint(prefix)
    sums = np.array([t[name].astype(np.float64).sum() for name, _ in GA3C_VARIABLES])
    np.savez_compressed(os.path.join(GOLDEN, "ga3c_iros18.npz"), weights=w, sums=sums, step=t["step:0"])
    print("wrote ga3c_iros18.npz", w.shape, int(t["step:0"]))
