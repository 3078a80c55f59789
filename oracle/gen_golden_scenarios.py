"""TEST INFRASTRUCTURE -- generates tests/golden/scenarios.npz by running the UNMODIFIED reference generators
train_agents_swap_circle / train_agents_pairwise_swap / train_agents_random_positions
(gym_collision_avoidance/envs/test_cases.py:1192-1463) with
explicit seeds in the build container.  Run through ``python oracle/gen_golden.py scenarios``."""
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(_HERE), "tests", "golden")

POLICY_ID = {"RVOPolicy": 8, "NonCooperativePolicy": 1, "LearningPolicy": 3, "ig_greedy": 5}
DYN_ID = {"UnicycleDynamics": 0, "FirstOrderDynamics": 1, "UnicycleDynamicsMaxAcc": 4}
SENSOR_BIT = {"OtherAgentsStatesSensor": 1, "LaserScanSensor": 2, "OccupancyGridSensor": 4}


def main(ns, A=12):
    from oracle import ref_harness
    with ref_harness.quiet():
        from gym_collision_avoidance.envs import test_cases as tc
    funcs = [tc.train_agents_swap_circle, tc.train_agents_random_positions, tc.train_agents_pairwise_swap]
    rec = {k: [] for k in ("kind", "seed", "number_of_agents", "init", "types", "n_agents", "coop")}
    for kind, f in enumerate(funcs):
        for n in (2, 4, 5, 6, 10, 12):
            for seed in (1, 2, 3, 7, 1234, 99991, 2 ** 31 + 5):
                with ref_harness.quiet():
                    agents, obstacles = f(number_of_agents=n, ego_agent_policy=ns.LearningPolicy,
                                          ego_agent_dynamics=ns.FirstOrderDynamics, seed=seed)
                assert len(obstacles) == 0 and len(agents) <= A
                init = np.zeros((A, 7))
                types = np.zeros(A, np.int32)
                coop = np.zeros(A)
                for k, a in enumerate(agents):
                    assert a.id == k
                    mask = 0
                    for s in a.sensors:
                        mask |= SENSOR_BIT[type(s).__name__]
                    init[k] = [a.pos_global_frame[0], a.pos_global_frame[1], a.goal_global_frame[0], a.goal_global_frame[1],
                               a.radius, a.pref_speed, a.heading_global_frame]       # heading: towards the goal (agent.py)
                    types[k] = POLICY_ID[type(a.policy).__name__] | (DYN_ID[type(a.dynamics_model).__name__] << 8) | (mask << 16)
                    coop[k] = a.cooperation_coef
                rec["kind"].append(kind); rec["seed"].append(seed); rec["number_of_agents"].append(n)
                rec["init"].append(init); rec["types"].append(types); rec["n_agents"].append(len(agents)); rec["coop"].append(coop)
    out = {k: np.array(v) for k, v in rec.items()}
    out["seed"] = out["seed"].astype(np.uint32)
    np.savez_compressed(os.path.join(GOLDEN, "scenarios.npz"), **out)
    nc = sum(int(((t & 0xff) == 1).sum()) for t in rec["types"])
    print("wrote scenarios.npz: %d scenarios (3 generators x 6 sizes x 7 seeds), %d NonCooperative agents" % (len(rec["kind"]), nc))
