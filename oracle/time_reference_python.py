"""TEST / BENCH INFRASTRUCTURE -- times the UNMODIFIED reference's own Python `CollisionAvoidanceEnv.step`
(/root/reference behind oracle/ref_harness.py) on the bench's synthetic scenarios, ONE core of the
build container, and writes profiles/reference_python_rates.json.  bench.py quotes that file next to the C-port figure
in `cpu_baseline` ("reference_python"): the GPU box has no /root/reference, so the Python step cannot be timed there.

    python oracle/time_reference_python.py [c2 c3 full c4 rvo]

Per workload: E_s environments of the bench's scenario generator (same seeds), each stepped `steps` times after 2
warm-up steps with `env.step({})`; animation off, stdout suppressed; the rate is agent-steps / wall time.  C5 cannot be
run: TensorFlow 1.15 is absent (the GA3C network of the reference cannot be built)."""
import json
import os
import platform
import sys
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
sys.path.insert(0, _ROOT)
sys.path.insert(0, os.path.join(_ROOT, "gym-exploration-2d_b200"))

from oracle import ref_harness as rh  # noqa: E402
from oracle import gen_golden as G  # noqa: E402
from casim import layout as L  # noqa: E402


def run(ns, wl, n_envs, steps):
    import bench
    E, A = n_envs, bench.WORKLOADS[wl]["A"]
    init, types, n, rects = bench.scenario(wl, E, A, seed=1234)
    obstacles = G.CROSSING if rects is not None else []
    if wl == "rvo":
        sys.path.insert(0, os.path.join(_HERE, "_ref"))        # the reference's own rvo2 binding, built by `make ref`
        sys.modules.pop("rvo2", None)
        import rvo2  # noqa: F401
        from gym_collision_avoidance.envs.policies import RVOPolicy as RP
        RP.rvo2 = rvo2
    pol, dyn = G.policy_classes(ns)
    if wl == "rvo":
        from gym_collision_avoidance.envs.policies.RVOPolicy import RVOPolicy
        pol[L.POLICY_RVO] = RVOPolicy
    if wl == "c4":
        pol[L.POLICY_IG_GREEDY] = ns.ig_greedy
    envs = []
    for e in range(E):
        agents = []
        union = 0
        for i in range(int(n[e])):
            p, d, m = L.decode_type(types[e, i])
            union |= m
            h = init[e, i, 6]
            with rh.quiet():
                ag = ns.Agent(float(init[e, i, 0]), float(init[e, i, 1]), float(init[e, i, 2]), float(init[e, i, 3]),
                              float(init[e, i, 4]), float(init[e, i, 5]), None if np.isnan(h) else np.float64(h),
                              pol[p], dyn[d], G.sensor_list(ns, m), i)
            agents.append(ag)
        env = rh.make_env(ns, agents, obstacles, G.states_for(union))
        if wl == "c4":
            for ag in agents:
                if type(ag.policy).__name__ == "ig_greedy":
                    with rh.quiet():                       # experiments/src/dmcts.py:73-78
                        ag.policy.init_maps(ag, env.map, (30, 30), 0.1, 60.0, 5.0, dt=0.1)
                    # ig_greedy.py:39 calls targetMap.update(pose, targets) with a bare pose, which raises as soon as a
                    # target is detected; the list form is what ig_mcts uses (documented deviation)
                    _u = ag.policy.targetMap.update
                    ag.policy.targetMap.update = lambda pose, targets, frame='global', _u=_u: _u([pose], [targets], frame=frame)
        envs.append(env)
    for env in envs:
        for _ in range(2):
            rh.step(env)
    t0 = time.perf_counter()
    agent_steps = 0
    for env in envs:
        for _ in range(steps):
            rh.step(env)
            agent_steps += len(env.agents)
    dt = time.perf_counter() - t0
    return agent_steps / dt, dt / (len(envs) * steps)


def main():
    want = sys.argv[1:] or ["c2", "c3", "full", "c4", "rvo"]
    ns = rh.modules()
    out = {"host": platform.processor() or platform.machine(), "cores_used": 1, "python": platform.python_version(),
           "numpy": np.__version__, "how": "oracle/time_reference_python.py: unmodified reference env.step({}) behind the import "
           "stubs, bench.scenario(workload, seed=1234), 2 warm-up + timed steps per env, one core of the build container",
           "workloads": {}}
    path = os.path.join(_ROOT, "profiles", "reference_python_rates.json")
    if os.path.exists(path):
        out["workloads"] = json.load(open(path)).get("workloads", {})
    plan = {"c2": (8, 20), "c3": (4, 10), "full": (2, 5), "c4": (2, 5), "rvo": (4, 10)}
    for wl in want:
        try:
            rate, sps = run(ns, wl, *plan[wl])
            out["workloads"][wl] = {"agent_steps_per_s_per_core": rate, "s_per_env_step": sps, "envs": plan[wl][0],
                                    "steps": plan[wl][1]}
            print(wl, "%.1f agent-steps/s/core, %.1f ms per env-step" % (rate, sps * 1e3))
        except Exception as ex:          # noqa: BLE001 -- record why a workload cannot run on the reference
            out["workloads"][wl] = {"unavailable": "%s: %s" % (type(ex).__name__, ex)}
            print(wl, "unavailable:", ex)
    out["workloads"]["c5"] = {"unavailable": "TensorFlow 1.15 (GA3C_CADRL/network.py) is not installable in this image"}
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
