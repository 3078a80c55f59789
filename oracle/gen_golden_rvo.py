"""Pins oracle/casim_oracle_rvo.c against the reference's own RVO2 library and writes tests/golden/rvo.npz.

Runs only where /root/reference exists (the build container): `make -C oracle ref` compiles the reference's C++
sources into oracle/_ref/librvo2_ref.so; this script drives it through oracle/rvo2_ref_shim.cpp (the calls
RVOPolicy.find_next_action makes) on seeded random crowds and stores inputs + the simulator's new ego position.
"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for _p in (ROOT, os.path.join(ROOT, "gym-exploration-2d_b200"), os.path.join(HERE, "_ref")):
    if _p not in sys.path:
        sys.path.insert(0, _p)
from oracle import ref_harness as rh  # noqa: E402


def ref_lib():
    lib = ctypes.CDLL(os.path.join(HERE, "_ref", "librvo2_ref.so"))
    return lib.rvo2_ref_step


def oracle_fn():
    lib = ctypes.CDLL(os.path.join(HERE, "libcasim_oracle.so"))
    return lib.casim_oracle_rvo_ref_step


def call(fn, pos, vel, pref, radius, ms, ego, collab, dt, nd, mn, th):
    P = ctypes.POINTER(ctypes.c_double)
    out = np.zeros(2)
    arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (pos, vel, pref, radius, ms)]
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_int, P, P, P, P, P, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_double,
                   ctypes.c_int, ctypes.c_double, P]
    rc = fn(len(radius), *[a.ctypes.data_as(P) for a in arrs], int(ego), float(collab), float(dt), float(nd), int(mn),
            float(th), out.ctypes.data_as(P))
    assert rc == 0
    return out


def cases(seed=0, count=600):
    rng = np.random.default_rng(seed)
    for c in range(count):
        n = int(rng.integers(2, 21))
        spread = float(rng.choice([1.5, 3.0, 6.0, 12.0]))          # dense crowds exercise linearProgram3 / collisions
        pos = rng.uniform(-spread, spread, (n, 2))
        goal = rng.uniform(-spread, spread, (n, 2))
        speed = rng.uniform(0.5, 1.5, n)
        head = rng.uniform(-np.pi, np.pi, n)
        v = rng.uniform(0, 1, n) * speed
        vel = np.stack([v * np.cos(head), v * np.sin(head)], 1)
        if c % 7 == 0:
            vel[:] = 0
        d = goal - pos
        pref = speed[:, None] / np.linalg.norm(d, axis=1, keepdims=True) * d
        radius = 1.15 * rng.uniform(0.2, 0.8, n)
        collab = float(rng.choice([1.0, 0.5, 0.0, -0.5, -2.0], p=[0.6, 0.2, 0.05, 0.1, 0.05]))
        mn = int(rng.choice([10, 10, 3, 32]))
        nd = float(rng.choice([np.inf, np.inf, 4.0]))
        ego = int(rng.integers(0, n))
        # the state the GPU path holds is fp32: use fp32-representable inputs
        f = lambda a: np.asarray(a, np.float32).astype(np.float64)
        yield dict(pos=f(pos), vel=f(vel), pref=f(pref), radius=f(radius), ms=f(speed), ego=ego, collab=collab, dt=0.1,
                   nd=nd, mn=mn, th=5.0)


def main():
    ref, orc = ref_lib(), oracle_fn()
    rows, bad = [], 0
    for k in cases():
        a = call(ref, **k)
        b = call(orc, **k)
        if not np.array_equal(a, b):
            bad += 1
            print("MISMATCH", k["ego"], k["collab"], k["mn"], k["nd"], a, b)
        rows.append((k, a))
    print("cases %d, oracle != reference: %d" % (len(rows), bad))
    A = 20
    out = dict(n=np.array([len(k["radius"]) for k, _ in rows], np.int32), ego=np.array([k["ego"] for k, _ in rows], np.int32),
               collab=np.array([k["collab"] for k, _ in rows]), mn=np.array([k["mn"] for k, _ in rows], np.int32),
               nd=np.array([k["nd"] for k, _ in rows]), new_pos=np.stack([a for _, a in rows]))
    for name in ("pos", "vel", "pref"):
        arr = np.zeros((len(rows), A, 2))
        for i, (k, _) in enumerate(rows):
            arr[i, :len(k["radius"])] = k[name]
        out[name] = arr
    for name in ("radius", "ms"):
        arr = np.zeros((len(rows), A))
        for i, (k, _) in enumerate(rows):
            arr[i, :len(k["radius"])] = k[name]
        out[name] = arr
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "rvo.npz"), **out)
    return bad


CROSSING = [[(10, 10), (2, 10), (2, 2), (10, 2)], [(-2, 10), (-10, 10), (-10, 2), (-2, 2)],
            [(10, -2), (2, -2), (2, -10), (10, -10)], [(-2, -2), (-10, -2), (-10, -10), (-2, -10)]]


def main_policy_obstacles(ns, ncases=120, seed=7):
    """As main_policy, with the crossing-map obstacles (envs/test_cases.py:3081-3085) handed to find_next_action: the
    policy adds them to its simulator before the first processObstacles (RVOPolicy.py:56-63, :44)."""
    from oracle import gen_golden as G
    from casim import layout as L
    import rvo2
    assert hasattr(rvo2, "PyRVOSimulator") and "Mock" not in type(rvo2).__name__
    from gym_collision_avoidance.envs.policies.RVOPolicy import RVOPolicy
    rng = np.random.default_rng(seed)
    A = 10
    states = np.zeros((ncases, A, L.STATE_WORDS))
    coop = np.ones((ncases, A))
    n_agents = np.zeros((ncases,), np.int32)
    actions = np.zeros((ncases, A, 2))
    t = L.encode_type(L.POLICY_RVO, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS)
    tb = L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS)
    for k in range(ncases):
        n = int(rng.integers(1, 7))
        n_agents[k] = n
        init = np.zeros((n, 7))
        arm = rng.integers(0, 4, n)
        along, across = rng.uniform(0.5, 11.0, n), rng.uniform(-1.6, 1.6, n)      # inside the free cross, near the walls
        init[:, 0] = np.where(arm == 0, along, np.where(arm == 1, -along, across))
        init[:, 1] = np.where(arm == 2, along, np.where(arm == 3, -along, across))
        init[:, 2:4] = rng.uniform(-11, 11, (n, 2))                                # goals anywhere: through the blocks
        init[:, 4] = rng.uniform(0.2, 0.6, n)
        init[:, 5] = rng.uniform(0.5, 1.5, n)
        init[:, 6] = rng.uniform(-np.pi, np.pi, n)
        init = np.float32(init).astype(np.float64)
        agents = G.build_agents(ns, init, np.full((n,), tb, np.int32))
        for i, ag in enumerate(agents):
            s = G.extract_state(ag, t)
            sp = rng.uniform(0, 1) * s[L.PREF_SPEED]
            s[L.VX], s[L.VY] = sp * np.cos(s[L.HEADING]), sp * np.sin(s[L.HEADING])
            s = np.float32(s).astype(np.float64)
            s[L.TYPE] = t
            G.inject_state(ag, s)
            states[k, i] = s
            coop[k, i] = float(rng.choice([1.0, 1.0, 0.5, 0.0]))
            ag.cooperation_coef = coop[k, i]
        for i in range(n):
            with rh.quiet():
                pol = RVOPolicy()
                actions[k, i] = pol.find_next_action({}, agents, i, [list(p) for p in CROSSING])
    path = os.path.join(ROOT, "tests", "golden", "rvo.npz")
    old = dict(np.load(path)) if os.path.exists(path) else {}
    old.update(obst_state=states, obst_coop=coop, obst_n_agents=n_agents, obst_actions=actions,
               obst_polygons=np.array(CROSSING, np.float64))
    np.savez_compressed(path, **old)
    print("wrote rvo.npz policy cases with obstacles: %d" % ncases)


def main_policy(ns, ncases=200, seed=99):
    """Policy level: the UNMODIFIED RVOPolicy.find_next_action (RVOPolicy.py:53-117) over the reference's own Cython
    binding (oracle/_ref/rvo2*.so, built by `make -C oracle ref`) -> tests/golden/rvo.npz keys state / coop /
    n_agents / actions."""
    from oracle import gen_golden as G
    from casim import layout as L
    import rvo2
    assert "MagicMock" not in type(rvo2).__name__ and hasattr(rvo2, "PyRVOSimulator"), "real rvo2 binding not built"
    from gym_collision_avoidance.envs.policies.RVOPolicy import RVOPolicy
    rng = np.random.default_rng(seed)
    A = 10          # Config.MAX_NUM_AGENTS_IN_ENVIRONMENT = maxNeighbors of the policy's simulator
    states = np.zeros((ncases, A, L.STATE_WORDS))
    coop = np.ones((ncases, A))
    n_agents = np.zeros((ncases,), np.int32)
    actions = np.zeros((ncases, A, 2))
    t = L.encode_type(L.POLICY_RVO, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS)
    tb = L.encode_type(L.POLICY_NONCOOPERATIVE, L.DYN_UNICYCLE, L.SENSOR_OTHER_AGENTS)
    for k in range(ncases):
        n = int(rng.integers(2, A + 1))
        n_agents[k] = n
        spread = float(rng.choice([2.0, 4.0, 8.0]))
        init = np.zeros((n, 7))
        init[:, 0:2] = rng.uniform(-spread, spread, (n, 2))
        init[:, 2:4] = rng.uniform(-spread, spread, (n, 2))
        init[:, 4] = rng.uniform(0.2, 0.8, n)
        init[:, 5] = rng.uniform(0.5, 1.5, n)
        init[:, 6] = rng.uniform(-np.pi, np.pi, n)
        init = np.float32(init).astype(np.float64)
        agents = G.build_agents(ns, init, np.full((n,), tb, np.int32))
        for i, ag in enumerate(agents):
            s = G.extract_state(ag, t)
            sp = rng.uniform(0, 1) * s[L.PREF_SPEED] * (k % 5 != 0)
            s[L.VX], s[L.VY] = sp * np.cos(s[L.HEADING]), sp * np.sin(s[L.HEADING])
            s = np.float32(s).astype(np.float64)
            s[L.TYPE] = t
            G.inject_state(ag, s)
            states[k, i] = s
            coop[k, i] = float(rng.choice([1.0, 1.0, 0.5, 0.0, -0.5]))
            ag.cooperation_coef = coop[k, i]
        for i in range(n):
            with rh.quiet():
                pol = RVOPolicy()                       # every agent owns a policy object with its own simulator
                actions[k, i] = pol.find_next_action({}, agents, i, [])
    path = os.path.join(ROOT, "tests", "golden", "rvo.npz")
    old = dict(np.load(path)) if os.path.exists(path) else {}
    old.update(state=states, coop=coop, n_agents=n_agents, actions=actions)
    np.savez_compressed(path, **old)
    print("wrote rvo.npz policy cases: %d" % ncases)


if __name__ == "__main__":
    sys.exit(1 if main() else 0)
