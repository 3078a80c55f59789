"""TEST / BENCH INPUT -- tests/golden/c2_10_agents_500_cases.npz: the reference's own 10-agent scenario suite
(gym_collision_avoidance/envs/test_cases/10_agents_500_cases.p; rows [start_x, start_y, goal_x, goal_y, pref_speed,
radius], loaded by envs/test_cases.py:get_testcase_fixed_initial_conditions) as a float32 array, so that BASELINE.json's
config C2 can run on the GPU box where /root/reference does not exist (SURVEY.md 8(d) "C2 input").

    python oracle/gen_bench_scenarios.py        (build container)

The pickle is read with a restricted Unpickler (numpy globals only); its SHA-256 is stored next to the data and
checked by tests/test_host_cpu.py when the reference tree is present."""
import hashlib
import os
import pickle

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(_HERE), "tests", "golden")
SRC = "/root/reference/gym_collision_avoidance/envs/test_cases/10_agents_500_cases.p"


class _NumpyOnly(pickle.Unpickler):
    def find_class(self, module, name):
        if module.split(".")[0] == "numpy":
            return super().find_class(module, name)
        raise pickle.UnpicklingError("refusing %s.%s" % (module, name))


def load_cases(path=SRC):
    with open(path, "rb") as f:
        raw = f.read()
    import io
    cases = np.asarray(_NumpyOnly(io.BytesIO(raw), encoding="latin1").load(), dtype=np.float64)
    return cases, hashlib.sha256(raw).hexdigest()


def main():
    cases, sha = load_cases()
    assert cases.shape == (500, 10, 6)
    np.savez_compressed(os.path.join(GOLDEN, "c2_10_agents_500_cases.npz"), cases=cases.astype(np.float32),
                        source_sha256=np.array(sha), source=np.array("envs/test_cases/10_agents_500_cases.p"))
    print("wrote c2_10_agents_500_cases.npz", cases.shape, sha[:16])


if __name__ == "__main__":
    main()
