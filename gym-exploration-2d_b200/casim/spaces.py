"""Observation / action spaces of the gym interface (collision_avoidance_env.py:103-134) without a hard dependency on
``gym``: when gym is importable its own ``gym.spaces`` classes are used (trainers do isinstance checks on them), else the
minimal stand-ins below with the attributes RL code reads (``low``, ``high``, ``shape``, ``dtype``, ``spaces``, ``n``,
``sample()``, ``contains()``)."""
import numpy as np

try:                                            # pragma: no cover - gym is absent from the build image
    from gym.spaces import Box, Dict, Discrete  # noqa: F401
    HAVE_GYM = True
except Exception:
    HAVE_GYM = False

    class Box:
        def __init__(self, low, high, shape=None, dtype=np.float32):
            if shape is None:
                low, high = np.asarray(low), np.asarray(high)
                shape = low.shape
            self.shape = tuple(int(x) for x in np.atleast_1d(shape)) if not isinstance(shape, tuple) else shape
            self.dtype = np.dtype(dtype)
            self.low = np.broadcast_to(np.asarray(low, dtype=self.dtype), self.shape).copy()
            self.high = np.broadcast_to(np.asarray(high, dtype=self.dtype), self.shape).copy()

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

        def sample(self):
            lo = np.where(np.isfinite(self.low), self.low, -1.0)
            hi = np.where(np.isfinite(self.high), self.high, 1.0)
            return np.random.uniform(lo, hi).astype(self.dtype)

        def __repr__(self):
            return "Box%s" % (self.shape,)

    class Discrete:
        def __init__(self, n, dtype=np.int64):
            self.n = int(n)
            self.shape = ()
            self.dtype = np.dtype(dtype)

        def contains(self, x):
            return 0 <= int(x) < self.n

        def sample(self):
            return int(np.random.randint(self.n))

    class Dict:
        def __init__(self, spaces=None):
            self.spaces = dict(spaces or {})

        def __getitem__(self, k):
            return self.spaces[k]

        def keys(self):
            return self.spaces.keys()

        def contains(self, x):
            return all(k in x and s.contains(x[k]) for k, s in self.spaces.items())

        def sample(self):
            return {k: s.sample() for k, s in self.spaces.items()}


def observation_space_from_config(config):
    """collision_avoidance_env.py:128-132: one Box per entry of Config.STATES_IN_OBS with the bounds / size / dtype of
    Config.STATE_INFO_DICT (envs/config.py:104-215)."""
    d = Dict({})
    info = config.STATE_INFO_DICT
    for state in config.STATES_IN_OBS:
        size = info[state]['size']
        lo, hi = info[state]['bounds']
        d.spaces[state] = Box(lo * np.ones(size), hi * np.ones(size), dtype=info[state]['dtype'])
    return d


def action_space_from_env(env):
    """collision_avoidance_env.py:104-112 (continuous action space: [speed, delta heading] within +-4)."""
    low = np.array([env.min_speed, env.min_heading_change])
    high = np.array([env.max_speed, env.max_heading_change])
    return Box(low, high, dtype=np.float32)
