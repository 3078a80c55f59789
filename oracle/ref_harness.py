"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Harness that imports the UNMODIFIED reference (mlodel/gym-exploration-2d) from
/root/reference inside the build container, so that golden input/output vectors
can be generated from the reference itself (``oracle/gen_golden.py``).

/root/reference does not exist on the GPU box: nothing in ``tests -m gpu``,
``smoke()`` or ``bench.py`` imports this module.  It is used only by
``oracle/gen_golden.py`` and by the (auto-skipped when the reference is absent)
cross-checks in ``tests/test_oracle_vs_reference.py``.

The reference cannot be imported as-is here because several third-party
modules are missing (gym, matplotlib, tensorflow, rvo2, ...).  Three shims,
no source edits:
  1. a meta-path finder that serves MagicMock packages for the missing modules,
     with real minimal classes where the reference subclasses/instantiates them
     (gym.Env, gym.spaces.Box/Dict/Discrete, DummyVecEnv ...);
  2. ``env._take_action_dmcts`` guarded for an empty agent list
     (reference bug at collision_avoidance_env.py:329,346);
  3. ``cv2.getRotationMatrix2D`` wrapped to cast ``center`` to Python floats
     (cv2 4.13 rejects numpy ints; call site OccupancyGridSensor.py:140).
"""
import contextlib
import importlib.abc
import importlib.machinery
import io
import os
import sys
import types
from unittest.mock import MagicMock

REFERENCE_ROOT = os.environ.get("CASIM_REFERENCE_ROOT", "/root/reference")

_MISSING = ("gym", "matplotlib", "imageio", "moviepy", "pypoman", "pylab", "rvo2",
            "tensorflow", "stable_baselines", "mpl_toolkits", "pykalman", "baselines",
            "shapely", "PIL", "networkx_stub")

_installed = False


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "gym_collision_avoidance"))


class _Loader(importlib.abc.Loader):
    def create_module(self, spec):
        m = MagicMock(name=spec.name)
        m.__path__ = []
        m.__name__ = spec.name
        m.__spec__ = spec
        m.__loader__ = self
        return m

    def exec_module(self, module):
        pass


class _Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, name, path, target=None):
        root = name.split(".")[0]
        if root in _MISSING:
            try:  # serve the real module if it happens to exist
                for f in sys.meta_path:
                    if f is self:
                        continue
                    s = f.find_spec(name, path, target) if hasattr(f, "find_spec") else None
                    if s is not None:
                        return s
            except Exception:
                pass
            return importlib.machinery.ModuleSpec(name, _Loader(), is_package=True)
        return None


def install():
    """Install the shims and put the reference on sys.path (idempotent)."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError("reference tree not found at %s" % REFERENCE_ROOT)
    sys.meta_path.insert(0, _Finder())

    import gym  # noqa: the mock

    class Env(object):
        pass

    class ObservationWrapper(object):
        def __init__(self, env=None):
            self.env = env

    class Box(object):
        def __init__(self, low=None, high=None, shape=None, dtype=None):
            self.low, self.high, self.shape, self.dtype = low, high, shape, dtype

    class Dict(object):
        def __init__(self, spaces=None):
            self.spaces = dict(spaces or {})

    class Discrete(object):
        def __init__(self, n=0, dtype=None):
            self.n = n

    gym.Env = Env
    gym.ObservationWrapper = ObservationWrapper
    import gym.spaces
    gym.spaces.Box = Box
    gym.spaces.Dict = Dict
    gym.spaces.Discrete = Discrete
    import gym.envs.registration
    gym.envs.registration.register = lambda **kw: None

    import stable_baselines.common.vec_env.dummy_vec_env as dve

    class DummyVecEnv(object):
        def __init__(self, env_fns=None):
            pass

    dve.DummyVecEnv = DummyVecEnv

    import cv2
    _g = cv2.getRotationMatrix2D

    def _rot(center, angle, scale):
        return _g((float(center[0]), float(center[1])), float(angle), float(scale))

    cv2.getRotationMatrix2D = _rot

    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    _installed = True


@contextlib.contextmanager
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield


def modules():
    """Import and return the reference modules the oracle cross-checks use."""
    install()
    ns = types.SimpleNamespace()
    with quiet():
        from gym_collision_avoidance.envs.config import Config
        from gym_collision_avoidance.envs.collision_avoidance_env import CollisionAvoidanceEnv
        from gym_collision_avoidance.envs.agent import Agent
        from gym_collision_avoidance.envs.Map import Map
        from gym_collision_avoidance.envs.policies.NonCooperativePolicy import NonCooperativePolicy
        from gym_collision_avoidance.envs.policies.StaticPolicy import StaticPolicy
        from gym_collision_avoidance.envs.policies.ExternalPolicy import ExternalPolicy
        from gym_collision_avoidance.envs.policies.LearningPolicy import LearningPolicy
        from gym_collision_avoidance.envs.policies.CARRLPolicy import CARRLPolicy
        from gym_collision_avoidance.envs.policies.ig_greedy import ig_greedy
        from gym_collision_avoidance.envs.dynamics.UnicycleDynamics import UnicycleDynamics
        from gym_collision_avoidance.envs.dynamics.FirstOrderDynamics import FirstOrderDynamics
        from gym_collision_avoidance.envs.dynamics.ExternalDynamics import ExternalDynamics
        from gym_collision_avoidance.envs.sensors.OtherAgentsStatesSensor import OtherAgentsStatesSensor
        from gym_collision_avoidance.envs.sensors.LaserScanSensor import LaserScanSensor
        from gym_collision_avoidance.envs.sensors.OccupancyGridSensor import OccupancyGridSensor
        from gym_collision_avoidance.envs.information_models.edfMap import edfMap
        from gym_collision_avoidance.envs.information_models.targetMap import targetMap
    for k, v in list(locals().items()):
        if k not in ("ns",):
            setattr(ns, k, v)
    Config.ANIMATE_EPISODES = False
    Config.SAVE_EPISODE_PLOTS = False
    Config.SHOW_EPISODE_PLOTS = False
    Config.TRAIN_SINGLE_AGENT = False
    return ns


def make_env(ns, agents, obstacles, states_in_obs):
    """Build a reference env around hand-made agents, bypassing reset()/_init_agents
    (they ignore set_agents())."""
    ns.Config.STATES_IN_OBS = list(states_in_obs)
    with quiet():
        env = ns.CollisionAvoidanceEnv()
        env.agents = agents
        env.obstacles = obstacles
        env._init_static_map()
        env.episode_number = 2
        _o = env._take_action_dmcts
        env._take_action_dmcts = lambda a: _o(a) if a else {}
        env._get_obs()
    return env


def step(env, actions=None):
    with quiet():
        return env.step(actions if actions is not None else {})
