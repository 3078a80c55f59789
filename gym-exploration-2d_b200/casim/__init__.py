"""casim -- B200-native batched simulator for the CollisionAvoidanceEnv stepping path of
mlodel/gym-exploration-2d.  Host code in Python, compute in libcasim.so (hand-written sm_100a CUDA behind the
C ABI of include/casim.h).  No CPU fallback: modules that compute import ``casim.lib`` which fails loudly
when the library is missing."""
from . import layout  # noqa: F401
from .config import Config, SimConfig, make_sim_config  # noqa: F401
from . import plugins  # noqa: F401
from .plugins import Agent  # noqa: F401

__all__ = ["layout", "plugins", "Agent", "Config", "SimConfig", "make_sim_config"]
