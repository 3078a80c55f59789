"""ctypes binding of libcasim.so (include/casim.h).

The library is the product's only compute path: there is no Python/CPU fallback.  Importing this
module on a machine where ``libcasim.so`` has not been built raises immediately.
"""
import ctypes
import os

from .config import SimConfig, NUM_STATS

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcasim.so")

ABI_VERSION = 1
EINVAL = 10001

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_double = ctypes.c_double


class Obs(ctypes.Structure):
    """``struct casim_obs``: device (or host, for the *_host call) pointers; 0 = not wanted."""
    _fields_ = [("obs_misc", c_void_p), ("other_agents_states", c_void_p), ("laserscan", c_void_p),
                ("local_grid", c_void_p), ("rewards", c_void_p), ("done", c_void_p), ("game_over", c_void_p),
                ("collision_flags", c_void_p), ("nearest_dist", c_void_p), ("stats", c_void_p)]


OBS_FIELDS = tuple(n for n, _ in Obs._fields_)


class RvoParams(ctypes.Structure):
    """``struct casim_rvo_params``: what RVOPolicy.__init__ reads from Config (RVOPolicy.py:13-28)."""
    _fields_ = [("neighbor_dist", c_double), ("time_horizon", c_double), ("radius_scale", c_double),
                ("max_delta_heading", c_double), ("max_neighbors", ctypes.c_int32), ("reserved", ctypes.c_int32 * 3)]


class MctsTrees(ctypes.Structure):
    """``struct casim_mcts_trees``: dimensions, constants and DEVICE pointers of T Dec-MCTS trees (include/casim.h)."""
    _fields_ = [("T", ctypes.c_int32), ("M", ctypes.c_int32), ("H", ctypes.c_int32), ("nsims", ctypes.c_int32),
                ("n_comms", ctypes.c_int32), ("comm_n", ctypes.c_int32), ("c_p", c_double), ("gamma", c_double),
                ("slot", c_void_p), ("n_nodes", c_void_p), ("node", c_void_p), ("parent", c_void_p), ("child", c_void_p),
                ("nchild", c_void_p), ("stage", c_void_p), ("has_rollout", c_void_p), ("mu", c_void_p), ("N", c_void_p),
                ("best_reward", c_void_p), ("path", c_void_p), ("best_seq", c_void_p), ("comm_q", c_void_p),
                ("comm_masks", c_void_p), ("comm_valid", c_void_p), ("overflow", c_void_p)]


# every symbol include/casim.h declares: name -> (restype, argtypes)
_PCFG = ctypes.POINTER(SimConfig)
_POBS = ctypes.POINTER(Obs)
SYMBOLS = {
    "casim_last_error": (ctypes.c_char_p, []),
    "casim_abi_version": (c_int, []),
    "casim_default_config": (None, [_PCFG]),
    "casim_create": (c_int, [_PCFG, c_int, c_int, ctypes.POINTER(c_void_p)]),
    "casim_destroy": (c_int, [c_void_p]),
    "casim_set_config": (c_int, [c_void_p, _PCFG]),
    "casim_set_obstacles": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "casim_set_map": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
    "casim_rasterize_noalloc": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "casim_set_stats_event": (c_int, [c_void_p, c_void_p]),
    "casim_set_grid_packed_output": (c_int, [c_void_p, c_void_p]),
    "casim_set_trajectories": (c_int, [c_void_p, c_void_p, c_int]),
    "casim_reset": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, _POBS, c_void_p]),
    "casim_reset_envs": (c_int, [c_void_p, c_void_p, c_void_p, _POBS, c_void_p]),
    "casim_observe": (c_int, [c_void_p, c_void_p, _POBS, c_void_p]),
    "casim_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_double, _POBS, c_void_p]),
    "casim_step_n": (c_int, [c_void_p, c_void_p, c_void_p, c_double, c_int, _POBS, c_void_p]),
    "casim_run_sensor": (c_int, [c_void_p, c_void_p, _POBS, c_int, c_void_p]),
    "casim_step_host": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_double, _POBS, _POBS,
                                c_void_p]),
    "casim_step_host_pipelined": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_double, _POBS, _POBS,
                                          c_void_p, c_void_p, c_void_p]),
    "casim_host_flush": (c_int, [c_void_p]),
    "casim_explore_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_void_p]),
    "casim_generate_scenario": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p]),
    "casim_explore_team_update": (c_int, [c_void_p] * 11),
    "casim_explore_rollouts": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int,
                                       c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "casim_explore_mi_table": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "casim_explore_set_mi_table": (c_int, [c_void_p, c_void_p]),
    "casim_explore_set_rollout_slots": (c_int, [c_void_p, c_void_p, c_int]),
    "casim_mcts_select": (c_int, [c_void_p, ctypes.POINTER(MctsTrees), c_void_p, ctypes.c_uint64, c_int, c_void_p, c_void_p,
                          c_void_p, c_int, c_void_p]),
    "casim_mcts_update": (c_int, [c_void_p, ctypes.POINTER(MctsTrees), c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "casim_ga3c_set_weights": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
    "casim_ga3c_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "casim_rvo_default_params": (None, [ctypes.POINTER(RvoParams)]),
    "casim_rvo_set_obstacles": (c_int, [c_void_p, ctypes.POINTER(c_double), ctypes.POINTER(ctypes.c_int32), c_int]),
    "casim_rvo_step": (c_int, [c_void_p, c_void_p, c_void_p, ctypes.POINTER(RvoParams), c_void_p, c_void_p]),
    "casim_flat_row_size": (c_int, [c_void_p, ctypes.POINTER(ctypes.c_int32), c_int]),
    "casim_flatten_obs": (c_int, [c_void_p, c_void_p, _POBS, ctypes.POINTER(ctypes.c_int32), c_int, c_void_p, c_void_p]),
    "casim_launch_count": (ctypes.c_ulonglong, [c_void_p]),
}

_lib = None


class CasimError(RuntimeError):
    pass


def load():
    """Load libcasim.so and bind every declared symbol.  Raises if the library is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CasimError("libcasim.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` "
                         "or `make -C gym-exploration-2d_b200/casim/csrc`. There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export a declared symbol
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.casim_abi_version() != ABI_VERSION:
        raise CasimError("libcasim.so ABI version %d, binding expects %d" % (lib.casim_abi_version(), ABI_VERSION))
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().casim_last_error()
        raise CasimError("%s failed (code %d): %s" % (what, rc, msg.decode() if msg else ""))


def default_config():
    c = SimConfig()
    load().casim_default_config(ctypes.byref(c))
    return c


def rvo_default_params():
    p = RvoParams()
    load().casim_rvo_default_params(ctypes.byref(p))
    return p


def ptr(t):
    """Device (or pinned host) pointer of a torch tensor / numpy array; None -> NULL."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        if not t.is_contiguous():
            raise CasimError("tensor passed to libcasim must be contiguous")
        return c_void_p(t.data_ptr())
    return c_void_p(t.ctypes.data)
