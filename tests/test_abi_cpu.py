"""CPU: the C-ABI library builds, loads without a GPU, and exports every symbol include/casim.h declares."""
import ctypes
import os
import re

from casim import lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "casim.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(casim_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 15
    l = lib.load()
    for n in names:
        assert hasattr(l, n), n
    assert set(names) == set(lib.SYMBOLS), set(names) ^ set(lib.SYMBOLS)
    assert l.casim_abi_version() == 1


def test_default_config_matches_oracle_and_python_snapshot(oracle_lib):
    from casim.config import Config, make_sim_config
    a, b, c = lib.default_config(), oracle_lib.default_config(), make_sim_config(Config)
    assert bytes(a) == bytes(b)
    for name, _ in a._fields_:
        if name in ("game_over_mode", "reserved", "laser_angles", "laser_ranges"):
            continue
        assert getattr(a, name) == getattr(c, name), name
    assert list(a.laser_angles) == list(c.laser_angles) and list(a.laser_ranges) == list(c.laser_ranges)


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        return
    h = ctypes.c_void_p()
    cfg = lib.default_config()
    rc = lib.load().casim_create(ctypes.byref(cfg), 4, 2, ctypes.byref(h))
    assert rc != 0 and lib.load().casim_last_error()
    from casim.sim import BatchedSim
    try:
        BatchedSim(4, 2)
    except lib.CasimError:
        pass
    else:
        raise AssertionError("BatchedSim must refuse to run without a CUDA device")


def test_invalid_arguments_are_rejected():
    h = ctypes.c_void_p()
    cfg = lib.default_config()
    l = lib.load()
    assert l.casim_create(ctypes.byref(cfg), 4, 33, ctypes.byref(h)) == lib.EINVAL
    assert l.casim_create(ctypes.byref(cfg), 0, 2, ctypes.byref(h)) == lib.EINVAL
    cfg.map_cell = 0.2
    assert l.casim_create(ctypes.byref(cfg), 4, 2, ctypes.byref(h)) == lib.EINVAL
    assert b"300x300" in l.casim_last_error()
