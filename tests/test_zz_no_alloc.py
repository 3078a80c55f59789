"""casim_rasterize_noalloc (include/casim.h) -- the allocation-free variant of the map build: its intermediates live in a
file-scope `static __device__` array of libcasim.so instead of cudaMalloc'ed buffers.  No counterpart in the reference.
Checked through the C-ABI against the existing entry point
(casim_set_obstacles' map_out, bit for bit) and against the reference-generated distance fields of
tests/golden/explore.npz (information_models/edfMap.py:11-12), on two calls in a row that share the scratch array.

Named test_zz_* so that pytest collects it last."""
import ctypes

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

MAX_MAPS = 4          # CASIM_NOALLOC_MAX_MAPS
MAX_RECTS = 256       # CASIM_NOALLOC_MAX_RECTS
EDF_RTOL = 1e-14      # fp64 sqrt * cell on both sides


def _rect_batch(z, ids):
    """-> rects [len(ids), R, 4] float32 (zero padded), n_rects [len(ids)] int32"""
    R = max(len(z["rects_%d" % k]) for k in ids)
    rects = np.zeros((len(ids), R, 4), np.float32)
    n = np.zeros((len(ids),), np.int32)
    for j, k in enumerate(ids):
        r = z["rects_%d" % k]
        rects[j, :len(r)] = r
        n[j] = len(r)
    return rects, n


def _existing(ids, rects, n):
    """casim_set_obstacles on a handle of its own (n_maps == E) -> map_out"""
    import torch
    from casim.sim import BatchedSim
    sim = BatchedSim(len(ids), 2)
    out = sim.set_obstacles(rects, n, want_map=True)
    torch.cuda.synchronize()
    out = out.cpu().numpy()
    sim.close()
    return out


def _noalloc(sim, rects, n, want_edf=True):
    import torch
    from casim import lib as _lib
    dev = torch.device("cuda:0")
    r = torch.from_numpy(rects).to(dev)
    nr = torch.from_numpy(n).to(dev)
    m = r.shape[0]
    map_out = torch.full((m, 300, 300), 7, dtype=torch.uint8, device=dev)
    edf_out = torch.full((m, 300, 300), -1.0, dtype=torch.float64, device=dev) if want_edf else None
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = sim.lib.casim_rasterize_noalloc(sim.handle, _lib.ptr(r), _lib.ptr(nr), m, r.shape[1], _lib.ptr(map_out),
                                         _lib.ptr(edf_out), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, "casim_rasterize_noalloc")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return map_out.cpu().numpy(), (None if edf_out is None else edf_out.cpu().numpy()), free0 - free1


def test_noalloc_rasterize_matches_set_obstacles_two_calls_in_a_row():
    from casim import lib as _lib
    from casim.sim import BatchedSim
    z = H.load("explore.npz")
    maps, d2 = H.explore_maps(z)
    sim = BatchedSim(4, 2)
    sim.set_obstacles(_rect_batch(z, [1])[0][0])        # the handle's own map: must survive the calls below
    cell = float(sim.cfg.map_cell)
    # call 1: three maps; call 2 RIGHT AFTER it: two maps in another order (call 1's contents still in the scratch);
    # call 3: one map, no distance field.  Nothing else touches the GPU in between.
    plan = (([0, 1, 2], True), ([2, 0], True), ([1], False))
    calls = [(_rect_batch(z, ids), want_edf) for ids, want_edf in plan]
    results = [_noalloc(sim, rects, n, want_edf) for (rects, n), want_edf in calls]
    used = [r[2] for r in results]
    for (ids, want_edf), ((rects, n), _), (got_map, got_edf, _g) in zip(plan, calls, results):
        ref_map = _existing(ids, rects, n)
        assert np.array_equal(got_map, ref_map), ids                              # == the existing entry point, bit for bit
        for j, k in enumerate(ids):
            assert np.array_equal(got_map[j], maps[k]), k                         # == the reference's Map.static_map
            if want_edf:
                want = np.sqrt(d2[k].astype(np.float64)) * cell
                np.testing.assert_allclose(got_edf[j], want, rtol=EDF_RTOL, atol=0, err_msg="edf map %d" % k)
    # the calls after the first (which may load the module lazily) took no device memory
    assert used[1] == 0 and used[2] == 0, used
    # the handle's own map is untouched: set_obstacles' products still drive the sensors
    assert sim.has_map
    # limits are refused with a message, nothing launched
    rects, n = _rect_batch(z, [0, 1, 2, 0, 1])
    import torch
    r = torch.from_numpy(rects).cuda()
    nr = torch.from_numpy(n).cuda()
    out = torch.zeros((5, 300, 300), dtype=torch.uint8, device="cuda")
    rc = sim.lib.casim_rasterize_noalloc(sim.handle, _lib.ptr(r), _lib.ptr(nr), MAX_MAPS + 1, r.shape[1], _lib.ptr(out),
                                         None, None)
    assert rc == 10001 and b"CASIM_NOALLOC_MAX_MAPS" in sim.lib.casim_last_error()
    rc = sim.lib.casim_rasterize_noalloc(sim.handle, _lib.ptr(r), _lib.ptr(nr), 1, MAX_RECTS + 1, _lib.ptr(out), None, None)
    assert rc == 10001 and b"CASIM_NOALLOC_MAX_RECTS" in sim.lib.casim_last_error()
    rc = sim.lib.casim_rasterize_noalloc(sim.handle, _lib.ptr(r), _lib.ptr(nr), 1, r.shape[1], None, None, None)
    assert rc == 10001
    sim.close()
