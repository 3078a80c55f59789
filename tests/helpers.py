"""Shared helpers for the parity tests."""
import os

import numpy as np

from casim import layout as L

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# continuous tolerance stated by BASELINE.json north_star: 1e-5 relative (fp32 vs the reference's fp64)
RTOL = 1e-5
ATOL = 1e-5     # absolute floor for values near zero (angles, ego-frame projections)

CONT_WORDS = [k for k in range(L.STATE_WORDS) if k not in L.INT_WORDS]


def load(name):
    return np.load(os.path.join(GOLDEN, name))


def unpack_maps():
    z = load("maps.npz")
    maps = {0: None}
    rects = {0: np.zeros((0, 4))}
    for mid in (1, 2):
        maps[mid] = np.unpackbits(z["map_%d" % mid])[:300 * 300].reshape(300, 300).astype(np.uint8)
        rects[mid] = z["rects_%d" % mid]
    return maps, rects


def assert_close(a, b, what, rtol=RTOL, atol=ATOL):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    err = np.abs(a - b)
    err[both_inf] = 0
    tol = atol + rtol * np.abs(b)
    tol[both_inf] = 1
    bad = ~(err <= tol)
    if bad.any():
        i = np.argwhere(bad)[0]
        raise AssertionError("%s: %d/%d mismatches, first at %s: %r vs %r" %
                             (what, bad.sum(), bad.size, tuple(i), a[tuple(i)], b[tuple(i)]))


def assert_state_close(s, ref, n_agents, what):
    """continuous words within tolerance, integer words (step_num, flags, type) exact"""
    for e in range(s.shape[0]):
        n = int(n_agents[e])
        for k in L.INT_WORDS:
            assert np.array_equal(s[e, :n, k], ref[e, :n, k]), (what, "word", k, "env", e, s[e, :n, k], ref[e, :n, k])
        # angles may differ by the wrap convention only within tolerance; compare directly
        assert_close(s[e, :n][:, CONT_WORDS], ref[e, :n][:, CONT_WORDS], "%s env %d" % (what, e))
