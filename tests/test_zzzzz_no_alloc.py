"""GPU: b200lb_field_stats_noalloc, the one entry point that allocates nothing (static
device scratch; no counterpart in the reference),
against b200lb_field_stats through the C ABI: same kernels, so the two must
agree to the bit, and certainly within the suite's tolerance; on two calls in a row whose
data and lattice shape differ, so stale scratch of the first call would show in the second.

Collected last on purpose (a failure here hides nothing else).
"""

import ctypes

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu]

TOL = 1e-12


def field(nlocal, seed, scale, nf=1):
    """[nsite(, nf)] host array in the reference's layout, random in the interior."""
    rng = np.random.default_rng(seed)
    nall = tuple(n + 2 for n in nlocal)
    a = np.zeros(nall + ((nf,) if nf > 1 else ()))
    a[1:-1, 1:-1, 1:-1] = scale*(rng.random(tuple(nlocal) + ((nf,) if nf > 1 else ())) - 0.3)
    return np.ascontiguousarray(a.reshape((-1, nf) if nf > 1 else (-1,)))


def free_bytes():
    """cudaMemGetInfo through the runtime, or None where it cannot be loaded."""
    for name in ("libcudart.so.12", "libcudart.so", "/usr/local/cuda/lib64/libcudart.so.12"):
        try:
            rt = ctypes.CDLL(name)
        except OSError:
            continue
        free, total = ctypes.c_size_t(0), ctypes.c_size_t(0)
        if rt.cudaMemGetInfo(ctypes.byref(free), ctypes.byref(total)) == 0:
            return free.value
    return None


def agree(a, b):
    scale = max(1.0, float(np.max(np.abs(a))))
    assert np.max(np.abs(a - b)) <= TOL*scale, (a, b)


def test_no_alloc_matches_field_stats_two_calls_in_a_row():
    import ludwig_b200 as lb
    # first call: many blocks, large values
    big = lb.Lattice(19, (24, 21, 140), ndist=2)
    phi = field((24, 21, 140), 11, 1000.0)
    big.field_h2d(lb.FIELD_PHI, phi)
    want = big.field_stats(lb.FIELD_PHI)
    before = free_bytes()
    got = big.field_stats_noalloc(lb.FIELD_PHI)
    after = free_bytes()
    agree(want, got)
    assert np.array_equal(want, got)                  # same kernels, same order of the fold
    m = np.zeros((26, 23, 142), dtype=bool)
    m[1:-1, 1:-1, 1:-1] = True
    v = phi[m.reshape(-1)]
    agree(np.array([[v.sum(), (v*v).sum(), v.min(), v.max()]]), got)
    if before is not None:
        assert after == before, "the no-alloc entry point changed the free device memory"

    # second call straight after: fewer blocks, small values, three components -- whatever
    # the first call left in the scratch must not reach the result
    small = lb.Lattice(19, (5, 3, 6))
    u = field((5, 3, 6), 12, 1.0e-3, nf=3)
    small.field_h2d(lb.FIELD_U, u)
    want = small.field_stats(lb.FIELD_U)
    got = small.field_stats_noalloc(lb.FIELD_U)
    agree(want, got)
    assert np.array_equal(want, got)
    m = np.zeros((7, 5, 8), dtype=bool)
    m[1:-1, 1:-1, 1:-1] = True
    for c in range(3):
        v = u[m.reshape(-1), c]
        agree(np.array([v.sum(), (v*v).sum(), v.min(), v.max()]), got[c])

    # and the first lattice again with new data: the array is shared by every handle
    phi2 = field((24, 21, 140), 13, 0.5)
    big.field_h2d(lb.FIELD_PHI, phi2)
    agree(big.field_stats(lb.FIELD_PHI), big.field_stats_noalloc(lb.FIELD_PHI))
    small.close()
    big.close()


def test_no_alloc_refuses_what_it_cannot_do_without_allocating():
    import ludwig_b200 as lb
    lat = lb.Lattice(19, (6, 6, 6))
    with pytest.raises(lb.B200LBError, match="allocates nothing"):
        lat.field_stats_noalloc(lb.FIELD_RHO)         # planes never created
    f0 = np.zeros((8*8*8, 1, 19))
    f0[:, 0, :] = 1.0/19.0
    lat.lb_memcpy_h2d(f0)
    lat.set_option("hydro_write", 0)
    lat.field_h2d(lb.FIELD_U, field((6, 6, 6), 1, 0.01, nf=3))      # planes exist from here on
    lat.step(1)                                       # rho,u of this collision not stored
    with pytest.raises(lb.B200LBError, match="lazily"):
        lat.field_stats_noalloc(lb.FIELD_U)
    lat.field_stats(lb.FIELD_U)                       # the existing entry point materialises them
    agree(lat.field_stats(lb.FIELD_U), lat.field_stats_noalloc(lb.FIELD_U))
    lat.close()
    # one block of 128 x 2 sites per x-plane and row pair: 65537 x-planes are one too many
    lat = lb.Lattice(19, (65537, 2, 2), ndist=2)
    with pytest.raises(lb.B200LBError, match="B200LB_NOALLOC_MAXBLOCK"):
        lat.field_stats_noalloc(lb.FIELD_PHI)
    lat.close()
