"""GPU: single-lattice streaming (option lattices = 1, the AA pattern; SURVEY 7 step 3 and
"Hard parts: memory at 1024^3 on 2 GPUs"). One array instead of lb_t's f and fprime
(src/propagation.c:223-252 swaps two): lb_collide alternates between an in-place sweep that
leaves every population in its opposite's slot and a sweep that pulls from there, collides
and stores one propagation ahead. Everything the API shows must be what the two-lattice
path shows: bit for bit after every lb_propagation, and within 1e-12 of the oracle.
"""

import ctypes

import numpy as np
import pytest

import lbinit
from goldenlib import relerr

pytestmark = [pytest.mark.gpu]

TOL = 1e-12
CASES = [(19, (10, 7, 9)), (19, (6, 5, 34)), (27, (5, 6, 8)), (9, (14, 11, 1))]


def run(nvel, nlocal, f0, counts, lattices, vec=None, nrelax=0, hydro_write=None, force=(0.0, 0.0, 0.0)):
    import ludwig_b200 as lb
    lat = lb.Lattice(nvel, nlocal)
    lat.set_param(eta_shear=0.07, eta_bulk=0.11, nrelax=nrelax, force=force)
    if vec:
        lat.set_option("vec", vec)
    if hydro_write is not None:
        lat.set_option("hydro_write", hydro_write)
    lat.lb_memcpy_h2d(f0)
    lat.set_option("lattices", lattices)
    assert lat.get_option("lattices") == lattices
    out, done = [], 0
    for n in counts:
        lat.step(n - done)
        done = n
        out.append((lat.lb_memcpy_d2h(), lat.field_d2h(lb.FIELD_RHO), lat.field_d2h(lb.FIELD_U), lat.stats()))
    lat.close()
    return out


@pytest.mark.parametrize("nvel,nlocal", CASES)
@pytest.mark.parametrize("vec", [1, 2])
@pytest.mark.parametrize("nrelax", [0, 1])
def test_single_lattice_equals_two_lattices(nvel, nlocal, vec, nrelax):
    from ludwig_b200 import models_host
    cv, wv = models_host.cv(nvel), models_host.wv(nvel)
    ndim = 2 if nvel == 9 else 3
    f0 = lbinit.make_state(cv, wv, nlocal, ndim=ndim)
    counts = (1, 2, 3, 6, 11)             # odd and even: observed after even and after odd sweeps
    frc = (1e-4, -2e-4, 0.0 if ndim == 2 else 3e-4)
    two = run(nvel, nlocal, f0, counts, 2, vec, nrelax, force=frc)
    one = run(nvel, nlocal, f0, counts, 1, vec, nrelax, force=frc)
    nall = tuple(n + 2 for n in nlocal)
    m = np.zeros(nall, dtype=bool)
    m[1:-1, 1:-1, 1:-1] = True
    m = m.reshape(-1)
    for k, (a, b) in enumerate(zip(two, one)):
        assert np.array_equal(a[0][m], b[0][m]), "populations differ after %d steps" % counts[k]
        assert np.array_equal(a[1][m], b[1][m]), "rho differs after %d steps" % counts[k]
        assert np.array_equal(a[2][m], b[2][m]), "u differs after %d steps" % counts[k]
        assert np.array_equal(a[3], b[3])


@pytest.mark.parametrize("nvel,nlocal", CASES)
@pytest.mark.parametrize("hydro_write", [0, 1])
def test_single_lattice_vs_oracle(nvel, nlocal, hydro_write):
    from oracle import pyoracle
    orc = pyoracle.Oracle(nvel, nlocal)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndim=orc.ndim)
    nsteps = 7
    got = run(nvel, nlocal, f0, (nsteps,), 1, hydro_write=hydro_write)[0]
    prm = pyoracle.make_param(eta_shear=0.07, eta_bulk=0.11)
    f, fp = f0.copy(), np.zeros_like(f0)
    n = orc.nsite
    rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    orc.step(prm, nsteps, f, fp, rho, u, frc)
    m = orc.interior_mask()
    assert relerr(got[0][m], f[m]) < TOL
    assert relerr(got[1][m], rho[m]) < TOL
    assert np.max(np.abs(got[2][m] - u[m])) < TOL
    assert abs(got[3][0] - f0[m].sum()) < 1e-9*n


@pytest.mark.parametrize("case", ["serial-dist-3du", "serial-dist-1dp", "serial-dist-2kh",
                                  "serial-dist-2sw", "serial-2khz-bm1"])
def test_single_lattice_reproduces_reference_log(case):
    """The reference's own single-fluid regression logs through the single-lattice path:
    every number of the final statistics block as printed, within 1e-12
    (tests/test_regression_logs.py has the two-lattice runs)."""
    import ludwig_b200 as lb
    from ludwig_b200 import models_host
    import test_regression_logs as trl
    rec = trl.GOLD[case]
    nvel = rec["nvel"]
    nt, f = trl.initial_state(rec["input"], models_host.cv(nvel), models_host.wv(nvel))
    eta, zeta = trl.viscosities(rec["input"])
    lat = lb.Lattice(nvel, nt)
    lat.set_option("halo", lb.HALO_REDUCED)
    lat.set_param(eta_shear=eta, eta_bulk=zeta, force=trl.body_force(rec["input"]))
    lat.lb_memcpy_h2d(f)
    lat.set_option("lattices", 1)
    lat.step(int(rec["input"]["N_cycles"]))
    sums = lat.stats()
    u = lat.field_d2h(lb.FIELD_U)
    m = np.zeros(tuple(n + 2 for n in nt), dtype=bool)
    m[1:-1, 1:-1, 1:-1] = True
    trl.check(rec, nt, sums, u[m.reshape(-1)])
    lat.close()


def test_single_lattice_state_machine_and_refusals():
    import ludwig_b200 as lb
    from ludwig_b200 import models_host
    nvel, nlocal = 19, (8, 6, 10)
    f0 = lbinit.make_state(models_host.cv(nvel), models_host.wv(nvel), nlocal)

    def free_bytes():
        for name in ("libcudart.so.12", "libcudart.so", "/usr/local/cuda/lib64/libcudart.so.12"):
            try:
                rt = ctypes.CDLL(name)
            except OSError:
                continue
            free, total = ctypes.c_size_t(0), ctypes.c_size_t(0)
            if rt.cudaMemGetInfo(ctypes.byref(free), ctypes.byref(total)) == 0:
                return free.value
        return None

    # half the memory: the second lattice is released
    big = lb.Lattice(nvel, (64, 64, 64))
    before = free_bytes()
    big.set_option("lattices", 1)
    after = free_bytes()
    if before is not None:
        assert after - before >= 19*66*66*64*8, "the second lattice was not released"
    big.close()

    ref = run(nvel, nlocal, f0, (4, 9), 2)
    m = np.zeros(tuple(n + 2 for n in nlocal), dtype=bool)
    m[1:-1, 1:-1, 1:-1] = True
    m = m.reshape(-1)             # halo sites of a streamed lattice are stale, as in the reference
    lat = lb.Lattice(nvel, nlocal)
    lat.set_param(eta_shear=0.07, eta_bulk=0.11)
    lat.lb_memcpy_h2d(f0)
    lat.set_option("lattices", 1)
    # the calls of the time loop one by one (src/ludwig.c:802-860), with the in-sweep sums
    lat.set_option("diag", 1)
    for k in range(4):
        lat.hydro_f_zero()
        lat.hydro_u_zero()
        lat.lb_collide()
        if k == 0:
            with pytest.raises(lb.B200LBError, match="between lb_collide and lb_propagation"):
                lat.lb_memcpy_d2h()
            with pytest.raises(lb.B200LBError, match="twice"):
                lat.lb_collide()
        d = lat.step_diag()
        assert abs(d[0] - ref[0][3][0]) < 1e-9*f0.shape[0]        # mass of the state collided
        lat.lb_halo()
        lat.lb_propagation()
    with pytest.raises(lb.B200LBError, match="follows lb_collide"):
        lat.lb_propagation()
    assert np.array_equal(lat.lb_memcpy_d2h()[m], ref[0][0][m])
    # back to two lattices in mid run, and on: still the same lattice
    lat.set_option("lattices", 2)
    lat.step(2)
    lat.set_option("lattices", 1)
    lat.step(3)
    assert np.array_equal(lat.lb_memcpy_d2h()[m], ref[1][0][m])
    lat.close()

    lat = lb.Lattice(nvel, (6, 6, 6), ndist=2)
    with pytest.raises(lb.B200LBError, match="single-fluid"):
        lat.set_option("lattices", 1)
    assert lat.get_option("lattices") == 2
    lat.close()
    lat = lb.Lattice(nvel, (6, 6, 6), periodic=(1, 1, 0))
    with pytest.raises(lb.B200LBError, match="periodic"):
        lat.set_option("lattices", 1)
    lat.close()
