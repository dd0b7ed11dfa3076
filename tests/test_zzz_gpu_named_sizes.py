"""GPU: parity at the sizes BASELINE.json names, where row pitch, grid dimensions and 64-bit
offsets differ from the small boxes of the other tests:

  * D3Q27 256^3 and D2Q9 4096^2: 5 steps against the oracle, 1e-12 of max|f|;
  * D3Q19 512^3 (config 4's block per GPU; 514^3 x 19 values overflow the reference's int
    indexing, src/memory.h:187-188, so no CPU code can run it whole): a state that depends
    on y and z with period 8 must evolve into the 512 x 8 x 8 result repeated 64 x 64
    times, bit for bit, and that small result is checked against the oracle; plus mass.
  * binary 128^3 with the 7-point gradient, 100 steps: tests/test_regression_logs.py, case
    named-spin-lb1-7pt-128 (golden log made by the unmodified reference executable).

Input: a seeded block with pairwise coprime edges (37 x 41 x 43), repeated and cropped --
generated in seconds, and no two sites a power of two apart see the same data.
"""

import numpy as np
import pytest

import lbinit
from goldenlib import relerr

pytestmark = [pytest.mark.gpu]

TOL = 1e-12


def mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                return float(line.split()[1])/1048576.0
    except OSError:
        pass
    return 0.0


def tiled_state(cv, wv, nlocal, ndim):
    from ludwig_b200 import synthetic
    return synthetic.tiled_state(cv, wv, nlocal, ndim=ndim)


@pytest.mark.parametrize("nvel,nlocal", [(27, (256, 256, 256)), (9, (4096, 4096, 1))])
def test_velocity_set_configs_vs_oracle(nvel, nlocal):
    """BASELINE config 5: D3Q27 256^3 and D2Q9 4096^2, five steps of the time loop."""
    import ludwig_b200 as lb
    from oracle import pyoracle
    if mem_available_gb() < 48.0:
        pytest.skip("needs 48 GB of host memory for the oracle's copy of the lattice")
    nsteps = 5
    orc = pyoracle.Oracle(nvel, nlocal)
    f0 = tiled_state(orc.cv(), orc.wv(), nlocal, orc.ndim)
    prm = pyoracle.make_param(eta_shear=0.1, eta_bulk=0.1)
    lat = lb.Lattice(nvel, nlocal)
    lat.set_param(eta_shear=0.1, eta_bulk=0.1)
    lat.lb_memcpy_h2d(f0)
    lat.step(nsteps)
    fd = lat.lb_memcpy_d2h()
    st = lat.stats()
    lat.close()
    f, fp = f0, np.zeros_like(f0)
    n = orc.nsite
    rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    nall = orc.nall
    mass0 = f0.reshape(nall + (nvel,))[1:-1, 1:-1, 1:-1].sum()
    orc.step(prm, nsteps, f, fp, rho, u, frc)
    a = fd.reshape(nall + (nvel,))[1:-1, 1:-1, 1:-1]
    b = f.reshape(nall + (nvel,))[1:-1, 1:-1, 1:-1]
    scale = np.max(np.abs(b))
    err = 0.0
    for i in range(0, a.shape[0], 64):              # in slabs: no lattice-sized temporaries
        err = max(err, float(np.max(np.abs(a[i:i + 64] - b[i:i + 64]))))
    assert err/scale < TOL, err/scale
    assert abs(st[0] - mass0) < 1e-9*float(np.prod(nlocal))


def test_d3q19_512_cubed_equals_repeated_slab():
    """BASELINE config 4's block per GPU. 514^3 sites x 19 populations = 2.58e9 values: every
    plane offset beyond the first 2^31 doubles is exercised."""
    import ludwig_b200 as lb
    from oracle import pyoracle
    if mem_available_gb() < 96.0:
        pytest.skip("needs 96 GB of host memory (two host copies of a 20.6 GB lattice)")
    nvel, nsteps, per = 19, 2, 8
    small_n, big_n = (512, per, per), (512, 512, 512)
    orc = pyoracle.Oracle(nvel, small_n)
    fs0 = lbinit.make_state(orc.cv(), orc.wv(), small_n)          # periodic images in its halos
    sall = orc.nall
    si = np.ascontiguousarray(fs0.reshape(sall + (nvel,))[1:-1, 1:-1, 1:-1])

    def run(nlocal, f0):
        lat = lb.Lattice(nvel, nlocal)
        lat.set_param(eta_shear=0.1, eta_bulk=0.1)
        lat.lb_memcpy_h2d(f0)
        lat.step(nsteps)
        out = lat.lb_memcpy_d2h()
        st = lat.stats()
        lat.close()
        return out, st

    # the slab on the GPU and on the oracle
    gs, _ = run(small_n, fs0)
    gsi = np.ascontiguousarray(gs.reshape(sall + (nvel,))[1:-1, 1:-1, 1:-1])
    prm = pyoracle.make_param(eta_shear=0.1, eta_bulk=0.1)
    f, fp = fs0.copy(), np.zeros_like(fs0)
    n = orc.nsite
    rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    orc.step(prm, nsteps, f, fp, rho, u, frc)
    m = orc.interior_mask()
    assert relerr(gs[m], f[m]) < TOL

    # the cube: the slab's data repeated 64 x 64 times across y and z
    ball = tuple(x + 2 for x in big_n)
    big = np.zeros(ball + (nvel,))
    nrep = big_n[1]//per
    for jy in range(nrep):
        for kz in range(nrep):
            big[1:-1, 1 + per*jy:1 + per*(jy + 1), 1 + per*kz:1 + per*(kz + 1)] = si
    mass0 = float(si.sum())*nrep*nrep
    gb, st = run(big_n, big.reshape(-1, 1, nvel))
    del big
    gb = gb.reshape(ball + (nvel,))
    for jy in range(nrep):
        for kz in range(nrep):
            blk = gb[1:-1, 1 + per*jy:1 + per*(jy + 1), 1 + per*kz:1 + per*(kz + 1)]
            assert np.array_equal(blk, gsi), (jy, kz)
    assert abs(st[0] - mass0) < 1e-9*float(np.prod(big_n))


@pytest.mark.parametrize("nvel,nlocal,ndist", [(19, (21, 10, 12), 2), (9, (40, 33, 1), 1)])
def test_upload_in_pieces_equals_upload_whole(nvel, nlocal, ndist):
    """b200lb_f_upload_range (lb_io_read through a buffer): the pieces of
    synthetic.tiled_slabs give the lattice lb_memcpy H2D of the whole array gives."""
    import ludwig_b200 as lb
    from ludwig_b200 import models_host, synthetic
    cv, wv = models_host.cv(nvel), models_host.wv(nvel)
    ndim = 2 if nvel == 9 else 3
    whole = synthetic.tiled_state(cv, wv, nlocal, ndist=ndist, ndim=ndim, block=(7, 5, 3))
    out = []
    for pieces in (False, True):
        lat = lb.Lattice(nvel, nlocal, ndist=ndist)
        if pieces:
            for first, part in synthetic.tiled_slabs(cv, wv, nlocal, ndist=ndist, ndim=ndim,
                                                     block=(7, 5, 3), planes=4):
                lat.lb_memcpy_h2d_range(part, first)
        else:
            lat.lb_memcpy_h2d(whole)
        lat.step(3)
        out.append(lat.lb_memcpy_d2h())
        lat.close()
    assert np.array_equal(out[0], out[1])
