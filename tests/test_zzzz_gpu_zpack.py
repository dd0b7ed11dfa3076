"""GPU: the packed z-face halo of peer-memory sweeps (KParams::zoff, option zpack).

With the GPU as its own neighbour (option loopback + b200lb_p2p_import(nranks = 1)) the sweeps
push what crosses a z face into dense planes behind the lattice and read their own z-face halo
from there. Every run is compared bit for bit with the run in which all directions wrap within
the rank (no halo, no push), and every way out of the packed state -- a download, a propagation
on its own, a full halo exchange, switching the option off -- with the row-end halo words the
unpacked path leaves. On several GPUs the same kernels run with the neighbours' pointers:
tests/mgpu_worker.py (option zpack is on by default there)."""

import numpy as np
import pytest

import lbinit

pytestmark = pytest.mark.gpu


def lattice(nvel, nlocal, mask, zpack, fused=1, nrelax=0):
    import ludwig_b200 as lb
    from ludwig_b200 import models_host
    lat = lb.Lattice(nvel, nlocal)
    lat.set_param(eta_shear=0.1, eta_bulk=0.07, nrelax=nrelax, force=(1e-4, -2e-4, 3e-4))
    lat.set_option("fused", fused)
    if mask:
        lat.set_option("halo", lb.HALO_REDUCED)
        lat.set_option("loopback", 8 | mask)
        lat.set_option("zpack", zpack)
        lat.p2p_import(None, 1)
        assert lat.get_option("p2p_ready") == 1
    return lat


def interior(nvel, nlocal):
    m = np.zeros((nlocal[0] + 2, nlocal[1] + 2, nlocal[2] + 2), dtype=bool)
    m[1:-1, 1:-1, 1:-1] = True
    return m.reshape(-1)


def state(nvel, nlocal):
    from ludwig_b200 import models_host
    return lbinit.make_state(models_host.cv(nvel), models_host.wv(nvel), nlocal,
                             ndim=2 if nvel == 9 else 3)


@pytest.mark.parametrize("nvel,nlocal,mask", [
    (19, (12, 10, 8), 7), (19, (8, 7, 70), 4), (19, (6, 9, 16), 5), (19, (5, 6, 130), 6),
    (19, (2, 2, 2), 7), (27, (8, 6, 10), 7), (27, (5, 7, 6), 4), (9, (16, 12, 1), 3), (9, (7, 33, 1), 2)])
@pytest.mark.parametrize("fused", [1, 0])
@pytest.mark.parametrize("nrelax", [0, 2])
def test_packed_planes_equal_wrap(nvel, nlocal, mask, fused, nrelax):
    if nrelax == 2 and nvel != 19:
        nrelax = 1            # TRT is D3Q19's (as in the reference); BGK keeps ghost modes too
    f0 = state(nvel, nlocal)
    m = interior(nvel, nlocal)
    ref = lattice(nvel, nlocal, 0, 0, fused, nrelax)
    ref.lb_memcpy_h2d(f0)
    ref.step(7)
    want = ref.lb_memcpy_d2h()
    ref.step(7)
    want14 = ref.lb_memcpy_d2h()
    ref.close()
    got = {}
    for zpack in (0, 1, 2):
        lat = lattice(nvel, nlocal, mask, zpack, fused, nrelax)
        lat.lb_memcpy_h2d(f0)
        lat.step(7)
        # the contiguous device axis is z in three dimensions, y in two
        zdec = (mask & 4) if nvel != 9 else (mask & 2)
        if fused:
            assert (lat.get_option("zpacked") != 0) == bool(zpack and zdec)
        got[zpack] = lat.lb_memcpy_d2h()
        assert lat.get_option("zpacked") == 0
        if zpack:
            # switching the arrangement between two sweeps goes through the halo words
            lat.step(2)
            lat.set_option("zpack", 3 - zpack)
            lat.step(2)
            lat.set_option("zpack", 0)
            lat.step(3)
            more = lat.lb_memcpy_d2h()
        lat.sync()
        lat.close()
        assert np.array_equal(got[zpack][m], want[m]), "zpack %d" % zpack
        if zpack:
            assert np.array_equal(more[m], want14[m]), "zpack %d and on" % zpack


def test_every_way_out_of_the_packed_state():
    import ludwig_b200 as lb
    nvel, nlocal = 19, (10, 6, 12)
    f0 = state(nvel, nlocal)
    m = interior(nvel, nlocal)

    def run(zpack, how):
        lat = lattice(nvel, nlocal, 7, zpack)
        lat.lb_memcpy_h2d(f0)
        lat.step(3)
        if how == "collide_then_download":
            # post-collision lattice, halo sites as pushed: observed without a propagation
            lat.set_option("fused", 0)
            lat.lb_collide()
            out = lat.lb_memcpy_d2h().reshape(-1)
        elif how == "full_halo":
            lat.lb_collide()
            lat.lb_halo_swap(lb.HALO_FULL)
            lat.lb_propagation()
            out = lat.lb_memcpy_d2h().reshape(-1)
        elif how == "option_off":
            lat.lb_collide()
            lat.lb_halo()
            lat.lb_propagation()
            lat.set_option("zpack", 0)
            lat.step(2)
            out = lat.lb_memcpy_d2h().reshape(-1)
        elif how == "p2p_off":
            lat.lb_collide()
            lat.lb_halo()
            lat.lb_propagation()
            lat.set_option("p2p", 0)
            lat.step(2)
            out = lat.lb_memcpy_d2h().reshape(-1)
        elif how == "stats_then_on":
            st = lat.stats()
            lat.step(2)
            out = np.concatenate([lat.lb_memcpy_d2h().reshape(-1), st])
        lat.sync()
        lat.close()
        return out

    for how, zpack in [(h, z) for h in ("collide_then_download", "full_halo", "option_off", "p2p_off",
                                         "stats_then_on") for z in (1, 2)]:
        a, b = run(0, how), run(zpack, how)
        if how == "collide_then_download":
            assert np.array_equal(a, b), how          # halo sites too
        else:
            n = m.size * nvel
            am, bm = a[:n].reshape(m.size, nvel), b[:n].reshape(m.size, nvel)
            assert np.array_equal(am[m], bm[m]), how
            assert np.array_equal(a[n:], b[n:]), how


def test_packed_planes_refused_where_they_do_not_apply():
    """Binary fluid, walls and the single lattice keep the row-end halo words: the option
    stays on and is simply not used."""
    import ludwig_b200 as lb
    from ludwig_b200 import models_host
    nlocal = (8, 8, 6)
    lat = lb.Lattice(19, nlocal, ndist=2)
    lat.set_param(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75, a=-0.00625, b=0.00625, kappa=0.004)
    lat.set_option("halo", lb.HALO_REDUCED)
    lat.set_option("loopback", 1)
    lat.p2p_import(None, 1)
    f0 = lbinit.make_state(models_host.cv(19), models_host.wv(19), nlocal, ndist=2)
    lat.lb_memcpy_h2d(f0)
    lat.step(3)
    assert lat.get_option("zpack") == 1 and lat.get_option("zpacked") == 0
    lat.sync()
    lat.close()
