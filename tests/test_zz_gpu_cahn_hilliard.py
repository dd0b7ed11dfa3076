"""GPU parity of the finite-difference order parameter ("free_energy symmetric"):
b200lb_phi_force + b200lb_cahn_hilliard (ludwig_b200/csrc/ch_kernels.cu) through the C ABI
against the reference-produced golden vectors (tests/golden/ref_ch.npz) and the CPU oracle.

Bars: with option ch_exact the kernel build without fused multiply-add, which keeps the
reference's order of operations, gives the force on the fluid and the updated order
parameter BIT-EXACT against the reference; the default (fast) build agrees to a few ulp;
the time loop (collision in between) within 1e-12; the order parameter is conserved to
round-off in both.

(The file name sorts last on purpose: these are the newest kernels.)
"""

import numpy as np
import pytest

import goldenlib
import lbinit
from goldenlib import relerr

pytestmark = pytest.mark.gpu

P = goldenlib.CH_PHYS
TAGS = sorted(goldenlib.CH_CASES)


def mk(nlocal, order, method, gmu, npt, **opts):
    import ludwig_b200 as lb
    lat = lb.Lattice(19, nlocal)
    for k, v in opts.items():
        lat.set_option(k, v)
    lat.set_param(eta_shear=P["eta"], eta_bulk=P["eta"], mobility=P["mobility"],
                  a=P["a"], b=P["b"], kappa=P["kappa"])
    lat.phi_ch_create(order, method, gmu)
    lat.field_grad_set(npt)
    return lat


def interior(nlocal):
    m = np.zeros(tuple(n + 2 for n in nlocal), dtype=bool)
    m[1:-1, 1:-1, 1:-1] = True
    return m.reshape(-1)


def close(a, b, exact):
    if exact:
        return np.array_equal(a, b)
    return np.max(np.abs(a - b)) <= 1e-15 * np.max(np.abs(b))


@pytest.mark.parametrize("tag", TAGS)
@pytest.mark.parametrize("mode", ["fused", "two_launches", "loopback"])
@pytest.mark.parametrize("exact", [1, 0])
def test_force_and_update_vs_reference(tag, mode, exact):
    """One phi_force_calculation + phi_cahn_hilliard on seeded phi and u: with ch_exact
    force and new phi equal the reference's bit for bit, whether the two are one launch
    or two, and when every direction goes through the multi-rank exchange (loopback);
    the default build (fused multiply-add) within 1e-15 of the largest value."""
    import ludwig_b200 as lb
    g = goldenlib.load_ch()
    nlocal, order, method, npt, gmu = goldenlib.CH_CASES[tag]
    opts = {"ch_exact": exact}
    if mode == "two_launches":
        opts["ch_fuse"] = 0
    if mode == "loopback":
        opts["loopback"] = 1
    lat = mk(nlocal, order, method, gmu, npt, **opts)
    m = interior(nlocal)
    phi, u = goldenlib.ch_fields(nlocal)
    lat.field_h2d(lb.FIELD_PHI, phi)
    lat.field_h2d(lb.FIELD_U, u)
    lat.hydro_f_zero()
    lat.field_halo()
    lat.field_grad_compute()
    n0 = lat.launch_count()
    lat.phi_force_calculation()
    if mode == "loopback" and order == 3:
        # phi two sites away is only reachable around directions that wrap within the rank
        with pytest.raises(lb.B200LBError, match="third-order"):
            lat.phi_cahn_hilliard()
        lat.close()
        return
    lat.phi_cahn_hilliard()
    if mode == "fused":
        assert lat.launch_count() - n0 == 1          # one pass over the lattice for both
    force = lat.field_d2h(lb.FIELD_FORCE)
    phi1 = lat.field_d2h(lb.FIELD_PHI)
    assert close(force[m], g[tag + "_force"], exact)
    assert close(phi1[m], g[tag + "_phi1"], exact)
    assert abs(phi1[m].sum() - phi[m].sum()) < 1e-13 * m.sum()
    lat.close()


@pytest.mark.parametrize("tag", TAGS)
@pytest.mark.parametrize("hydro_write,exact", [(1, 0), (0, 0), (1, 1)])
def test_time_loop_vs_reference(tag, hydro_write, exact):
    """Eight steps of the src/ludwig.c loop (b200lb_step) against the reference's output
    and the oracle: populations 1e-12, order parameter 1e-13."""
    import ludwig_b200 as lb
    from oracle import pyoracle
    g = goldenlib.load_ch()
    nlocal, order, method, npt, gmu = goldenlib.CH_CASES[tag]
    lat = mk(nlocal, order, method, gmu, npt, hydro_write=hydro_write, ch_exact=exact)
    m = interior(nlocal)
    orc = pyoracle.Oracle(19, nlocal)
    phi, _ = goldenlib.ch_fields(nlocal)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    lat.lb_memcpy_h2d(f0)
    lat.field_h2d(lb.FIELD_PHI, phi)
    lat.step(8)
    f = lat.lb_memcpy_d2h()
    assert relerr(f[m], g[tag + "_step8_f"]) < 1e-12
    assert relerr(lat.field_d2h(lb.FIELD_PHI)[m], g[tag + "_step8_phi"]) < 1e-13
    assert relerr(lat.field_d2h(lb.FIELD_U)[m], g[tag + "_step8_u"]) < 1e-11
    # and the oracle, run here on the same input
    prm = pyoracle.make_param(eta_shear=P["eta"], eta_bulk=P["eta"], mobility=P["mobility"],
                              a=P["a"], b=P["b"], kappa=P["kappa"])
    n = orc.nsite
    fo, fp = f0.copy(), np.zeros_like(f0)
    rho, u, force = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    grad, delsq = np.zeros((n, 3)), np.zeros(n)
    orc.gradient_select(npt)
    orc.ch_attach(orc.ch(order, method, gmu))
    try:
        orc.step(prm, 8, fo, fp, rho, u, force, phi, grad, delsq)
    finally:
        orc.ch_attach(None)
        orc.gradient_select(7)
    assert relerr(f[m], fo[m]) < 1e-12
    lat.close()


def test_call_sequence_equals_step():
    """The reference's calls one at a time (src/ludwig.c:537-575, 727-768, 791-862), unfused
    propagation and eager halos, give the same lattice as b200lb_step bit for bit."""
    import ludwig_b200 as lb
    from ludwig_b200 import models_host
    nlocal, order, method, npt, gmu = goldenlib.CH_CASES["o2s7"]
    phi, _ = goldenlib.ch_fields(nlocal)
    f0 = lbinit.make_state(models_host.cv(19), models_host.wv(19), nlocal)
    out = []
    for manual in (0, 1):
        # ch_exact: the one-launch and two-launch kernels are different instantiations;
        # without fused multiply-add they cannot differ
        opts = dict(fused=0, halo_lazy=0, ch_fuse=0, ch_exact=1) if manual else dict(ch_exact=1)
        lat = mk(nlocal, order, method, gmu, npt, **opts)
        lat.lb_memcpy_h2d(f0)
        lat.field_h2d(lb.FIELD_PHI, phi)
        if manual:
            for _ in range(5):
                lat.hydro_f_zero()
                lat.field_halo()
                lat.field_grad_compute()
                lat.phi_force_calculation()
                lat.phi_cahn_hilliard()
                lat.hydro_u_zero()
                lat.lb_collide()
                lat.lb_halo()
                lat.lb_propagation()
        else:
            lat.step(5)
        out.append((lat.lb_memcpy_d2h(), lat.field_d2h(lb.FIELD_PHI)))
        lat.close()
    m = interior(nlocal)
    assert np.array_equal(out[0][0][m], out[1][0][m])
    assert np.array_equal(out[0][1][m], out[1][1][m])


def test_force_seen_by_the_host_and_without_update():
    """phi_force_calculation recorded lazily is evaluated when the host asks for the force
    field, and by lb_collide when no Cahn-Hilliard update follows."""
    import ludwig_b200 as lb
    g = goldenlib.load_ch()
    tag = "o1s27"
    nlocal, order, method, npt, gmu = goldenlib.CH_CASES[tag]
    m = interior(nlocal)
    phi, u = goldenlib.ch_fields(nlocal)
    lat = mk(nlocal, order, method, gmu, npt, ch_exact=1)
    lat.field_h2d(lb.FIELD_PHI, phi)
    lat.field_h2d(lb.FIELD_U, u)
    lat.hydro_f_zero()
    lat.field_halo()
    lat.field_grad_compute()
    lat.phi_force_calculation()
    assert np.array_equal(lat.field_d2h(lb.FIELD_FORCE)[m], g[tag + "_force"])
    # a constant added by hydro_f_zero is part of the field the force lands on
    lat.hydro_f_zero((1e-3, 0.0, -2e-3))
    lat.phi_force_calculation()
    frc = lat.field_d2h(lb.FIELD_FORCE)[m]
    assert np.array_equal(frc, np.array([1e-3, 0.0, -2e-3]) + g[tag + "_force"])
    # the order parameter has not moved
    assert np.array_equal(lat.field_d2h(lb.FIELD_PHI)[m], phi[m])
    lat.close()


def test_refusals():
    """What is not on this path fails loudly, as a b200lb_last_error message."""
    import ludwig_b200 as lb
    lat = lb.Lattice(19, (8, 8, 8), ndist=2)
    with pytest.raises(lb.B200LBError, match="ndist = 1"):
        lat.phi_ch_create(1)
    lat.close()
    lat = lb.Lattice(19, (8, 8, 8), periodic=(1, 0, 1))
    with pytest.raises(lb.B200LBError, match="walls"):
        lat.phi_ch_create(1)
    lat.close()
    lat = lb.Lattice(19, (8, 8, 8))
    with pytest.raises(lb.B200LBError, match="b200lb_ch_create"):
        lat.phi_force_calculation()
    with pytest.raises(lb.B200LBError, match="orders 1 to 4"):
        lat.phi_ch_create(5)
    lat.phi_ch_create(1)
    with pytest.raises(lb.B200LBError, match="field_grad_compute"):
        lat.phi_force_calculation()
    lat.set_option("le_fd", 0)
    with pytest.raises(lb.B200LBError, match="Lees-Edwards"):
        lat.le_commit(2, 0.01)
    lat.close()


def test_conservation_128():
    """128^3 spinodal start, 20 steps, third-order advection: total phi and mass to
    round-off (the two sites of a face compute identical fluxes), phi stays bounded."""
    import ludwig_b200 as lb
    from ludwig_b200 import models_host
    nt = (128, 128, 128)
    lat = mk(nt, 3, lb.FORCE_STRESS_DIVERGENCE, None, 27)
    rng = np.random.default_rng(8361235)
    nall = tuple(n + 2 for n in nt)
    phi = np.zeros(nall)
    phi[1:-1, 1:-1, 1:-1] = 0.1 * (rng.random(nt) - 0.5)
    f = np.zeros(nall + (1, 19))
    f[1:-1, 1:-1, 1:-1, 0, :] = models_host.wv(19)
    lat.lb_memcpy_h2d(np.ascontiguousarray(f.reshape(-1, 1, 19)))
    lat.field_h2d(lb.FIELD_PHI, np.ascontiguousarray(phi.reshape(-1)))
    s0 = lat.field_stats(lb.FIELD_PHI)[0]
    lat.step(20)
    s1 = lat.field_stats(lb.FIELD_PHI)[0]
    st = lat.stats()
    vol = float(np.prod(nt))
    assert abs(s1[0] - s0[0]) < 1e-16 * vol * 20
    assert abs(st[0] - vol) < 1e-6
    assert s1[2] >= -0.051 and s1[3] <= 0.051 and s1[1] < s0[1]     # coarsening smooths the noise
    lat.close()
