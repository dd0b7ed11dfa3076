"""GPU parity of the viscosity model: lb_collide with a per-site viscosity
(src/collision.c:386-404, hydro->eta) and viscosity_model arrhenius (src/visc_arrhenius.c),
through the C ABI against the CPU oracle (pinned against the reference build:
tests/test_oracle_vs_ref.py::test_viscosity_model_*).

Bars: eta from the device pow within 1e-14 relative of the host's; populations 1e-12.
"""

import numpy as np
import pytest

import goldenlib
import lbinit
from goldenlib import relerr

pytestmark = pytest.mark.gpu

VISC = dict(eta_plus=0.625, eta_minus=0.00625, phistar=1.0)
PHYS = dict(eta_shear=0.00625, eta_bulk=0.0125, mobility=1.25, a=-0.00625, b=0.00625, kappa=0.004)


def setup(nlocal, nrelax=0, **opts):
    import ludwig_b200 as lb
    from oracle import pyoracle
    lat = lb.Lattice(19, nlocal)
    for k, v in opts.items():
        lat.set_option(k, v)
    lat.set_param(nrelax=nrelax, **PHYS)
    orc = pyoracle.Oracle(19, nlocal)
    prm = pyoracle.make_param(nrelax=nrelax, **PHYS)
    return lat, orc, prm


@pytest.mark.parametrize("nrelax", [0, 1, 2])
@pytest.mark.parametrize("fused", [1, 0])
def test_collision_with_site_viscosity(nrelax, fused):
    """visc update + one collision (after a propagation when fused): every scheme, the
    ghost rates of BGK and TRT following the site's viscosity."""
    import ludwig_b200 as lb
    nlocal = (12, 10, 8)
    lat, orc, prm = setup(nlocal, nrelax, fused=fused)
    n, m = orc.nsite, orc.interior_mask()
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    force = lbinit.make_force(nlocal)
    phi, _ = goldenlib.ch_fields(nlocal)
    phi *= 2.0
    lat.phi_ch_create(1)
    lat.lb_memcpy_h2d(f0)
    lat.field_h2d(lb.FIELD_FORCE, force)
    lat.field_h2d(lb.FIELD_PHI, phi)
    lat.visc_arrhenius_commit(**VISC)
    lat.visc_update()
    eta = np.zeros(n)
    orc.visc_arrhenius_update(VISC["eta_plus"], VISC["eta_minus"], VISC["phistar"], phi, eta)
    eta_dev = lat.field_d2h(lb.FIELD_ETA)
    assert relerr(eta_dev[m], eta[m]) < 1e-14
    f, fp = f0.copy(), np.zeros_like(f0)
    rho, u = np.zeros(n), np.zeros((n, 3))
    if fused:
        # collide, halo, propagate, then the fused stream-and-collide pass
        lat.lb_collide(); lat.lb_halo(); lat.lb_propagation()
        lat.lb_collide()
        orc.visc_attach(eta)
        try:
            orc.collide_mrt(prm, f, force, rho, u)
            orc.halo(f, 19)
            orc.propagate(f, fp)
            orc.collide_mrt(prm, fp, force, rho, u)
        finally:
            orc.visc_attach(None)
        f = fp
    else:
        lat.lb_collide()
        orc.visc_attach(eta)
        try:
            orc.collide_mrt(prm, f, force, rho, u)
        finally:
            orc.visc_attach(None)
    fd = lat.lb_memcpy_d2h()
    assert relerr(fd[m], f[m]) < 1e-12
    assert np.max(np.abs(lat.field_d2h(lb.FIELD_U)[m] - u[m])) < 1e-14
    # the uniform-viscosity collision is something else
    g, r2, u2 = f0.copy(), np.zeros(n), np.zeros((n, 3))
    orc.collide_mrt(prm, g, force, r2, u2)
    if not fused:
        assert relerr(fd[m], g[m]) > 1e-6
    lat.close()


def test_viscosity_from_the_host():
    """A model evaluated by the caller: eta uploaded as B200LB_FIELD_ETA, visc_model_set."""
    import ludwig_b200 as lb
    nlocal = (9, 7, 11)
    lat, orc, prm = setup(nlocal)
    n, m = orc.nsite, orc.interior_mask()
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    force = np.zeros((n, 3))
    ids = lbinit.global_ids(nlocal).reshape(-1)
    eta = 0.05 + 0.04 * lbinit.uniform(ids, 33)
    lat.lb_memcpy_h2d(f0)
    lat.visc_model_set(1)
    lat.field_h2d(lb.FIELD_ETA, eta)
    lat.lb_collide()
    f = f0.copy()
    rho, u = np.zeros(n), np.zeros((n, 3))
    orc.visc_attach(eta)
    try:
        orc.collide_mrt(prm, f, force, rho, u)
    finally:
        orc.visc_attach(None)
    assert relerr(lat.lb_memcpy_d2h()[m], f[m]) < 1e-13
    # switched off again: the uniform viscosity of set_param
    lat.visc_model_set(0)
    lat.lb_memcpy_h2d(f0)
    lat.lb_collide()
    g = f0.copy()
    orc.collide_mrt(prm, g, force, rho, u)
    assert relerr(lat.lb_memcpy_d2h()[m], g[m]) < 1e-13
    lat.close()


@pytest.mark.parametrize("hydro_write", [1, 0])
def test_time_loop_with_viscosity_model(hydro_write):
    """free_energy symmetric + viscosity_model arrhenius, ten steps of b200lb_step."""
    import ludwig_b200 as lb
    from oracle import pyoracle
    nlocal = (10, 8, 12)
    lat, orc, prm = setup(nlocal, hydro_write=hydro_write)
    n, m = orc.nsite, orc.interior_mask()
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    phi, _ = goldenlib.ch_fields(nlocal)
    phi *= 2.0
    lat.phi_ch_create(2, lb.FORCE_STRESS_DIVERGENCE)
    lat.field_grad_set(27)
    lat.visc_arrhenius_commit(**VISC)
    lat.lb_memcpy_h2d(f0)
    lat.field_h2d(lb.FIELD_PHI, phi)
    lat.step(10)
    f, fp = f0.copy(), np.zeros_like(f0)
    rho, u, force = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    grad, delsq, eta = np.zeros((n, 3)), np.zeros(n), np.zeros(n)
    orc.gradient_select(27)
    orc.ch_attach(orc.ch(2, pyoracle.FORCE_STRESS_DIVERGENCE))
    orc.visc_attach(eta, **VISC)
    try:
        orc.step(prm, 10, f, fp, rho, u, force, phi, grad, delsq)
    finally:
        orc.ch_attach(None)
        orc.visc_attach(None)
        orc.gradient_select(7)
    assert relerr(lat.lb_memcpy_d2h()[m], f[m]) < 1e-12
    assert relerr(lat.field_d2h(lb.FIELD_PHI)[m], phi[m]) < 1e-13
    assert relerr(lat.field_d2h(lb.FIELD_ETA)[m], eta[m]) < 1e-13
    assert relerr(lat.field_d2h(lb.FIELD_U)[m], u[m]) < 1e-11
    lat.close()


def test_refusals():
    import ludwig_b200 as lb
    lat = lb.Lattice(19, (8, 8, 8))
    with pytest.raises(lb.B200LBError, match="composition"):
        lat.visc_arrhenius_commit(**VISC)
    with pytest.raises(lb.B200LBError, match="b200lb_visc_arrhenius_commit"):
        lat.visc_update()
    lat.visc_model_set(1)
    lat.noise_init(0)
    lat.noise_present_set(1)
    lat.set_param(kt=1e-5)
    with pytest.raises(lb.B200LBError, match="fluctuations"):
        lat.lb_collide()
    lat.close()
    lat = lb.Lattice(19, (8, 8, 8), ndist=2)
    with pytest.raises(lb.B200LBError, match="single-fluid"):
        lat.visc_model_set(1)
    lat.close()
