"""The CPU restatement against the committed reference-produced golden vectors
(tests/golden/ref_*.npz). Runs anywhere gcc is (no /root/reference needed)."""

import numpy as np
import pytest

from oracle import pyoracle
import goldenlib
import lbinit
from goldenlib import relerr


@pytest.mark.parametrize("nvel", [19, 27, 9])
def test_collide_halo_propagate(nvel):
    o0 = pyoracle.Oracle(nvel, (2, 2, 1 if nvel == 9 else 2))
    g, nlocal, f0, force, fg = goldenlib.inputs(nvel, o0.cv(), o0.wv())
    orc = pyoracle.Oracle(nvel, nlocal)
    m = orc.interior_mask()
    for nrelax, name in ((0, "m10"), (1, "bgk"), (2, "trt")):
        if "collide_%s_f" % name not in g:
            continue
        prm = pyoracle.make_param(force=fg, nrelax=nrelax, **goldenlib.PHYS)
        f = f0.copy()
        rho, u = np.zeros(orc.nsite), np.zeros((orc.nsite, 3))
        orc.collide_mrt(prm, f, force, rho, u)
        assert relerr(f[m], g["collide_%s_f" % name]) < 1e-13
        assert relerr(rho[m], g["collide_%s_rho" % name]) < 1e-14
        assert np.max(np.abs(u[m] - g["collide_%s_u" % name])) < 1e-15
        if name == "m10":
            orc.halo(f, nvel)
            # halo is pure data movement: equal wherever the collided values are
            fh = g["halo_f"]
            assert relerr(f, fh) < 1e-13
            f = fh.copy()
            fp = np.zeros_like(f)
            orc.propagate(f, fp)
            assert np.array_equal(fp[m], g["prop_f"])      # bit-exact


@pytest.mark.parametrize("nvel", [19, 27, 9])
def test_ten_steps(nvel):
    o0 = pyoracle.Oracle(nvel, (2, 2, 1 if nvel == 9 else 2))
    g, nlocal, f0, force, fg = goldenlib.inputs(nvel, o0.cv(), o0.wv())
    orc = pyoracle.Oracle(nvel, nlocal)
    m = orc.interior_mask()
    prm = pyoracle.make_param(eta_shear=0.1, eta_bulk=0.1)
    f, fp = f0.copy(), np.zeros_like(f0)
    n = orc.nsite
    rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    orc.step(prm, 10, f, fp, rho, u, frc)
    assert relerr(f[m], g["step10_f"]) < 1e-12
    assert relerr(rho[m], g["step10_rho"]) < 1e-12
    assert np.max(np.abs(u[m] - g["step10_u"])) < 1e-14


def test_binary():
    nvel = 19
    o0 = pyoracle.Oracle(nvel, (2, 2, 2))
    g, nlocal, f0, force, fg = goldenlib.inputs(nvel, o0.cv(), o0.wv(), ndist=2)
    B = goldenlib.BIN
    orc = pyoracle.Oracle(nvel, nlocal, ndist=2)
    m = orc.interior_mask()
    n = orc.nsite
    prm = pyoracle.make_param(eta_shear=B["eta"], eta_bulk=B["eta"],
                              mobility=B["mobility"], a=B["a"], b=B["b"],
                              kappa=B["kappa"])
    phi, grad, delsq = np.zeros(n), np.zeros((n, 3)), np.zeros(n)
    orc.phi_from_g(f0, phi)
    orc.halo(phi, 1)
    assert np.array_equal(phi, g["bin_phi"])
    orc.grad_7pt(phi, grad, delsq)
    assert np.array_equal(grad[m], g["bin_grad"])
    assert np.array_equal(delsq[m], g["bin_delsq"])
    f, u = f0.copy(), np.zeros((n, 3))
    orc.collide_binary(prm, f, force, u, phi, grad, delsq)
    assert relerr(f[m], g["bin_collide_f"]) < 1e-13
    assert np.max(np.abs(u[m] - g["bin_collide_u"])) < 1e-15
    f, fp = f0.copy(), np.zeros_like(f0)
    rho, frc = np.zeros(n), np.zeros((n, 3))
    orc.step(prm, 10, f, fp, rho, u, frc, phi, grad, delsq)
    assert relerr(f[m], g["bin_step10_f"]) < 1e-12
    # ref phi is the field at the START of step 10; recompute from step-9 state not
    # available here, so compare the order parameter of the final state instead
    assert abs(f[m][:, 1].sum() - g["bin_step10_f"][:, 1].sum()) < 1e-10


@pytest.mark.parametrize("nvel", [19, 27])
def test_gradient_27pt_golden(nvel):
    """orc_grad_27pt against reference-produced vectors (tests/golden/ref_grad27.npz)."""
    g, nlocal, f0 = goldenlib.inputs_grad27(nvel)
    orc = pyoracle.Oracle(nvel, nlocal, ndist=2)
    m = orc.interior_mask()
    n = orc.nsite
    tag = "q%d_" % nvel
    phi = np.zeros(n)
    orc.phi_from_g(f0, phi)
    orc.halo(phi, 1)
    assert np.array_equal(phi, g[tag + "phi"])
    grad, delsq = np.zeros((n, 3)), np.zeros(n)
    orc.grad_27pt(phi, grad, delsq)
    assert np.array_equal(grad[m], g[tag + "grad"])
    assert np.array_equal(delsq[m], g[tag + "delsq"])
    B = goldenlib.BIN
    prm = pyoracle.make_param(eta_shear=B["eta"], eta_bulk=B["eta"], mobility=B["mobility"],
                              a=B["a"], b=B["b"], kappa=B["kappa"])
    f = f0.copy()
    u, force = np.zeros((n, 3)), np.zeros((n, 3))
    orc.collide_binary(prm, f, force, u, phi, grad, delsq)
    assert relerr(f[m], g[tag + "collide_f"]) < 1e-13
    f = f0.copy()
    fp, rho = np.zeros_like(f), np.zeros(n)
    orc.gradient_select(27)
    try:
        orc.step(prm, 10, f, fp, rho, u, force, phi, grad, delsq)
    finally:
        orc.gradient_select(7)
    assert relerr(f[m], g[tag + "step10_f"]) < 1e-12
    assert np.max(np.abs(phi[m] - g[tag + "step10_phi"])) < 1e-13


@pytest.mark.parametrize("tag", ["q19", "q9", "q19b"])
def test_fluctuations_golden(tag):
    """Isothermal fluctuations against reference-produced vectors
    (tests/golden/ref_noise.npz): generator state bit for bit, populations to round-off."""
    g, nvel, ndist, nlocal, f0, force = goldenlib.inputs_noise(tag)
    kt = float(g["kt"])
    orc = pyoracle.Oracle(nvel, nlocal, ndist=ndist)
    m = orc.interior_mask()
    n = orc.nsite
    kw = dict(goldenlib.NOISE_PHYS[ndist])
    if ndist == 2:
        kw.update(a=-0.00625, b=0.00625, kappa=0.004)
    assert np.array_equal(orc.noise_init(0), g[tag + "_state0"])
    if tag == "q19":
        o2 = pyoracle.Oracle(19, (4, 3, 5))
        assert np.array_equal(o2.noise_init(8361235), g["seeded_state0"])
    try:
        for nrelax, name in ((0, "m10"), (1, "bgk")):
            if tag + "_%s_collide_f" % name not in g:
                continue
            prm = pyoracle.make_param(nrelax=nrelax, **kw)
            state = orc.noise_init(0)
            orc.noise_attach(state, kt)
            f = f0.copy()
            rho, u = np.zeros(n), np.zeros((n, 3))
            phi, grad, delsq = np.zeros(n), np.zeros((n, 3)), np.zeros(n)
            if ndist == 2:
                orc.phi_from_g(f, phi)
                orc.halo(phi, 1)
                orc.grad_7pt(phi, grad, delsq)
                orc.collide_binary(prm, f, force, u, phi, grad, delsq)
            else:
                orc.collide_mrt(prm, f, force, rho, u)
            assert np.array_equal(state[m], g[tag + "_%s_collide_state" % name])
            assert relerr(f[m], g[tag + "_%s_collide_f" % name]) < 1e-13
            assert np.max(np.abs(u[m] - g[tag + "_%s_collide_u" % name])) < 1e-15
            state = orc.noise_init(0)
            orc.noise_attach(state, kt)
            f = f0.copy()
            fp, frc = np.zeros_like(f), np.zeros((n, 3))
            orc.step(prm, 10, f, fp, rho, u, frc, phi, grad, delsq)
            assert np.array_equal(state[m], g[tag + "_%s_step10_state" % name])
            assert relerr(f[m], g[tag + "_%s_step10_f" % name]) < 1e-12
    finally:
        orc.noise_attach(None, 0.0)


@pytest.mark.parametrize("tag", sorted(goldenlib.CH_CASES))
def test_cahn_hilliard_golden(tag):
    """The finite-difference order parameter (oracle/lb_oracle_ch.c) against
    reference-produced vectors (tests/golden/ref_ch.npz): force and updated phi bit for
    bit, eight time steps to 1e-12."""
    g = goldenlib.load_ch()
    nlocal, order, method, npt, gmu = goldenlib.CH_CASES[tag]
    P = goldenlib.CH_PHYS
    orc = pyoracle.Oracle(19, nlocal)
    prm = pyoracle.make_param(eta_shear=P["eta"], eta_bulk=P["eta"], mobility=P["mobility"],
                              a=P["a"], b=P["b"], kappa=P["kappa"])
    ch = orc.ch(order, method, gmu)
    m = orc.interior_mask()
    n = orc.nsite
    phi, u = goldenlib.ch_fields(nlocal)
    orc.halo(phi, 1)
    grad, delsq, force = np.zeros((n, 3)), np.zeros(n), np.zeros((n, 3))
    (orc.grad_27pt if npt == 27 else orc.grad_7pt)(phi, grad, delsq)
    orc.phi_force(prm, ch, phi, grad, delsq, force)
    assert np.array_equal(force[m], g[tag + "_force"])
    orc.cahn_hilliard(prm, ch, phi, delsq, u)
    assert np.array_equal(phi[m], g[tag + "_phi1"])

    phi, _ = goldenlib.ch_fields(nlocal)
    f = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    fp = np.zeros_like(f)
    rho, u, force = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    orc.gradient_select(npt)
    orc.ch_attach(ch)
    try:
        orc.step(prm, 8, f, fp, rho, u, force, phi, grad, delsq)
    finally:
        orc.ch_attach(None)
        orc.gradient_select(7)
    assert goldenlib.relerr(f[m], g[tag + "_step8_f"]) < 1e-12
    assert goldenlib.relerr(phi[m], g[tag + "_step8_phi"]) < 1e-13
    assert goldenlib.relerr(u[m], g[tag + "_step8_u"]) < 1e-11
