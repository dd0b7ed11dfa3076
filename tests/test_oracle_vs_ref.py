"""Pins the CPU restatement (oracle/lb_oracle*.c) against the UNMODIFIED reference
compiled in place (oracle/_ref/libref_*.so, built by oracle/Makefile from
/root/reference). Skipped where that build is absent; the committed golden vectors
(tests/golden, test_oracle_golden.py) carry the same pin to such machines.

Tolerances: data movement bit-exact; arithmetic 1e-13 relative to max|f| (the
restatement sums the same terms, the reference's unrolled D3Q19 code in a different
association).
"""

import numpy as np
import pytest

from oracle import pyoracle, pyref
import lbinit

CASES = [(19, (12, 10, 8)), (27, (8, 6, 10)), (9, (16, 12, 1))]


def need_ref(nvel):
    if not pyref.available(nvel):
        pytest.skip("reference build oracle/_ref not present")


def relerr(a, b):
    scale = max(np.max(np.abs(b)), 1e-300)
    return np.max(np.abs(a - b)) / scale


@pytest.mark.parametrize("nvel,nlocal", CASES)
def test_model_constants(nvel, nlocal):
    need_ref(nvel)
    orc = pyoracle.Oracle(nvel, nlocal)
    with pyref.Ref(nvel, nlocal) as ref:
        wv, ma, mi, cv = ref.model()
        assert np.array_equal(cv, orc.cv())
        assert np.array_equal(wv, orc.wv())
        assert np.max(np.abs(ma - orc.ma())) < 1e-15
        assert np.max(np.abs(mi - orc.mi())) < 1e-15
        assert ref.nall == orc.nall and ref.nsite == orc.nsite
        assert ref.index(1, 2, 1) == orc.index(1, 2, 1)
        # host layout is LB_ADDR AOS (src/memory.h:124-130)
        assert ref.addr_f(5, 0, 3) == 5 * nvel + 3


@pytest.mark.parametrize("nvel,nlocal", CASES)
@pytest.mark.parametrize("nrelax", [pyoracle.RELAX_M10, pyoracle.RELAX_BGK,
                                    pyoracle.RELAX_TRT])
def test_collide_mrt(nvel, nlocal, nrelax):
    need_ref(nvel)
    if nrelax == pyoracle.RELAX_TRT and nvel != 19:
        pytest.skip("reference TRT is D3Q15/D3Q19 only")
    orc = pyoracle.Oracle(nvel, nlocal)
    ndim = orc.ndim
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndim=ndim)
    force = lbinit.make_force(nlocal, ndim=ndim)
    fg = (1e-4, -2e-4, 3e-4 if ndim == 3 else 0.0)
    prm = pyoracle.make_param(eta_shear=0.05, eta_bulk=0.12, force=fg,
                              nrelax=nrelax)
    with pyref.Ref(nvel, nlocal) as ref:
        ref.physics(eta_shear=0.05, eta_bulk=0.12, force=fg)
        ref.relaxation(nrelax)
        assert np.allclose(ref.rtau()[orc.model.nhydro:nvel],
                           orc.rtau(prm)[orc.model.nhydro:nvel], rtol=0, atol=1e-15)
        ref.f()[...] = f0
        ref.force()[...] = force
        ref.collide()
        f = f0.copy()
        rho = np.zeros(orc.nsite)
        u = np.zeros((orc.nsite, 3))
        orc.collide_mrt(prm, f, force, rho, u)
        m = orc.interior_mask()
        assert relerr(f[m], ref.f()[m]) < 1e-13
        assert relerr(rho[m], ref.rho()[m]) < 1e-14
        assert np.max(np.abs(u[m] - ref.u()[m])) < 1e-15


@pytest.mark.parametrize("nvel,nlocal", CASES)
@pytest.mark.parametrize("ndist", [1, 2])
def test_halo_and_propagation_bit_exact(nvel, nlocal, ndist):
    need_ref(nvel)
    if ndist == 2 and nvel == 9:
        pytest.skip("two distributions are 3-d only in the hot path")
    orc = pyoracle.Oracle(nvel, nlocal, ndist=ndist)
    rng = np.random.default_rng(5)
    f0 = rng.standard_normal((orc.nsite, ndist, nvel))
    with pyref.Ref(nvel, nlocal, ndist=ndist) as ref:
        ref.f()[...] = f0
        ref.halo()
        f = f0.copy()
        orc.halo(f, ndist * nvel)
        assert np.array_equal(f, ref.f())
        ref.propagate()
        fp = np.zeros_like(f)
        orc.propagate(f, fp)
        m = orc.interior_mask()
        assert np.array_equal(fp[m], ref.f()[m])


def test_halo_openmp_schemes_match_target():
    nvel, nlocal = 19, (8, 6, 4)
    need_ref(nvel)
    orc = pyoracle.Oracle(nvel, nlocal)
    rng = np.random.default_rng(6)
    f0 = rng.standard_normal((orc.nsite, 1, nvel))
    f = f0.copy()
    orc.halo(f, nvel)
    with pyref.Ref(nvel, nlocal, halo=pyref.HALO_OPENMP_FULL) as ref:
        ref.f()[...] = f0
        ref.halo()
        assert np.array_equal(f, ref.f())


def test_binary_stages():
    nvel, nlocal = 19, (10, 8, 6)
    need_ref(nvel)
    orc = pyoracle.Oracle(nvel, nlocal, ndist=2)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=2)
    force = lbinit.make_force(nlocal)
    A, B, K, M, eta = -0.00625, 0.00625, 0.004, 3.75, 0.00625
    prm = pyoracle.make_param(eta_shear=eta, eta_bulk=eta, mobility=M, a=A, b=B,
                              kappa=K)
    with pyref.Ref(nvel, nlocal, ndist=2) as ref:
        ref.physics(eta_shear=eta, eta_bulk=eta, mobility=M)
        ref.fe_symm(A, B, K)
        ref.f()[...] = f0
        ref.force()[...] = force
        ref.phi_from_lb()
        phi = np.zeros(orc.nsite)
        orc.phi_from_g(f0, phi)
        m = orc.interior_mask()
        assert np.array_equal(phi[m], ref.phi()[m])
        ref.phi_halo()
        orc.halo(phi, 1)
        assert np.array_equal(phi, ref.phi())
        ref.grad_compute()
        grad = np.zeros((orc.nsite, 3))
        delsq = np.zeros(orc.nsite)
        orc.grad_7pt(phi, grad, delsq)
        assert np.array_equal(grad[m], ref.grad()[m])
        assert np.array_equal(delsq[m], ref.delsq()[m])
        ref.collide()
        f = f0.copy()
        u = np.zeros((orc.nsite, 3))
        orc.collide_binary(prm, f, force, u, phi, grad, delsq)
        assert relerr(f[m][:, 0], ref.f()[m][:, 0]) < 1e-13
        assert relerr(f[m][:, 1], ref.f()[m][:, 1]) < 1e-13
        assert np.max(np.abs(u[m] - ref.u()[m])) < 1e-15


@pytest.mark.parametrize("nvel,nlocal,ndist", [(19, (8, 8, 8), 1), (27, (6, 6, 6), 1),
                                               (9, (12, 12, 1), 1), (19, (8, 8, 8), 2)])
def test_time_loop(nvel, nlocal, ndist):
    need_ref(nvel)
    nsteps = 5
    orc = pyoracle.Oracle(nvel, nlocal, ndist=ndist)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=ndist, ndim=orc.ndim)
    kw = dict(eta_shear=0.1, eta_bulk=0.1)
    if ndist == 2:
        kw = dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75)
    prm = pyoracle.make_param(a=-0.00625, b=0.00625, kappa=0.004, **kw)
    with pyref.Ref(nvel, nlocal, ndist=ndist) as ref:
        ref.physics(**kw)
        if ndist == 2:
            ref.fe_symm(-0.00625, 0.00625, 0.004)
        ref.f()[...] = f0
        ref.step(nsteps)
        f = f0.copy()
        fp = np.zeros_like(f)
        n = orc.nsite
        rho, u, force = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        phi, grad, delsq = np.zeros(n), np.zeros((n, 3)), np.zeros(n)
        orc.step(prm, nsteps, f, fp, rho, u, force, phi, grad, delsq)
        m = orc.interior_mask()
        assert relerr(f[m], ref.f()[m]) < 1e-12
        assert abs(f[m][:, 0].sum() - f0[m][:, 0].sum()) < 1e-9


@pytest.mark.parametrize("nvel,nlocal", [(19, (10, 8, 6)), (27, (5, 7, 6))])
def test_gradient_27pt(nvel, nlocal):
    """grad_3d_27pt_fluid_d2 selected through field_grad_set: bit-exact, alone and
    inside the time loop."""
    need_ref(nvel)
    orc = pyoracle.Oracle(nvel, nlocal, ndist=2)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=2)
    A, B, K, M, eta = -0.00625, 0.00625, 0.004, 3.75, 0.00625
    prm = pyoracle.make_param(eta_shear=eta, eta_bulk=eta, mobility=M, a=A, b=B, kappa=K)
    m = orc.interior_mask()
    rng = np.random.default_rng(27)
    with pyref.Ref(nvel, nlocal, ndist=2) as ref:
        ref.physics(eta_shear=eta, eta_bulk=eta, mobility=M)
        ref.fe_symm(A, B, K)
        ref.grad_select(27)
        phi = rng.standard_normal(orc.nsite)         # every site, halo too
        ref.phi()[...] = phi
        ref.grad_compute()
        grad, delsq = np.zeros((orc.nsite, 3)), np.zeros(orc.nsite)
        orc.grad_27pt(phi, grad, delsq)
        assert np.array_equal(grad[m], ref.grad()[m])
        assert np.array_equal(delsq[m], ref.delsq()[m])
        # and it is not the 7-point result
        g7, d7 = np.zeros_like(grad), np.zeros_like(delsq)
        orc.grad_7pt(phi, g7, d7)
        assert not np.allclose(g7[m], grad[m])
        ref.f()[...] = f0
        ref.step(5)
        f = f0.copy()
        fp = np.zeros_like(f)
        n = orc.nsite
        rho, u, force = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        phi2 = np.zeros(n)
        orc.gradient_select(27)
        try:
            orc.step(prm, 5, f, fp, rho, u, force, phi2, grad, delsq)
        finally:
            orc.gradient_select(7)
        assert relerr(f[m], ref.f()[m]) < 1e-12


KT = 2.13333e-05       # temperature of tests/regression/d2q9/serial-flct-t01.inp


@pytest.mark.parametrize("nvel,nlocal", [(19, (6, 5, 4)), (9, (8, 7, 1))])
@pytest.mark.parametrize("seed", [0, 8361235])
def test_noise_generator(nvel, nlocal, seed):
    """noise_init / noise_uniform / noise_reap_n (src/noise.c): integer state and the
    discrete variates, bit for bit, every site of the halo-inclusive box."""
    need_ref(nvel)
    orc = pyoracle.Oracle(nvel, nlocal)
    state = orc.noise_init(seed)
    with pyref.Ref(nvel, nlocal) as ref:
        ref.noise(1, KT, reseed=seed)
        assert np.array_equal(state, ref.noise_state())
        for index in (0, 7, orc.nsite - 1):
            for nmax in (6, 10, 3):
                a = orc.noise_reap_n(state, index, nmax)
                b = ref.noise_reap_n(index, nmax)
                assert np.array_equal(a, b)
        assert np.array_equal(state, ref.noise_state())
    # levels have zero mean and unit variance over the eight equally likely indices
    lv = np.array([-np.sqrt(2 + np.sqrt(2)), -np.sqrt(2 - np.sqrt(2)), 0, 0, 0, 0,
                   np.sqrt(2 - np.sqrt(2)), np.sqrt(2 + np.sqrt(2))])
    assert abs(lv.mean()) < 1e-15 and abs((lv**2).mean() - 1.0) < 1e-15


@pytest.mark.parametrize("nvel,nlocal,ndist,nrelax,ghost_off", [
    (19, (6, 5, 4), 1, 0, 0), (19, (6, 5, 4), 1, 1, 0), (19, (6, 5, 4), 1, 2, 0),
    (9, (8, 7, 1), 1, 0, 0), (9, (8, 7, 1), 1, 1, 0), (19, (6, 5, 4), 2, 0, 0),
    (19, (6, 5, 4), 1, 0, 1), (19, (6, 5, 4), 1, 1, 1), (19, (6, 5, 4), 2, 0, 1),
    (19, (6, 5, 4), 2, 1, 1), (9, (8, 7, 1), 1, 0, 1)])
def test_fluctuating_collision(nvel, nlocal, ndist, nrelax, ghost_off):
    """lb_collide with isothermal fluctuations on: one collision, then the time loop;
    generator state stays bit-identical, populations to round-off."""
    need_ref(nvel)
    orc = pyoracle.Oracle(nvel, nlocal, ndist=ndist)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=ndist, ndim=orc.ndim)
    force = lbinit.make_force(nlocal, ndim=orc.ndim)
    kw = dict(eta_shear=0.1, eta_bulk=0.07)
    if ndist == 2:
        kw = dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75)
    prm = pyoracle.make_param(a=-0.00625, b=0.00625, kappa=0.004, nrelax=nrelax,
                              ghost_off=ghost_off, **kw)
    m = orc.interior_mask()
    n = orc.nsite
    state = orc.noise_init(0)
    orc.noise_attach(state, KT)
    try:
        with pyref.Ref(nvel, nlocal, ndist=ndist) as ref:
            ref.physics(**kw)
            ref.relaxation(nrelax)
            ref.ghost_modes(not ghost_off)
            if ndist == 2:
                ref.fe_symm(-0.00625, 0.00625, 0.004)
            ref.noise(1, KT)
            ref.f()[...] = f0
            ref.force()[...] = force
            f = f0.copy()
            rho, u = np.zeros(n), np.zeros((n, 3))
            phi, grad, delsq = np.zeros(n), np.zeros((n, 3)), np.zeros(n)
            if ndist == 2:
                ref.phi_from_lb()
                ref.phi_halo()
                ref.grad_compute()
                orc.phi_from_g(f, phi)
                orc.halo(phi, 1)
                orc.grad_7pt(phi, grad, delsq)
                ref.collide()
                orc.collide_binary(prm, f, force, u, phi, grad, delsq)
            else:
                ref.collide()
                orc.collide_mrt(prm, f, force, rho, u)
            assert np.array_equal(state, ref.noise_state())
            assert relerr(f[m], ref.f()[m]) < 1e-13
            # the noise is there: compare with the deterministic collision
            orc.noise_attach(None, 0.0)
            fd = f0.copy()
            if ndist == 2:
                orc.collide_binary(prm, fd, force, u, phi, grad, delsq)
            else:
                orc.collide_mrt(prm, fd, force, rho, u)
            assert relerr(fd[m][:, 0], f[m][:, 0]) > 1e-4
            # mass and momentum are untouched by the fluctuations
            assert np.max(np.abs(fd[m][:, 0].sum(-1) - f[m][:, 0].sum(-1))) < 1e-14
            cv = orc.cv().astype(float)
            assert np.max(np.abs((fd[m][:, 0] - f[m][:, 0]) @ cv)) < 1e-15
            # time loop
            orc.noise_attach(state, KT)
            ref.f()[...] = f0
            ref.step(8)
            f = f0.copy()
            fp, frc = np.zeros_like(f), np.zeros((n, 3))
            orc.step(prm, 8, f, fp, rho, u, frc, phi, grad, delsq)
            assert np.array_equal(state, ref.noise_state())
            assert relerr(f[m], ref.f()[m]) < 1e-12
    finally:
        orc.noise_attach(None, 0.0)


@pytest.mark.parametrize("nvel,nlocal,walls,ubot,utop", [
    (19, (6, 5, 4), (1, 1, 1), (0, 0, 0), (0, 0, 0)),
    (19, (4, 7, 6), (0, 1, 1), (0, 0, 0), (0, 0, 0)),
    (19, (5, 4, 6), (0, 0, 1), (-0.01, 0, 0), (0.02, 0, 0)),
    (27, (4, 5, 4), (1, 0, 1), (0, 0, 0), (0, 0, 0)),
    (9, (8, 7, 1), (0, 1, 0), (0.015, 0, 0), (-0.005, 0, 0)),
    (9, (8, 7, 1), (1, 1, 0), (0, 0, 0), (0, 0, 0))])
def test_wall_bounce_back(nvel, nlocal, walls, ubot, utop):
    """wall_commit + wall_bbl (src/wall.c): link count, the populations written into
    the wall layer, the momentum account, and the time loop with walls (collide, halo,
    bbl, propagate) with a body force driving a flow."""
    need_ref(nvel)
    periodic = tuple(0 if w else 1 for w in walls)
    orc = pyoracle.Oracle(nvel, nlocal)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndim=orc.ndim)
    fg = (2e-4, 0.0, 0.0) if not walls[0] else (0.0, 0.0, 0.0)
    kw = dict(eta_shear=0.1666, eta_bulk=0.1)
    prm = pyoracle.make_param(force=fg, **kw)
    m = orc.interior_mask()
    n = orc.nsite
    with pyref.Ref(nvel, nlocal, periodic=periodic) as ref:
        ref.physics(force=fg, **kw)
        nlink = ref.wall(walls, ubot, utop)
        w = orc.wall(walls, ubot, utop)
        assert orc.wall_nlink(w) == nlink and nlink > 0
        # one bounce-back on a post-collision state: the reversed populations of the wall
        # layer are exactly what the reference writes there (lb->param->rho0, which
        # scales the moving-wall term, is only set by a collision: src/collision.c:1200)
        ref.f()[...] = f0
        ref.collide()
        ref.f()[...] = f0
        ref.wall_bbl()
        f = f0.copy()
        orc.wall_bbl(prm, w, f)
        assert np.array_equal(f, ref.f())
        assert np.allclose(np.array(w.fnet[:]), ref.wall_momentum(), rtol=1e-12, atol=1e-15)
        # time loop
        ref.f()[...] = f0
        ref.step(12)
        f = f0.copy()
        fp = np.zeros_like(f)
        rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        w = orc.wall(walls, ubot, utop)
        orc.wall_attach(w)
        try:
            orc.step(prm, 12, f, fp, rho, u, frc)
        finally:
            orc.wall_attach(None)
        assert relerr(f[m], ref.f()[m]) < 1e-12
        assert np.max(np.abs(u[m] - ref.u()[m])) < 1e-14
        # wall_momentum() above moved the account of the single bbl to the host total;
        # the twelve steps add to it
        fn = ref.wall_momentum()
        w1 = orc.wall(walls, ubot, utop)
        orc.wall_bbl(prm, w1, f0.copy())
        assert np.allclose(np.array(w.fnet[:]) + np.array(w1.fnet[:]), fn, rtol=1e-10, atol=1e-14)
        # mass of the fluid is conserved by the bounce-back
        assert abs(f[m].sum() - f0[m].sum()) < 1e-10


@pytest.mark.parametrize("nvel,nlocal,walls,sbot,stop,force_active", [
    (19, (4, 7, 6), (0, 1, 1), (0, 0, 0), (0, 0, 0), 1),            # serial-rect-ct2: slip code, s = 0
    (19, (4, 7, 6), (0, 1, 1), (0, 0, 0), (0, 1.0, 1.0), 0),        # serial-rect-ct3
    (19, (6, 5, 4), (1, 1, 0), (0, 0, 0), (1.0, 1.0, 0), 0),        # serial-rect-ct4
    (19, (6, 5, 4), (1, 1, 1), (0.3, 0.1, 0.7), (0.2, 0.9, 0.5), 0),
    (27, (4, 5, 4), (1, 0, 1), (0.25, 0, 0.5), (1.0, 0, 0.75), 0),
    (9, (8, 7, 1), (1, 1, 0), (0.5, 0.2, 0), (0.4, 1.0, 0), 0)])
def test_wall_slip(nvel, nlocal, walls, sbot, stop, force_active):
    """wall_bbl_slip_kernel with the link geometry of wall_init_boundaries_slip: faces,
    edges (mean of the two walls) and corners (no slip), alone and in the time loop."""
    need_ref(nvel)
    periodic = tuple(0 if w else 1 for w in walls)
    orc = pyoracle.Oracle(nvel, nlocal)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndim=orc.ndim)
    fg = (2e-4, 0.0, 0.0) if not walls[0] else (0.0, 0.0, 0.0)
    kw = dict(eta_shear=0.1666, eta_bulk=0.1)
    prm = pyoracle.make_param(force=fg, **kw)
    m = orc.interior_mask()
    n = orc.nsite
    with pyref.Ref(nvel, nlocal, periodic=periodic) as ref:
        ref.physics(force=fg, **kw)
        nlink = ref.wall_slip(walls, sbot, stop, force_active)
        w = orc.wall(walls, sbot=sbot, stop=stop, slip_active=1)
        assert orc.wall_nlink(w) == nlink
        ref.f()[...] = f0
        ref.wall_bbl()
        f = f0.copy()
        orc.wall_bbl(prm, w, f)
        assert relerr(f, ref.f()) < 1e-15        # (1-s) fi + s fk: same expression
        assert np.allclose(np.array(w.fnet[:]), ref.wall_momentum(), rtol=1e-12, atol=1e-15)
        ref.f()[...] = f0
        ref.step(12)
        f = f0.copy()
        fp = np.zeros_like(f)
        rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        w = orc.wall(walls, sbot=sbot, stop=stop, slip_active=1)
        orc.wall_attach(w)
        try:
            orc.step(prm, 12, f, fp, rho, u, frc)
        finally:
            orc.wall_attach(None)
        assert relerr(f[m], ref.f()[m]) < 1e-12
        assert abs(f[m].sum() - f0[m].sum()) < 1e-10


@pytest.mark.parametrize("nvel,nlocal,ndist,le", [
    (19, (16, 6, 4), 1, dict(nplanes=2, uy=0.05)),
    (19, (12, 7, 5), 1, dict(nplanes=1, uy=-0.03)),
    (19, (12, 8, 4), 2, dict(nplanes=3, uy=0.02)),
    (19, (16, 6, 4), 1, dict(nplanes=2, uy=0.04, oscillatory=1, period=40, nt0=0)),
    (27, (8, 6, 5), 1, dict(nplanes=2, uy=0.05)),
    (9, (16, 12, 1), 1, dict(nplanes=4, uy=0.01))])
def test_lees_edwards(nvel, nlocal, ndist, le):
    """lb_le_apply_boundary_conditions (src/model_le.c:57): reprojection of the plane-
    crossing populations and their interpolated displacement, at several times (integer
    and fractional displacements, wrap-around), then inside the time loop."""
    need_ref(nvel)
    orc = pyoracle.Oracle(nvel, nlocal, ndist=ndist)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=ndist, ndim=orc.ndim)
    m = orc.interior_mask()
    n = orc.nsite
    kw = dict(eta_shear=0.1, eta_bulk=0.1)
    if ndist == 2:
        kw = dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75)
    prm = pyoracle.make_param(a=-0.00625, b=0.00625, kappa=0.004, **kw)
    ole = orc.le(**le)
    with pyref.Ref(nvel, nlocal, ndist=ndist, le=le) as ref:
        ref.physics(**kw)
        if ndist == 2:
            ref.fe_symm(-0.00625, 0.00625, 0.004)
        for k in range(le["nplanes"]):
            assert orc.le_plane_location(ole, k) == ref.le_plane_location(k)
        for t in (1, 7, 20, 133, 1000):
            ref.f()[...] = f0
            ref.le_apply(t)
            f = f0.copy()
            orc.le_apply(ole, t, f)
            assert relerr(f, ref.f()) < 1e-15, t
        assert relerr(f, f0) > 1e-4           # it does something
        ref.f()[...] = f0
        ref.time_set(5)
        ref.step(9)
        f = f0.copy()
        fp = np.zeros_like(f)
        rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        phi, grad, delsq = np.zeros(n), np.zeros((n, 3)), np.zeros(n)
        ole.t_current, ole.t_start = 5, 5
        orc.le_attach(ole)
        try:
            orc.step(prm, 9, f, fp, rho, u, frc, phi, grad, delsq)
        finally:
            orc.le_attach(None)
        # ndist 2: the gradient of phi across the planes feeds the binary collision
        assert relerr(f[m], ref.f()[m]) < 1e-12
        assert abs(f[m][:, 0].sum() - f0[m][:, 0].sum()) < 1e-10


@pytest.mark.parametrize("nlocal,le", [
    ((16, 6, 4), dict(nplanes=2, uy=0.05)),
    ((12, 7, 5), dict(nplanes=1, uy=-0.03)),
    ((12, 8, 4), dict(nplanes=3, uy=0.02)),
    ((8, 9, 6), dict(nplanes=4, uy=0.31)),
    ((16, 6, 4), dict(nplanes=2, uy=0.04, oscillatory=1, period=40, nt0=0))])
@pytest.mark.parametrize("npt", [7, 27])
def test_lees_edwards_order_parameter(nlocal, le, npt):
    """field_grad_compute with planes (src/field_grad.c:305-325): the buffer planes of
    field_leesedwards (src/field.c:500-600; four-point interpolation along y at the
    displacement the planes have reached) and the 7- and 27-point stencils that read
    them (lees_edw_ic_to_buff in src/gradient_3d_7pt_fluid.c, _27pt_fluid.c), bit for
    bit; then the binary time loop across the planes."""
    nvel = 19
    need_ref(nvel)
    orc = pyoracle.Oracle(nvel, nlocal, ndist=2)
    m = orc.interior_mask()
    n = orc.nsite
    nall = orc.nall
    rng = np.random.default_rng(12)
    phi0 = np.zeros(nall)
    phi0[1:-1, 1:-1, 1:-1] = 0.4*(rng.random(nlocal) - 0.5)
    phi0 = phi0.reshape(-1)
    orc.halo(phi0, 1)
    ole = orc.le(**le)
    kw = dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75)
    prm = pyoracle.make_param(a=-0.00625, b=0.00625, kappa=0.004, **kw)
    with pyref.Ref(nvel, nlocal, ndist=2, le=le) as ref:
        ref.physics(**kw)
        ref.fe_symm(-0.00625, 0.00625, 0.004)
        ref.grad_select(npt)
        for t in (0, 1, 7, 20, 133, 1000):
            ref.time_set(t)
            ref.phi_all()[...] = 0.0
            ref.phi()[...] = phi0
            ref.grad_compute()
            t = ref.time()                                    # t_start + t_current - 1
            buf = orc.le_field_buffers(ole, t, phi0)
            rbuf = ref.phi_all()[n:].reshape(buf.shape)
            assert np.array_equal(buf, rbuf), t
            assert np.array_equal(ref.phi(), phi0)           # the real planes are untouched
            grad, delsq = np.zeros((n, 3)), np.zeros(n)
            orc.le_field_grad(ole, t, npt, phi0, grad, delsq)
            assert np.array_equal(grad[m], ref.grad()[m]), t
            assert np.array_equal(delsq[m], ref.delsq()[m]), t
        # it differs from the stencil that knows nothing about the planes, next to them only
        g0, d0 = np.zeros((n, 3)), np.zeros(n)
        (orc.grad_27pt if npt == 27 else orc.grad_7pt)(phi0, g0, d0)
        x = np.arange(nall[0])[:, None, None]*np.ones(nall, dtype=int)
        near = np.zeros(nall, dtype=bool)
        for k in range(le["nplanes"]):
            loc = orc.le_plane_location(ole, k)
            near |= (x == loc) | (x == loc + 1)
        near = near.reshape(-1)
        assert np.array_equal(grad[m & ~near], g0[m & ~near])
        assert not np.array_equal(grad[m & near], g0[m & near])
        # the time loop
        f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=2)
        ref.f()[...] = f0
        ref.time_set(5)
        ref.step(9)
        f, fp = f0.copy(), np.zeros_like(f0)
        rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        phi, grad, delsq = np.zeros(n), np.zeros((n, 3)), np.zeros(n)
        ole.t_current, ole.t_start = 5, 5
        orc.le_attach(ole)
        orc.gradient_select(npt)
        try:
            orc.step(prm, 9, f, fp, rho, u, frc, phi, grad, delsq)
        finally:
            orc.gradient_select(7)
            orc.le_attach(None)
        assert relerr(f[m], ref.f()[m]) < 1e-12
        assert np.max(np.abs(grad[m] - ref.grad()[m])) < 1e-12


@pytest.mark.parametrize("nvel,nlocal,walls,u0", [
    (19, (12, 6, 6), (0, 1, 1), (0.0, 0.0, 0.0)), (19, (6, 6, 12), (1, 1, 0), (0.0, 0.0, 0.003)),
    (27, (8, 6, 5), (0, 1, 1), (0.004, 0.0, 0.0)), (19, (5, 9, 7), (1, 0, 1), (0.0, 0.002, 0.0))])
def test_open_boundaries(nvel, nlocal, walls, u0):
    """lb_bc_inflow_rhou / lb_bc_outflow_rhou update + impose (src/lb_bc_inflow_rhou.c:235-281,
    478-531 and the outflow twin): the imposed populations bit for bit, then the time loop
    with walls across the flow."""
    need_ref(nvel)
    flow = walls.index(0)
    orc = pyoracle.Oracle(nvel, nlocal)
    n = orc.nsite
    m = orc.interior_mask()
    kw = dict(eta_shear=0.1, eta_bulk=0.1)
    prm = pyoracle.make_param(**kw)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    rng = np.random.default_rng(9)
    rho0 = 1.0 + 0.1*rng.random(n)
    u_in = 0.05*(rng.random((n, 3)) - 0.5)
    with pyref.Ref(nvel, nlocal, periodic=(0, 0, 0)) as ref:
        ref.physics(**kw)
        ref.wall(walls)
        ref.open_bc(flow, u0, 1.01)
        o = orc.open_bc(flow, u0, 1.01)
        ref.f()[...] = f0
        ref.rho()[...] = rho0
        ref.u()[...] = u_in
        ref.open_apply()
        f, rho, u = f0.copy(), rho0.copy(), u_in.copy()
        orc.open_apply(o, rho, u, f)
        assert np.array_equal(f, ref.f())
        assert np.array_equal(rho, ref.rho()) and np.array_equal(u, ref.u())
        assert not np.array_equal(f, f0)
        ref.f()[...] = f0
        ref.step(10)
        f, fp = f0.copy(), np.zeros_like(f0)
        rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        w = orc.wall(walls)
        orc.wall_attach(w)
        orc.open_attach(o)
        try:
            orc.step(prm, 10, f, fp, rho, u, frc)
        finally:
            orc.wall_attach(None)
            orc.open_attach(None)
        assert relerr(f[m], ref.f()[m]) < 1e-12
        assert np.allclose(np.array(w.fnet[:]), ref.wall_momentum(), rtol=1e-10, atol=1e-14)


@pytest.mark.parametrize("nlocal,walls", [((8, 6, 10), (1, 1, 1)), ((6, 7, 5), (0, 1, 1)),
                                          ((9, 5, 6), (1, 0, 0)), ((5, 6, 8), (0, 0, 1))])
def test_gradient_next_to_walls(nlocal, walls):
    """grad_3d_27pt_solid_d2 (src/gradient_3d_27pt_solid.c:131-325) with neutral wetting
    next to flat walls, bit for bit, and the binary time loop between them."""
    nvel = 19
    need_ref(nvel)
    periodic = tuple(0 if w else 1 for w in walls)
    orc = pyoracle.Oracle(nvel, nlocal, ndist=2)
    m = orc.interior_mask()
    n = orc.nsite
    rng = np.random.default_rng(21)
    phi0 = np.zeros(orc.nall)
    phi0[1:-1, 1:-1, 1:-1] = 0.4*(rng.random(nlocal) - 0.5)
    phi0 = phi0.reshape(-1)
    orc.halo(phi0, 1)
    kw = dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75)
    prm = pyoracle.make_param(a=-0.00625, b=0.00625, kappa=0.004, **kw)
    with pyref.Ref(nvel, nlocal, ndist=2, periodic=periodic) as ref:
        ref.physics(**kw)
        ref.fe_symm(-0.00625, 0.00625, 0.004)
        ref.wall(walls)
        ref.grad_select(28)
        ref.phi()[...] = phi0
        ref.grad_compute()
        w = orc.wall(walls)
        grad, delsq = np.zeros((n, 3)), np.zeros(n)
        orc.grad_27pt_solid(w, phi0, grad, delsq)
        assert np.array_equal(grad[m], ref.grad()[m])
        assert np.array_equal(delsq[m], ref.delsq()[m])
        f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=2)
        ref.f()[...] = f0
        ref.step(8)
        f, fp = f0.copy(), np.zeros_like(f0)
        rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        phi = np.zeros(n)
        orc.gradient_select(28)
        orc.wall_attach(w)
        try:
            orc.step(prm, 8, f, fp, rho, u, frc, phi, grad, delsq)
        finally:
            orc.gradient_select(7)
            orc.wall_attach(None)
        assert relerr(f[m], ref.f()[m]) < 1e-12
        assert np.allclose(np.array(w.fnet[:]), ref.wall_momentum(), rtol=1e-10, atol=1e-14)


# ---------------------------------------------------------------------------
# "free_energy symmetric": the order parameter as a finite-difference field
# (phi_force_calculation + phi_cahn_hilliard), reference built with nhalo = 2
# ---------------------------------------------------------------------------

def fd_fields(nlocal):
    """phi, u on the halo-1 box (interior values; halos are filled by the codes)."""
    ids = lbinit.global_ids(nlocal).reshape(-1)
    phi = 0.4 * lbinit.uniform(ids, 50)
    u = np.zeros((ids.size, 3))
    for a in range(3):
        u[:, a] = 0.02 * lbinit.uniform(ids, 2 + a)
    return phi, u


FD_CASES = [
    (1, pyoracle.FORCE_STRESS_DIVERGENCE, 27, (0.0, 0.0, 0.0)),
    (2, pyoracle.FORCE_STRESS_DIVERGENCE, 7, (0.0, 0.0, 0.0)),
    (3, pyoracle.FORCE_STRESS_DIVERGENCE, 27, (0.0, 0.0, 0.0)),
    (1, pyoracle.FORCE_PHI_GRADMU, 27, (1e-5, 2e-5, 3e-5)),
    (3, pyoracle.FORCE_PHI_GRADMU, 7, (-2e-5, 0.0, 1e-5)),
    (2, pyoracle.FORCE_NONE, 7, (0.0, 0.0, 0.0)),
    (4, pyoracle.FORCE_STRESS_DIVERGENCE, 7, (0.0, 0.0, 0.0)),
]


@pytest.mark.parametrize("order,method,npt,gmu", FD_CASES)
def test_cahn_hilliard_stages(order, method, npt, gmu):
    ch_stages((10, 8, 6), order, method, npt, gmu)


@pytest.mark.parametrize("order,method,npt,gmu", [FD_CASES[0], FD_CASES[2], FD_CASES[4]])
def test_cahn_hilliard_one_site_thick(order, method, npt, gmu):
    """n_z = 1 with a three-dimensional model and stencil (as serial-le2d-fd1): the z
    fluxes stay out of the update (wz = 0, src/phi_cahn_hilliard.c:999)."""
    ch_stages((12, 9, 1), order, method, npt, gmu)


def ch_stages(nlocal, order, method, npt, gmu):
    nvel = 19
    need_ref(nvel)
    A, B, K, M, eta = -0.00625, 0.00625, 0.004, 1.25, 0.00625
    orc = pyoracle.Oracle(nvel, nlocal)
    prm = pyoracle.make_param(eta_shear=eta, eta_bulk=eta, mobility=M, a=A, b=B, kappa=K)
    ch = orc.ch(order, method, gmu)
    phi, u = fd_fields(nlocal)
    n = orc.nsite
    with pyref.Ref(nvel, nlocal, fd=dict(order=order, method=method)) as ref:
        assert ref.nhalo == 2
        ref.physics(eta_shear=eta, eta_bulk=eta, mobility=M)
        ref.fe_symm(A, B, K)
        ref.grad_mu(gmu)
        ref.grad_select(npt)
        ref.inner(ref.phi())[...] = phi.reshape(orc.nall)
        ref.inner(ref.u())[...] = u.reshape(orc.nall + (3,))
        m = orc.interior_mask()

        # field_halo + field_grad_compute (region nextra = 1 in the reference)
        ref.phi_halo()
        ref.grad_compute()
        orc.halo(phi, 1)
        grad, delsq = np.zeros((n, 3)), np.zeros(n)
        (orc.grad_27pt if npt == 27 else orc.grad_7pt)(phi, grad, delsq)
        assert np.array_equal(phi, ref.inner(ref.phi()).reshape(-1))

        # phi_force_calculation
        ref.hydro_f_zero()
        ref.phi_force()
        force = np.zeros((n, 3))
        orc.phi_force(prm, ch, phi, grad, delsq, force)
        if method != pyoracle.FORCE_NONE:
            # the gradient one site into the halo is the gradient of the periodic image
            assert np.array_equal(grad, ref.inner(ref.grad()).reshape(-1, 3))
            assert np.array_equal(delsq, ref.inner(ref.delsq()).reshape(-1))
            assert np.max(np.abs(force[m])) > 0.0
        fr = ref.inner(ref.force()).reshape(-1, 3)
        assert np.array_equal(force[m], fr[m])

        # phi_cahn_hilliard
        ref.cahn_hilliard()
        phi0 = phi.copy()
        orc.cahn_hilliard(prm, ch, phi, delsq, u)
        pr = ref.inner(ref.phi()).reshape(-1)
        assert np.max(np.abs(phi[m] - phi0[m])) > 1e-6
        assert np.array_equal(phi[m], pr[m])
        # conservation to round-off
        assert abs(phi[m].sum() - phi0[m].sum()) < 1e-13 * m.sum()


FD_LE_CASES = [
    ((16, 8, 6), dict(nplanes=2, uy=0.05), 1, pyoracle.FORCE_STRESS_DIVERGENCE, 7, (0.0, 0.0, 0.0)),
    ((16, 8, 6), dict(nplanes=2, uy=0.05), 2, pyoracle.FORCE_STRESS_DIVERGENCE, 27, (0.0, 0.0, 0.0)),
    ((12, 7, 5), dict(nplanes=1, uy=-0.03), 3, pyoracle.FORCE_STRESS_DIVERGENCE, 27, (0.0, 0.0, 0.0)),
    ((15, 9, 6), dict(nplanes=3, uy=0.31), 3, pyoracle.FORCE_PHI_GRADMU, 7, (1e-5, 2e-5, 3e-5)),
    ((16, 12, 1), dict(nplanes=2, uy=0.02), 1, pyoracle.FORCE_STRESS_DIVERGENCE, 27, (0.0, 0.0, 0.0)),
    ((12, 8, 6), dict(nplanes=2, uy=0.07), 4, pyoracle.FORCE_STRESS_DIVERGENCE, 7, (0.0, 0.0, 0.0)),
    ((16, 8, 6), dict(nplanes=2, uy=0.04, oscillatory=1, period=40, nt0=0), 2,
     pyoracle.FORCE_STRESS_DIVERGENCE, 7, (0.0, 0.0, 0.0)),
]


@pytest.mark.parametrize("nlocal,le,order,method,npt,gmu", FD_LE_CASES)
def test_cahn_hilliard_lees_edwards(nlocal, le, order, method, npt, gmu):
    """The finite-difference order parameter across Lees-Edwards planes: the gradient in
    the buffer region, the flux form of the force with its plane correction
    (src/phi_force.c:289-340), the velocity buffers, plane-aware advection and diffusive
    fluxes, phi_ch_le_fix_fluxes and the update, stage by stage at several times and in
    the time loop."""
    nvel = 19
    need_ref(nvel)
    A, B, K, M, eta = -0.00625, 0.00625, 0.004, 1.25, 0.00625
    orc = pyoracle.Oracle(nvel, nlocal)
    prm = pyoracle.make_param(eta_shear=eta, eta_bulk=eta, mobility=M, a=A, b=B, kappa=K)
    ch = orc.ch(order, method, gmu)
    ole = orc.le(**le)
    phi0, u0 = fd_fields(nlocal)
    n = orc.nsite
    m = orc.interior_mask()
    with pyref.Ref(nvel, nlocal, fd=dict(order=order, method=method), le=le) as ref:
        assert ref.nhalo == 2
        ref.physics(eta_shear=eta, eta_bulk=eta, mobility=M)
        ref.fe_symm(A, B, K)
        ref.grad_mu(gmu)
        ref.grad_select(npt)
        for ts in (0, 1, 4, 67):
            ref.time_set(ts)
            t = ref.time()
            ref.inner(ref.phi())[...] = phi0.reshape(orc.nall)
            ref.inner(ref.u())[...] = u0.reshape(orc.nall + (3,))
            ref.phi_halo()
            ref.grad_compute()
            ref.hydro_f_zero()
            ref.phi_force()
            phi, u = phi0.copy(), u0.copy()
            force, grad, delsq = np.zeros((n, 3)), np.zeros((n, 3)), np.zeros(n)
            orc.ch_le_step(prm, ch, ole, t, npt, phi, u, force, grad, delsq, do_update=0)
            assert np.array_equal(grad[m], ref.inner(ref.grad()).reshape(-1, 3)[m]), ts
            assert np.array_equal(delsq[m], ref.inner(ref.delsq()).reshape(-1)[m]), ts
            fr = ref.inner(ref.force()).reshape(-1, 3)
            assert np.array_equal(force[m], fr[m]), ts
            if method != pyoracle.FORCE_NONE:
                assert np.max(np.abs(force[m])) > 0.0
            ref.cahn_hilliard()
            orc.ch_le_step(prm, ch, ole, t, npt, phi, u, force, grad, delsq, do_force=0)
            pr = ref.inner(ref.phi()).reshape(-1)
            assert np.max(np.abs(phi[m] - phi0[m])) > 1e-6
            assert np.array_equal(phi[m], pr[m]), ts
            assert abs(phi[m].sum() - phi0[m].sum()) < 1e-12 * m.sum()
        # the time loop
        nsteps = 6
        f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
        ref.inner(ref.f())[...] = f0.reshape(orc.nall + (1, nvel))
        ref.inner(ref.phi())[...] = phi0.reshape(orc.nall)
        ref.hydro_u_zero()              # the stages above left their velocity field behind
        ref.time_set(3)
        ref.step(nsteps)
        f, fp = f0.copy(), np.zeros_like(f0)
        rho, u, force = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        grad, delsq = np.zeros((n, 3)), np.zeros(n)
        phi = phi0.copy()
        ole.t_current, ole.t_start = 3, 3
        orc.gradient_select(npt)
        orc.ch_attach(ch)
        orc.le_attach(ole)
        try:
            orc.step(prm, nsteps, f, fp, rho, u, force, phi, grad, delsq)
        finally:
            orc.gradient_select(7)
            orc.ch_attach(None)
            orc.le_attach(None)
        assert relerr(f[m], ref.inner(ref.f()).reshape(f.shape)[m]) < 1e-12
        assert np.max(np.abs(phi[m] - ref.inner(ref.phi()).reshape(-1)[m])) < 1e-12


@pytest.mark.parametrize("nlocal,walls,order", [((10, 8, 6), (1, 0, 0), 1), ((8, 6, 10), (0, 1, 1), 2),
                                                ((6, 7, 8), (1, 1, 1), 1)])
def test_cahn_hilliard_between_walls(nlocal, walls, order):
    """The finite-difference order parameter between flat walls (serial-spin-fd2): the
    stress of the site itself at a face shared with solid and what crosses it booked to the
    walls (pth_force_map_kernel, pth_force_wall_kernel: src/phi_force_colloid.c:490-790), no
    flux through such faces (src/advection_bcs.c:38-140), the stencil next to solid; stage by
    stage bit for bit, then the time loop."""
    nvel = 19
    need_ref(nvel)
    A, B, K, M, eta = -0.00625, 0.00625, 0.004, 1.25, 0.00625
    periodic = tuple(0 if w else 1 for w in walls)
    orc = pyoracle.Oracle(nvel, nlocal)
    prm = pyoracle.make_param(eta_shear=eta, eta_bulk=eta, mobility=M, a=A, b=B, kappa=K)
    ch = orc.ch(order, pyoracle.FORCE_STRESS_DIVERGENCE, (0.0, 0.0, 0.0))
    phi0, u0 = fd_fields(nlocal)
    n = orc.nsite
    m = orc.interior_mask()
    with pyref.Ref(nvel, nlocal, fd=dict(order=order, method=2), periodic=periodic) as ref:
        ref.physics(eta_shear=eta, eta_bulk=eta, mobility=M)
        ref.fe_symm(A, B, K)
        ref.wall(walls)
        ref.grad_select(28)
        ref.inner(ref.phi())[...] = phi0.reshape(orc.nall)
        ref.inner(ref.u())[...] = u0.reshape(orc.nall + (3,))
        ref.phi_halo()
        ref.grad_compute()
        ref.hydro_f_zero()
        ref.phi_force()
        w = orc.wall(walls)
        phi, u = phi0.copy(), u0.copy()
        orc.halo(phi, 1)
        grad, delsq, force = np.zeros((n, 3)), np.zeros(n), np.zeros((n, 3))
        orc.grad_27pt_solid(w, phi, grad, delsq)
        assert np.array_equal(grad[m], ref.inner(ref.grad()).reshape(-1, 3)[m])
        orc.wall_attach(w)
        try:
            orc.phi_force(prm, ch, phi, grad, delsq, force)
            assert np.array_equal(force[m], ref.inner(ref.force()).reshape(-1, 3)[m])
            assert np.allclose(np.array(w.fnet[:]), ref.wall_momentum(), rtol=1e-12, atol=1e-18)
            ref.cahn_hilliard()
            orc.cahn_hilliard(prm, ch, phi, delsq, u)
        finally:
            orc.wall_attach(None)
        assert np.array_equal(phi[m], ref.inner(ref.phi()).reshape(-1)[m])
        assert abs(phi[m].sum() - phi0[m].sum()) < 1e-13*m.sum()      # nothing leaves through a wall


@pytest.mark.parametrize("order,method,npt,gmu", FD_CASES[:4])
def test_cahn_hilliard_time_loop(order, method, npt, gmu):
    nvel, nlocal, nsteps = 19, (8, 8, 8), 6
    need_ref(nvel)
    A, B, K, M, eta = -0.00625, 0.00625, 0.004, 1.25, 0.00625
    orc = pyoracle.Oracle(nvel, nlocal)
    prm = pyoracle.make_param(eta_shear=eta, eta_bulk=eta, mobility=M, a=A, b=B, kappa=K)
    ch = orc.ch(order, method, gmu)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    phi, _ = fd_fields(nlocal)
    n = orc.nsite
    with pyref.Ref(nvel, nlocal, fd=dict(order=order, method=method)) as ref:
        ref.physics(eta_shear=eta, eta_bulk=eta, mobility=M)
        ref.fe_symm(A, B, K)
        ref.grad_mu(gmu)
        ref.grad_select(npt)
        ref.inner(ref.f())[...] = f0.reshape(orc.nall + (1, nvel))
        ref.inner(ref.phi())[...] = phi.reshape(orc.nall)
        ref.step(nsteps)
        f, fp = f0.copy(), np.zeros_like(f0)
        rho, u, force = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        grad, delsq = np.zeros((n, 3)), np.zeros(n)
        orc.gradient_select(npt)
        orc.ch_attach(ch)
        try:
            orc.step(prm, nsteps, f, fp, rho, u, force, phi, grad, delsq)
        finally:
            orc.ch_attach(None)
            orc.gradient_select(7)
        m = orc.interior_mask()
        fr = ref.inner(ref.f()).reshape(n, 1, nvel)
        assert relerr(f[m], fr[m]) < 1e-12
        assert relerr(phi[m], ref.inner(ref.phi()).reshape(-1)[m]) < 1e-13
        assert relerr(u[m], ref.inner(ref.u()).reshape(-1, 3)[m]) < 1e-11


# ---------------------------------------------------------------------------
# viscosity model: lb_collide with hydro->eta (src/collision.c:386-404), visc_arrhenius.c
# ---------------------------------------------------------------------------

VISC = dict(eta_plus=0.625, eta_minus=0.00625, phistar=1.0)


@pytest.mark.parametrize("nrelax", [pyoracle.RELAX_M10, pyoracle.RELAX_BGK, pyoracle.RELAX_TRT])
def test_viscosity_model_collision(nrelax):
    """visc_arrhenius_update + one collision with a per-site viscosity, every scheme."""
    nvel, nlocal = 19, (10, 8, 6)
    need_ref(nvel)
    A, B, K, M, eta, zeta = -0.00625, 0.00625, 0.004, 1.25, 0.00625, 0.0125
    orc = pyoracle.Oracle(nvel, nlocal)
    prm = pyoracle.make_param(eta_shear=eta, eta_bulk=zeta, mobility=M, a=A, b=B, kappa=K,
                              nrelax=nrelax)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    force = lbinit.make_force(nlocal)
    phi, _ = fd_fields(nlocal)
    phi *= 2.0                                      # spans most of (-phistar, phistar)
    n = orc.nsite
    m = orc.interior_mask()
    with pyref.Ref(nvel, nlocal, fd=dict(order=1, method=2)) as ref:
        ref.physics(eta_shear=eta, eta_bulk=zeta, mobility=M)
        ref.fe_symm(A, B, K)
        ref.relaxation(nrelax)
        ref.visc_arrhenius(**VISC)
        ref.inner(ref.f())[...] = f0.reshape(orc.nall + (1, nvel))
        ref.inner(ref.force())[...] = force.reshape(orc.nall + (3,))
        ref.inner(ref.phi())[...] = phi.reshape(orc.nall)
        ref.visc_update()
        eta_site = np.zeros(n)
        orc.visc_arrhenius_update(VISC["eta_plus"], VISC["eta_minus"], VISC["phistar"], phi, eta_site)
        er = ref.inner(ref.eta()).reshape(-1)
        assert np.array_equal(eta_site[m], er[m])
        assert eta_site[m].max() / eta_site[m].min() > 10.0
        ref.collide()
        f = f0.copy()
        rho, u = np.zeros(n), np.zeros((n, 3))
        orc.visc_attach(eta_site)
        try:
            orc.collide_mrt(prm, f, force, rho, u)
        finally:
            orc.visc_attach(None)
        fr = ref.inner(ref.f()).reshape(n, 1, nvel)
        assert relerr(f[m], fr[m]) < 1e-13
        # and it matters: the uniform-viscosity collision differs
        g = f0.copy()
        orc.collide_mrt(prm, g, force, rho, u)
        assert relerr(g[m], fr[m]) > 1e-6


def test_viscosity_model_time_loop():
    """free_energy symmetric + viscosity_model arrhenius, six steps (src/ludwig.c:727-803)."""
    nvel, nlocal, nsteps = 19, (8, 8, 8), 6
    need_ref(nvel)
    A, B, K, M, eta = -0.00625, 0.00625, 0.004, 1.25, 0.00625
    orc = pyoracle.Oracle(nvel, nlocal)
    prm = pyoracle.make_param(eta_shear=eta, eta_bulk=eta, mobility=M, a=A, b=B, kappa=K)
    ch = orc.ch(2, pyoracle.FORCE_STRESS_DIVERGENCE)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    phi, _ = fd_fields(nlocal)
    phi *= 2.0
    n = orc.nsite
    with pyref.Ref(nvel, nlocal, fd=dict(order=2, method=2)) as ref:
        ref.physics(eta_shear=eta, eta_bulk=eta, mobility=M)
        ref.fe_symm(A, B, K)
        ref.grad_select(27)
        ref.visc_arrhenius(**VISC)
        ref.inner(ref.f())[...] = f0.reshape(orc.nall + (1, nvel))
        ref.inner(ref.phi())[...] = phi.reshape(orc.nall)
        ref.step(nsteps)
        f, fp = f0.copy(), np.zeros_like(f0)
        rho, u, force = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        grad, delsq, eta_site = np.zeros((n, 3)), np.zeros(n), np.zeros(n)
        orc.gradient_select(27)
        orc.ch_attach(ch)
        orc.visc_attach(eta_site, **VISC)
        try:
            orc.step(prm, nsteps, f, fp, rho, u, force, phi, grad, delsq)
        finally:
            orc.ch_attach(None)
            orc.visc_attach(None)
            orc.gradient_select(7)
        m = orc.interior_mask()
        fr = ref.inner(ref.f()).reshape(n, 1, nvel)
        assert relerr(f[m], fr[m]) < 1e-12
        assert relerr(phi[m], ref.inner(ref.phi()).reshape(-1)[m]) < 1e-13
        assert np.array_equal(eta_site[m], ref.inner(ref.eta()).reshape(-1)[m]) or \
            relerr(eta_site[m], ref.inner(ref.eta()).reshape(-1)[m]) < 1e-13
