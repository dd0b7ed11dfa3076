"""GPU: the order parameter of the binary fluid across Lees-Edwards planes --
field_leesedwards (src/field.c:500-600) and the plane-aware 7- and 27-point stencils
(lees_edw_ic_to_buff in src/gradient_3d_7pt_fluid.c:264-265, _27pt_fluid.c:262-265) --
against the oracle, which is pinned to the reference bit for bit
(tests/test_oracle_vs_ref.py::test_lees_edwards_order_parameter): gradients bit-exact,
the sheared binary time loop within 1e-12.

First run on a B200 in round 2 (profiles/r2_pytest_gpu_formerly_gated.log): all green, the
opt-in gate of round 1 is gone. The fourth-order advection scheme of the
finite-difference order parameter is at the end."""

import numpy as np
import pytest

import lbinit
from goldenlib import relerr

pytestmark = [pytest.mark.gpu]

TOL = 1e-12
KW = dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75, a=-0.00625, b=0.00625, kappa=0.004)
CASES = [((16, 6, 4), dict(nplanes=2, uy=0.05)),
         ((12, 7, 5), dict(nplanes=1, uy=-0.03)),
         ((12, 8, 4), dict(nplanes=3, uy=0.02)),
         ((12, 9, 140), dict(nplanes=3, uy=0.31)),
         ((16, 6, 4), dict(nplanes=2, uy=0.04, oscillatory=1, period=40, nt0=0))]


def commit(lat, le):
    lat.le_commit(le["nplanes"], le["uy"], le.get("oscillatory", 0), le.get("period", 0), le.get("nt0", 0))


@pytest.mark.parametrize("nlocal,le", CASES)
@pytest.mark.parametrize("npt", [7, 27])
@pytest.mark.parametrize("loopback", [0, 1])
def test_gradient_across_planes_bit_exact(nlocal, le, npt, loopback):
    import ludwig_b200 as lb
    from oracle import pyoracle
    orc = pyoracle.Oracle(19, nlocal, ndist=2)
    lat = lb.Lattice(19, nlocal, ndist=2)
    lat.set_option("loopback", loopback)
    lat.set_param(**KW)
    commit(lat, le)
    ole = orc.le(**le)
    n, nall = orc.nsite, orc.nall
    m = orc.interior_mask()
    rng = np.random.default_rng(3)
    phi = np.zeros(nall)
    phi[1:-1, 1:-1, 1:-1] = 0.4*(rng.random(nlocal) - 0.5)
    phi = phi.reshape(-1)
    lat.field_grad_set(npt)
    for t in (0.0, 1.0, 7.0, 133.0, 1000.0):
        lat.field_h2d(lb.FIELD_PHI, phi)
        lat.field_halo()
        lat.le_field_time_set(t)
        lat.field_grad_compute()
        ph = phi.copy()
        orc.halo(ph, 1)
        grad, delsq = np.zeros((n, 3)), np.zeros(n)
        orc.le_field_grad(ole, t, npt, ph, grad, delsq)
        assert np.array_equal(lat.field_d2h(lb.FIELD_GRAD)[m], grad[m]), t
        assert np.array_equal(lat.field_d2h(lb.FIELD_DELSQ)[m], delsq[m]), t
    lat.close()


@pytest.mark.parametrize("nlocal,le", CASES)
@pytest.mark.parametrize("npt", [7, 27])
@pytest.mark.parametrize("fused", [1, 0])
def test_sheared_binary_time_loop_vs_oracle(nlocal, le, npt, fused):
    import ludwig_b200 as lb
    from oracle import pyoracle
    orc = pyoracle.Oracle(19, nlocal, ndist=2)
    m = orc.interior_mask()
    n = orc.nsite
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=2)
    prm = pyoracle.make_param(**KW)
    ole = orc.le(t_current=5, t_start=5, **le)
    nsteps = 9
    f, fp = f0.copy(), np.zeros_like(f0)
    rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    phi, grad, delsq = np.zeros(n), np.zeros((n, 3)), np.zeros(n)
    orc.le_attach(ole)
    orc.gradient_select(npt)
    try:
        orc.step(prm, nsteps, f, fp, rho, u, frc, phi, grad, delsq)
    finally:
        orc.gradient_select(7)
        orc.le_attach(None)
    lat = lb.Lattice(19, nlocal, ndist=2)
    lat.set_param(**KW)
    lat.set_option("fused", fused)
    lat.set_option("gradient", npt)
    commit(lat, le)
    lat.le_time_set(5)
    lat.lb_memcpy_h2d(f0)
    lat.step(nsteps)
    fd = lat.lb_memcpy_d2h()
    assert relerr(fd[m], f[m]) < TOL
    assert np.max(np.abs(lat.field_d2h(lb.FIELD_GRAD)[m] - grad[m])) < TOL
    assert np.max(np.abs(lat.field_d2h(lb.FIELD_U)[m] - u[m])) < TOL
    assert abs(fd[m][:, 0].sum() - f0[m][:, 0].sum()) < 1e-9
    # the same through the separate calls, in the reference's order (src/ludwig.c:552-860)
    lat.le_time_set(5)
    lat.lb_memcpy_h2d(f0)
    for s in range(nsteps):
        step = 6 + s                                      # physics_control_next_step
        lat.hydro_f_zero()
        lat.phi_lb_to_field()
        lat.field_halo()
        lat.le_field_time_set(5 + step - 1.0)             # physics_control_time
        lat.field_grad_compute()
        lat.hydro_u_zero()
        lat.lb_collide()
        lat.le_apply_boundary_conditions(step)
        lat.lb_halo()
        lat.lb_propagation()
    assert np.array_equal(lat.lb_memcpy_d2h()[m], fd[m])
    lat.close()


# ---------------------------------------------------------------------------
# fourth-order advection of the finite-difference order parameter (ch_face<4>,
# ludwig_b200/csrc/ch_kernels.cu): also written after the GPU allocation was spent,
# checked on the host against the oracle (tools/check_ch_kernel_on_host.py)
# ---------------------------------------------------------------------------

@pytest.mark.parametrize("nlocal,npt", [((10, 8, 6), 7), ((6, 7, 9), 27)])
@pytest.mark.parametrize("exact", [1, 0])
def test_fourth_order_advection_vs_oracle(nlocal, npt, exact):
    import ludwig_b200 as lb
    import goldenlib
    from oracle import pyoracle
    P = goldenlib.CH_PHYS
    gmu = (1e-5, -2e-5, 3e-5)
    orc = pyoracle.Oracle(19, nlocal)
    n = orc.nsite
    m = orc.interior_mask()
    prm = pyoracle.make_param(eta_shear=P["eta"], eta_bulk=P["eta"], mobility=P["mobility"],
                              a=P["a"], b=P["b"], kappa=P["kappa"])
    ch = orc.ch(4, pyoracle.FORCE_STRESS_DIVERGENCE, gmu)
    phi, u = goldenlib.ch_fields(nlocal)
    lat = lb.Lattice(19, nlocal)
    lat.set_option("ch_exact", exact)
    lat.set_param(eta_shear=P["eta"], eta_bulk=P["eta"], mobility=P["mobility"],
                  a=P["a"], b=P["b"], kappa=P["kappa"])
    lat.phi_ch_create(4, lb.FORCE_STRESS_DIVERGENCE, gmu)
    lat.field_grad_set(npt)
    lat.field_h2d(lb.FIELD_PHI, phi)
    lat.field_h2d(lb.FIELD_U, u)
    lat.hydro_f_zero()
    lat.field_halo()
    lat.field_grad_compute()
    lat.phi_force_calculation()
    lat.phi_cahn_hilliard()
    force = lat.field_d2h(lb.FIELD_FORCE)
    phi1 = lat.field_d2h(lb.FIELD_PHI)
    po, uo = phi.copy(), u.copy()
    orc.halo(po, 1)
    grad, delsq, fo = np.zeros((n, 3)), np.zeros(n), np.zeros((n, 3))
    (orc.grad_27pt if npt == 27 else orc.grad_7pt)(po, grad, delsq)
    orc.phi_force(prm, ch, po, grad, delsq, fo)
    orc.cahn_hilliard(prm, ch, po, delsq, uo)
    if exact:
        assert np.array_equal(force[m], fo[m])
        assert np.array_equal(phi1[m], po[m])
    else:
        assert np.max(np.abs(force[m] - fo[m])) <= 1e-15*np.max(np.abs(fo[m]))
        assert np.max(np.abs(phi1[m] - po[m])) <= 1e-15*np.max(np.abs(po[m]))
    assert abs(phi1[m].sum() - phi[m].sum()) < 1e-13*m.sum()
    lat.close()


# ---------------------------------------------------------------------------
# the finite-difference order parameter across the planes (option le_fd; ch_le_pass_a/_b,
# checked on the host by tools/check_ch_le_on_host.py)
# ---------------------------------------------------------------------------

FD_LE = [((16, 8, 6), dict(nplanes=2, uy=0.05), 1, 7),
         ((16, 8, 6), dict(nplanes=2, uy=0.05), 2, 27),
         ((12, 7, 5), dict(nplanes=1, uy=-0.03), 3, 27),
         ((12, 8, 6), dict(nplanes=2, uy=0.07), 4, 7),
         ((16, 8, 140), dict(nplanes=2, uy=0.04, oscillatory=1, period=40, nt0=0), 2, 7)]


@pytest.mark.parametrize("nlocal,le,order,npt", FD_LE)
@pytest.mark.parametrize("exact,fuse", [(1, 1), (0, 1), (0, 0)])
def test_finite_difference_order_parameter_across_planes(nlocal, le, order, npt, exact, fuse):
    """Stage parity (force 1e-15 of its largest value, phi bit for bit with ch_exact) at
    several times, then the sheared time loop within 1e-12."""
    import ludwig_b200 as lb
    import goldenlib
    from oracle import pyoracle
    P = goldenlib.CH_PHYS
    gmu = (1e-5, -2e-5, 3e-5)
    kw = dict(eta_shear=P["eta"], eta_bulk=P["eta"], mobility=P["mobility"], a=P["a"], b=P["b"], kappa=P["kappa"])
    orc = pyoracle.Oracle(19, nlocal)
    n = orc.nsite
    m = orc.interior_mask()
    prm = pyoracle.make_param(**kw)
    ch = orc.ch(order, pyoracle.FORCE_STRESS_DIVERGENCE, gmu)
    ole = orc.le(**le)
    phi0, u0 = goldenlib.ch_fields(nlocal)
    lat = lb.Lattice(19, nlocal)
    lat.set_option("ch_exact", exact)
    lat.set_option("ch_fuse", fuse)
    lat.set_param(**kw)
    commit(lat, le)
    lat.phi_ch_create(order, lb.FORCE_STRESS_DIVERGENCE, gmu)
    lat.field_grad_set(npt)
    for t in (0.0, 1.0, 6.0, 133.0):
        lat.field_h2d(lb.FIELD_PHI, phi0)
        lat.field_h2d(lb.FIELD_U, u0)
        lat.hydro_f_zero()
        lat.field_halo()
        lat.le_field_time_set(t)
        lat.field_grad_compute()
        lat.phi_force_calculation()
        lat.phi_cahn_hilliard()
        force = lat.field_d2h(lb.FIELD_FORCE)
        phi1 = lat.field_d2h(lb.FIELD_PHI)
        po, uo = phi0.copy(), u0.copy()
        fo, grad, delsq = np.zeros((n, 3)), np.zeros((n, 3)), np.zeros(n)
        orc.ch_le_step(prm, ch, ole, t, npt, po, uo, fo, grad, delsq)
        assert np.max(np.abs(force[m] - fo[m])) <= 1e-15*np.max(np.abs(fo[m])), t
        if exact:
            assert np.array_equal(phi1[m], po[m]), t
        else:
            assert np.max(np.abs(phi1[m] - po[m])) <= 1e-15*np.max(np.abs(po[m])), t
        assert abs(phi1[m].sum() - phi0[m].sum()) < 1e-13*m.sum()
    # the time loop
    nsteps = 6
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
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
    lat.le_time_set(3)
    lat.lb_memcpy_h2d(f0)
    lat.field_h2d(lb.FIELD_PHI, phi0)
    lat.field_h2d(lb.FIELD_U, np.zeros((n, 3)))
    lat.step(nsteps)
    assert relerr(lat.lb_memcpy_d2h()[m], f[m]) < TOL
    assert np.max(np.abs(lat.field_d2h(lb.FIELD_PHI)[m] - phi[m])) < TOL
    lat.close()


# ---------------------------------------------------------------------------
# open boundaries for the fluid (open_bc_kernel; tools/check_open_bc_on_host.py)
# ---------------------------------------------------------------------------

@pytest.mark.parametrize("nvel,nlocal,walls,u0", [
    (19, (24, 12, 12), (0, 1, 1), (0.0, 0.0, 0.0)),
    (19, (12, 12, 24), (1, 1, 0), (0.0, 0.0, 0.003)),
    (27, (10, 8, 6), (0, 1, 1), (0.004, 0.0, 0.0)),
    (19, (7, 9, 140), (1, 0, 1), (0.0, 0.002, 0.0))])
@pytest.mark.parametrize("fused,hydro_write", [(1, 1), (0, 1), (1, 0)])
def test_open_boundaries_vs_oracle(nvel, nlocal, walls, u0, fused, hydro_write):
    """lb_bc_open: one application against the oracle bit for bit (the imposed populations),
    then the time loop with walls across the flow within 1e-12."""
    import ludwig_b200 as lb
    from oracle import pyoracle
    flow = walls.index(0)
    periodic = (0, 0, 0)
    orc = pyoracle.Oracle(nvel, nlocal)
    n = orc.nsite
    m = orc.interior_mask()
    kw = dict(eta_shear=0.1, eta_bulk=0.1)
    prm = pyoracle.make_param(**kw)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
    lat = lb.Lattice(nvel, nlocal, periodic=periodic)
    lat.set_option("fused", fused)
    lat.set_option("hydro_write", hydro_write)
    lat.set_param(**kw)
    lat.wall_commit(walls)
    lat.open_bc_commit(flow, u0, 1.0)
    o = orc.open_bc(flow, u0, 1.0)
    w = orc.wall(walls)
    nsteps = 10
    f, fp = f0.copy(), np.zeros_like(f0)
    rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    orc.wall_attach(w)
    orc.open_attach(o)
    try:
        orc.step(prm, nsteps, f, fp, rho, u, frc)
    finally:
        orc.wall_attach(None)
        orc.open_attach(None)
    lat.lb_memcpy_h2d(f0)
    lat.step(nsteps)
    fd = lat.lb_memcpy_d2h()
    assert relerr(fd[m], f[m]) < TOL
    assert np.max(np.abs(lat.field_d2h(lb.FIELD_U)[m] - u[m])) < TOL
    assert abs(fd[m].sum() - f0[m].sum()) > 1e-9 or u0 == (0.0, 0.0, 0.0)     # an open system
    # the same through the separate calls, in the reference's order
    lat.lb_memcpy_h2d(f0)
    for s in range(nsteps):
        lat.hydro_f_zero()
        lat.hydro_u_zero()
        lat.lb_collide()
        lat.lb_halo()
        lat.open_bc_apply()
        lat.wall_bbl()
        lat.lb_propagation()
    assert np.array_equal(lat.lb_memcpy_d2h()[m], fd[m])
    lat.close()


# ---------------------------------------------------------------------------
# the binary fluid between flat walls (option binary_walls): bounce-back of both
# distributions folded into a propagation of its own, grad27_solid_kernel
# (tools/check_solid_grad_on_host.py, tools/sweep_emulate.cpp)
# ---------------------------------------------------------------------------

@pytest.mark.parametrize("nvel,nlocal,walls", [
    (19, (10, 8, 6), (1, 1, 1)), (19, (8, 6, 40), (0, 1, 1)), (27, (6, 7, 9), (1, 0, 0)),
    (19, (12, 10, 8), (0, 0, 1))])
@pytest.mark.parametrize("hydro_write", [1, 0])
def test_binary_fluid_between_walls_vs_oracle(nvel, nlocal, walls, hydro_write):
    import ctypes
    import ludwig_b200 as lb
    from oracle import pyoracle
    periodic = tuple(0 if w else 1 for w in walls)
    orc = pyoracle.Oracle(nvel, nlocal, ndist=2)
    n = orc.nsite
    m = orc.interior_mask()
    prm = pyoracle.make_param(**KW)
    f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=2)
    lat = lb.Lattice(nvel, nlocal, ndist=2, periodic=periodic)
    lat.set_option("hydro_write", hydro_write)
    lat.set_param(**KW)
    lat.wall_commit(walls)
    lat.field_grad_set(28)
    # the stencil next to the walls, bit for bit
    lat.lb_memcpy_h2d(f0)
    lat.phi_lb_to_field()
    lat.field_halo()
    lat.field_grad_compute()
    phi, grad, delsq = np.zeros(n), np.zeros((n, 3)), np.zeros(n)
    orc.phi_from_g(f0, phi)
    orc.halo(phi, 1)
    w = orc.wall(walls)
    dp = ctypes.POINTER(ctypes.c_double)
    orc.L.orc_grad_27pt_solid.argtypes = [ctypes.c_void_p, ctypes.c_void_p, dp, dp, dp]
    assert orc.L.orc_grad_27pt_solid(ctypes.byref(orc.box), ctypes.byref(w), phi.ctypes.data_as(dp),
                                     grad.ctypes.data_as(dp), delsq.ctypes.data_as(dp)) == 0
    assert np.array_equal(lat.field_d2h(lb.FIELD_GRAD)[m], grad[m])
    assert np.array_equal(lat.field_d2h(lb.FIELD_DELSQ)[m], delsq[m])
    # the time loop
    nsteps = 8
    f, fp = f0.copy(), np.zeros_like(f0)
    rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
    orc.gradient_select(28)
    orc.wall_attach(w)
    try:
        orc.step(prm, nsteps, f, fp, rho, u, frc, phi, grad, delsq)
    finally:
        orc.gradient_select(7)
        orc.wall_attach(None)
    lat.lb_memcpy_h2d(f0)
    lat.step(nsteps)
    fd = lat.lb_memcpy_d2h()
    assert relerr(fd[m], f[m]) < TOL
    assert np.max(np.abs(lat.field_d2h(lb.FIELD_U)[m] - u[m])) < TOL
    assert abs(fd[m][:, 0].sum() - f0[m][:, 0].sum()) < 1e-9          # closed to mass ...
    assert abs(fd[m][:, 1].sum() - f0[m][:, 1].sum()) < 1e-9          # ... and to composition
    assert np.allclose(lat.wall_momentum(), np.array(w.fnet[:]), rtol=1e-9, atol=1e-12)
    lat.close()
