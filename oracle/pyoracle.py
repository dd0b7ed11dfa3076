"""ctypes wrapper of oracle/liblb_oracle.so -- TEST INFRASTRUCTURE ONLY.

The CPU restatement of the Ludwig LB hot path (oracle/lb_oracle.c). Only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm
may import this module; the product (ludwig_b200/) never does.

Arrays use the reference's host layout (see oracle/lb_oracle.h).
"""

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liblb_oracle.so")

NVELMAX = 27
RELAX_M10, RELAX_BGK, RELAX_TRT = 0, 1, 2

c_double_p = ctypes.POINTER(ctypes.c_double)


class Model(ctypes.Structure):
    _fields_ = [
        ("ndim", ctypes.c_int),
        ("nvel", ctypes.c_int),
        ("nhydro", ctypes.c_int),
        ("cv", (ctypes.c_int * 3) * NVELMAX),
        ("wv", ctypes.c_double * NVELMAX),
        ("na", ctypes.c_double * NVELMAX),
        ("ma", (ctypes.c_double * NVELMAX) * NVELMAX),
        ("mi", (ctypes.c_double * NVELMAX) * NVELMAX),
        ("maf", (ctypes.c_double * NVELMAX) * NVELMAX),
    ]


class Param(ctypes.Structure):
    _fields_ = [
        ("rho0", ctypes.c_double),
        ("eta_shear", ctypes.c_double),
        ("eta_bulk", ctypes.c_double),
        ("force_global", ctypes.c_double * 3),
        ("nrelax", ctypes.c_int),
        ("mobility", ctypes.c_double),
        ("a", ctypes.c_double),
        ("b", ctypes.c_double),
        ("kappa", ctypes.c_double),
        ("ghost_off", ctypes.c_int),
    ]


class Wall(ctypes.Structure):
    _fields_ = [
        ("isboundary", ctypes.c_int * 3),
        ("ubot", ctypes.c_double * 3),
        ("utop", ctypes.c_double * 3),
        ("fnet", ctypes.c_double * 3),
        ("slip_active", ctypes.c_int),
        ("sbot", ctypes.c_double * 3),
        ("stop", ctypes.c_double * 3),
    ]


class OpenBC(ctypes.Structure):
    _fields_ = [("id", ctypes.c_int), ("u0", ctypes.c_double * 3), ("rho0", ctypes.c_double)]


class LeesEdwards(ctypes.Structure):
    _fields_ = [
        ("nplanes", ctypes.c_int),
        ("uy", ctypes.c_double),
        ("oscillatory", ctypes.c_int),
        ("period", ctypes.c_int),
        ("nt0", ctypes.c_int),
        ("t_current", ctypes.c_int),
        ("t_start", ctypes.c_int),
    ]


FORCE_NONE, FORCE_STRESS_DIVERGENCE, FORCE_PHI_GRADMU = 1, 2, 3


class CahnHilliard(ctypes.Structure):
    _fields_ = [
        ("order", ctypes.c_int),
        ("method", ctypes.c_int),
        ("grad_mu", ctypes.c_double * 3),
    ]


class Visc(ctypes.Structure):
    _fields_ = [("eta_plus", ctypes.c_double), ("eta_minus", ctypes.c_double),
                ("phistar", ctypes.c_double)]


class Box(ctypes.Structure):
    _fields_ = [
        ("nlocal", ctypes.c_int * 3),
        ("nall", ctypes.c_int * 3),
        ("nsite", ctypes.c_long),
    ]


def build():
    """Compile the restatement (gcc only)."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        _lib = ctypes.CDLL(_LIB)
        _lib.orc_index.restype = ctypes.c_long
        _lib.orc_noise_uniform.restype = ctypes.c_uint
        _lib.orc_noise_attach.argtypes = [ctypes.c_void_p, ctypes.c_double]
        _lib.orc_noise_reap_n.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_int,
                                          c_double_p]
    return _lib


def _p(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_double_p)


def make_param(rho0=1.0, eta_shear=0.1, eta_bulk=0.1, force=(0.0, 0.0, 0.0),
               nrelax=RELAX_M10, mobility=0.0, a=0.0, b=0.0, kappa=0.0, ghost_off=0):
    prm = Param()
    prm.ghost_off = ghost_off
    prm.rho0, prm.eta_shear, prm.eta_bulk = rho0, eta_shear, eta_bulk
    prm.force_global[:] = force
    prm.nrelax = nrelax
    prm.mobility, prm.a, prm.b, prm.kappa = mobility, a, b, kappa
    return prm


class Oracle:
    """One periodic local box + one velocity set."""

    def __init__(self, nvel, nlocal, ndist=1):
        self.L = lib()
        self.model = Model()
        if self.L.orc_model_init(nvel, ctypes.byref(self.model)) != 0:
            raise ValueError("bad nvel %d" % nvel)
        self.box = Box()
        self.L.orc_box_init((ctypes.c_int * 3)(*nlocal), ctypes.byref(self.box))
        self.nvel = nvel
        self.ndim = self.model.ndim
        self.ndist = ndist
        self.nlocal = tuple(nlocal)
        self.nall = tuple(self.box.nall)
        self.nsite = int(self.box.nsite)

    # model tables as numpy
    def cv(self):
        return np.array([[self.model.cv[p][a] for a in range(3)]
                         for p in range(self.nvel)], dtype=np.int64)

    def wv(self):
        return np.array(self.model.wv[:self.nvel])

    def ma(self):
        return np.array([[self.model.ma[m][p] for p in range(self.nvel)]
                         for m in range(self.nvel)])

    def maf(self):
        return np.array([[self.model.maf[m][p] for p in range(self.nvel)]
                         for m in range(self.nvel)])

    def mi(self):
        return np.array([[self.model.mi[p][m] for m in range(self.nvel)]
                         for p in range(self.nvel)])

    def index(self, ic, jc, kc):
        return (ic * self.nall[1] + jc) * self.nall[2] + kc

    def interior_mask(self):
        m = np.zeros(self.nall, dtype=bool)
        m[1:-1, 1:-1, 1:-1] = True
        return m.reshape(-1)

    def alloc_f(self):
        return np.zeros((self.nsite, self.ndist, self.nvel))

    def rtau(self, prm):
        out = (ctypes.c_double * NVELMAX)()
        rc = self.L.orc_rtau(ctypes.byref(self.model), ctypes.byref(prm), out)
        if rc != 0:
            raise ValueError("relaxation scheme not available")
        return np.array(out[:self.nvel])

    def equilib_set(self, f, index, rho, u):
        self.L.orc_equilib_set(ctypes.byref(self.model), self.ndist, _p(f),
                               ctypes.c_long(index), ctypes.c_double(rho),
                               (ctypes.c_double * 3)(*u))

    def collide_mrt(self, prm, f, force, rho, u):
        rc = self.L.orc_collide_mrt(ctypes.byref(self.model),
                                    ctypes.byref(self.box), ctypes.byref(prm),
                                    self.ndist, _p(f), _p(force), _p(rho), _p(u))
        assert rc == 0

    def collide_binary(self, prm, f, force, u, phi, grad, delsq):
        rc = self.L.orc_collide_binary(ctypes.byref(self.model),
                                       ctypes.byref(self.box),
                                       ctypes.byref(prm), _p(f), _p(force),
                                       _p(u), _p(phi), _p(grad), _p(delsq))
        assert rc == 0

    def halo(self, data, nfel):
        self.L.orc_halo_periodic(ctypes.byref(self.box), nfel, _p(data))

    def propagate(self, f, fprime):
        self.L.orc_propagate(ctypes.byref(self.model), ctypes.byref(self.box),
                             self.ndist, _p(f), _p(fprime))

    def phi_from_g(self, f, phi):
        self.L.orc_phi_from_g(ctypes.byref(self.model), ctypes.byref(self.box),
                              _p(f), _p(phi))

    def grad_7pt(self, phi, grad, delsq):
        self.L.orc_grad_7pt(ctypes.byref(self.box), _p(phi), _p(grad),
                            _p(delsq))

    def grad_27pt(self, phi, grad, delsq):
        self.L.orc_grad_27pt(ctypes.byref(self.box), _p(phi), _p(grad),
                             _p(delsq))

    def grad_27pt_solid(self, wall, phi, grad, delsq):
        """grad_3d_27pt_solid_d2 next to the flat walls of `wall` (None: none), neutral wetting."""
        self.L.orc_grad_27pt_solid.argtypes = [ctypes.c_void_p, ctypes.c_void_p, c_double_p,
                                               c_double_p, c_double_p]
        assert self.L.orc_grad_27pt_solid(ctypes.byref(self.box),
                                          None if wall is None else ctypes.byref(wall),
                                          _p(phi), _p(grad), _p(delsq)) == 0

    def gradient_select(self, npt):
        assert self.L.orc_gradient_select(int(npt)) == 0

    # -- Lees-Edwards planes (oracle/lb_oracle_le.c) --------------------------------
    # -- open boundaries (inflow / outflow "rhou") ------------------------------------
    def open_bc(self, flow_dir, u0=(0.0, 0.0, 0.0), rho0=1.0):
        o = OpenBC()
        o.id, o.rho0 = int(flow_dir), float(rho0)
        o.u0[:] = u0
        return o

    def open_attach(self, o):
        self._open = o
        self.L.orc_open_attach(None if o is None else ctypes.byref(o))

    def open_apply(self, o, rho, u, f):
        self.L.orc_open_apply.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                          ctypes.c_void_p, c_double_p, c_double_p, c_double_p]
        assert self.L.orc_open_apply(ctypes.byref(self.model), ctypes.byref(self.box), self.ndist,
                                     ctypes.byref(o), _p(rho), _p(u), _p(f)) == 0

    def le(self, nplanes, uy, oscillatory=0, period=0, nt0=0, t_current=0, t_start=0):
        le = LeesEdwards()
        le.nplanes, le.uy, le.oscillatory = nplanes, uy, oscillatory
        le.period, le.nt0, le.t_current, le.t_start = period, nt0, t_current, t_start
        return le

    def le_attach(self, le):
        self._le = le
        self.L.orc_le_attach(None if le is None else ctypes.byref(le))

    def le_apply(self, le, t, f):
        self.L.orc_le_apply.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                        ctypes.c_void_p, ctypes.c_double, c_double_p]
        rc = self.L.orc_le_apply(ctypes.byref(self.model), ctypes.byref(self.box), self.ndist,
                                 ctypes.byref(le), float(t), _p(f))
        assert rc == 0

    def le_field_buffers(self, le, t, phi):
        """field_leesedwards: buf[2 nplanes, nall_y, nall_z]."""
        buf = np.zeros((2*le.nplanes, self.nall[1], self.nall[2]))
        self.L.orc_le_field_buffers.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double,
                                                c_double_p, c_double_p]
        assert self.L.orc_le_field_buffers(ctypes.byref(self.box), ctypes.byref(le), float(t),
                                           _p(phi), _p(buf)) == 0
        return buf

    def le_field_grad(self, le, t, npt, phi, grad, delsq):
        """field_grad_compute with planes (field_leesedwards + the stencil across them)."""
        self.L.orc_le_field_grad.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_double,
                                             ctypes.c_int, c_double_p, c_double_p, c_double_p]
        assert self.L.orc_le_field_grad(ctypes.byref(self.box), ctypes.byref(le), float(t), int(npt),
                                        _p(phi), _p(grad), _p(delsq)) == 0

    def le_plane_location(self, le, n):
        return self.L.orc_le_plane_location(ctypes.byref(self.box), ctypes.byref(le), int(n))

    # -- finite-difference order parameter (oracle/lb_oracle_ch.c) -------------------
    def ch(self, order=1, method=FORCE_STRESS_DIVERGENCE, grad_mu=(0.0, 0.0, 0.0)):
        c = CahnHilliard()
        c.order, c.method = order, method
        c.grad_mu[:] = grad_mu
        return c

    def ch_le_step(self, prm, ch, le, t, npt, phi, u, force, grad, delsq, do_force=1, do_update=1):
        """One step of the finite-difference order parameter across Lees-Edwards planes."""
        self.L.orc_ch_le_step.argtypes = [ctypes.c_void_p]*4 + [ctypes.c_double, ctypes.c_int,
                                                                ctypes.c_int, ctypes.c_int] + [c_double_p]*5
        rc = self.L.orc_ch_le_step(ctypes.byref(self.box), ctypes.byref(prm), ctypes.byref(ch),
                                   ctypes.byref(le), float(t), int(npt), int(do_force), int(do_update),
                                   _p(phi), _p(u), _p(force), _p(grad), _p(delsq))
        assert rc == 0

    def ch_attach(self, ch):
        """orc_step (ndist = 1) runs the "free_energy symmetric" route while attached."""
        self._ch = ch
        self.L.orc_ch_attach(None if ch is None else ctypes.byref(ch))

    def ch_stress(self, prm, phi, grad, delsq):
        pth = np.zeros((self.nsite, 3, 3))
        self.L.orc_ch_stress(ctypes.byref(self.box), ctypes.byref(prm), _p(phi), _p(grad),
                             _p(delsq), _p(pth))
        return pth

    def ch_advection(self, order, u, phi):
        flux = np.zeros((4, self.nsite))
        rc = self.L.orc_ch_advection(ctypes.byref(self.box), int(order), _p(u), _p(phi), _p(flux))
        assert rc == 0
        return flux

    def phi_force(self, prm, ch, phi, grad, delsq, force):
        """phi_force_calculation: adds to force; fills the halos of grad and delsq."""
        rc = self.L.orc_phi_force(ctypes.byref(self.box), ctypes.byref(prm), ctypes.byref(ch),
                                  _p(phi), _p(grad), _p(delsq), _p(force))
        assert rc == 0

    def cahn_hilliard(self, prm, ch, phi, delsq, u):
        """phi_cahn_hilliard: u halo, fluxes, forward step of phi (interior)."""
        rc = self.L.orc_cahn_hilliard(ctypes.byref(self.box), ctypes.byref(prm), ctypes.byref(ch),
                                      _p(phi), _p(delsq), _p(u))
        assert rc == 0

    # -- viscosity model (src/visc_arrhenius.c, src/collision.c:386-404) --------------
    def visc_attach(self, eta, eta_plus=None, eta_minus=None, phistar=1.0):
        """Collisions take each site's viscosity from eta[nsite] (None: off); with the
        Arrhenius parameters orc_step refreshes eta from phi ahead of each collision."""
        self._eta = eta
        self._visc = None
        if eta is None:
            self.L.orc_visc_attach(None, None)
            return
        if eta_plus is not None:
            self._visc = Visc(eta_plus, eta_minus, phistar)
        self.L.orc_visc_attach(None if self._visc is None else ctypes.byref(self._visc), _p(eta))

    def visc_arrhenius_update(self, eta_plus, eta_minus, phistar, phi, eta):
        v = Visc(eta_plus, eta_minus, phistar)
        self.L.orc_visc_arrhenius_update(ctypes.byref(self.box), ctypes.byref(v), _p(phi), _p(eta))

    # -- flat walls (oracle/lb_oracle_wall.c) ---------------------------------------
    def wall(self, isboundary, ubot=(0.0, 0.0, 0.0), utop=(0.0, 0.0, 0.0), sbot=None, stop=None,
             slip_active=None):
        w = Wall()
        w.isboundary[:] = isboundary
        w.ubot[:] = ubot
        w.utop[:] = utop
        w.sbot[:] = sbot or (0.0, 0.0, 0.0)
        w.stop[:] = stop or (0.0, 0.0, 0.0)
        # wall_slip(), src/wall.c:284-294: active if any fraction is non-zero (or forced)
        active = any(x != 0.0 for x in list(w.sbot) + list(w.stop))
        w.slip_active = int(active if slip_active is None else slip_active)
        return w

    def wall_attach(self, wall):
        """orc_step applies bounce-back on links while a wall is attached (None: off)."""
        self._wall = wall                      # keep it alive
        self.L.orc_wall_attach(None if wall is None else ctypes.byref(wall))

    def wall_nlink(self, wall):
        return self.L.orc_wall_nlink(ctypes.byref(self.model), ctypes.byref(self.box),
                                     ctypes.byref(wall))

    def wall_bbl(self, prm, wall, f):
        self.L.orc_wall_bbl(ctypes.byref(self.model), ctypes.byref(self.box), ctypes.byref(prm),
                            self.ndist, ctypes.byref(wall), _p(f))

    # -- isothermal fluctuations (oracle/lb_oracle_noise.c) ------------------------
    def noise_init(self, master_seed=0, noffset=(0, 0, 0), ntotal=None):
        """KISS99 state of every site, state[index, 0:4] (uint32)."""
        ntotal = self.nlocal if ntotal is None else ntotal
        state = np.zeros((self.nsite, 4), dtype=np.uint32)
        rc = self.L.orc_noise_init(ctypes.byref(self.box), int(master_seed),
                                   (ctypes.c_int * 3)(*noffset), (ctypes.c_int * 3)(*ntotal),
                                   state.ctypes.data_as(ctypes.c_void_p))
        assert rc == 0
        return state

    def noise_attach(self, state, kt):
        """Collisions from now on fluctuate at temperature kt (state=None: off)."""
        if state is None:
            self.L.orc_noise_attach(None, 0.0)
        else:
            assert state.dtype == np.uint32 and state.flags["C_CONTIGUOUS"]
            self.L.orc_noise_attach(state.ctypes.data_as(ctypes.c_void_p), float(kt))

    def noise_reap_n(self, state, index, nmax):
        out = np.zeros(nmax)
        rc = self.L.orc_noise_reap_n(state.ctypes.data_as(ctypes.c_void_p), int(index),
                                     int(nmax), _p(out))
        assert rc == 0
        return out

    def stats(self, f):
        out = (ctypes.c_double * 8)()
        self.L.orc_stats(ctypes.byref(self.model), ctypes.byref(self.box),
                         self.ndist, _p(f), out)
        return np.array(out[:])

    def step(self, prm, nsteps, f, fprime, rho, u, force, phi=None, grad=None,
             delsq=None):
        rc = self.L.orc_step(ctypes.byref(self.model), ctypes.byref(self.box),
                             ctypes.byref(prm), self.ndist, nsteps, _p(f),
                             _p(fprime), _p(rho), _p(u), _p(force), _p(phi),
                             _p(grad), _p(delsq))
        assert rc == 0
