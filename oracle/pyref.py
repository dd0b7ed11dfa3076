"""ctypes wrapper of oracle/_ref/libref_<model>.so -- TEST INFRASTRUCTURE ONLY.

The UNMODIFIED reference (Ludwig v0.20.1) compiled in place from
/root/reference by oracle/Makefile plus oracle/ref_driver.c. Present only
where it was built (this container; the GPU box gets the prebuilt file with
the snapshot). `available(nvel)` says whether it can be used.

The reference keeps global state (physics singleton), so only one Ref object
per model may be alive at a time; use it as a context manager.
"""

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_NAMES = {19: "d3q19", 27: "d3q27", 9: "d2q9"}
_libs = {}

HALO_TARGET, HALO_OPENMP_FULL, HALO_OPENMP_REDUCED = 0, 1, 2
c_double_p = ctypes.POINTER(ctypes.c_double)


def _path(nvel):
    return os.path.join(_HERE, "_ref", "libref_%s.so" % _NAMES[nvel])


def build():
    """Build the reference libraries if the reference sources are here."""
    if os.path.isdir("/root/reference/src"):
        ncpu = str(os.cpu_count() or 4)
        subprocess.check_call(["make", "-s", "-j", ncpu, "-C", _HERE, "ref"],
                              stdout=subprocess.DEVNULL)


def available(nvel=19):
    return nvel in _NAMES and os.path.exists(_path(nvel))


def _lib(nvel):
    if nvel not in _libs:
        L = ctypes.CDLL(_path(nvel))
        for name in ("refdrv_f", "refdrv_rho", "refdrv_u", "refdrv_force",
                     "refdrv_phi", "refdrv_grad", "refdrv_delsq", "refdrv_eta"):
            getattr(L, name).restype = c_double_p
        L.refdrv_status.restype = ctypes.POINTER(ctypes.c_char)
        L.refdrv_step.restype = ctypes.c_double
        L.refdrv_noise_state.restype = ctypes.POINTER(ctypes.c_uint)
        L.refdrv_noise.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_int]
        L.refdrv_noise_reap_n.argtypes = [ctypes.c_int, ctypes.c_int, c_double_p]
        _libs[nvel] = L
    return _libs[nvel]


class Ref:
    """Reference objects on a periodic one-rank box ntotal, halo width 1."""

    def __init__(self, nvel, ntotal, ndist=1, halo=HALO_TARGET, periodic=(1, 1, 1), le=None,
                 fd=None):
        """le = dict(nplanes, uy, oscillatory=0, period=0, nt0=0): Lees-Edwards planes.
        fd = dict(order=1, method=2): "free_energy symmetric" -- halo width 2, phi a
        finite-difference field (phi_cahn_hilliard + phi_force_calculation);
        method = fe_force_method_enum_t (1 none, 2 stress divergence, 3 phi grad mu)."""
        self.L = _lib(nvel)
        assert self.L.refdrv_nvel() == nvel
        self.nvel, self.ndist = nvel, ndist
        self.L.refdrv_periodicity((ctypes.c_int * 3)(*periodic))
        if le:
            self.L.refdrv_le_options.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_int,
                                                 ctypes.c_int, ctypes.c_int]
            self.L.refdrv_le_options(int(le["nplanes"]), float(le["uy"]),
                                     int(le.get("oscillatory", 0)), int(le.get("period", 0)),
                                     int(le.get("nt0", 0)))
        if fd:
            self.L.refdrv_fd_options(int(fd.get("order", 1)), int(fd.get("method", 2)))
        self.L.refdrv_init((ctypes.c_int * 3)(*ntotal), ndist, halo)
        self.nhalo = self.L.refdrv_nhalo()
        nall = (ctypes.c_int * 3)()
        self.L.refdrv_nall(nall)
        self.nall = tuple(nall)
        self.nsite = self.L.refdrv_nsite()
        self.nlocal = tuple(ntotal)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def close(self):
        if self.L is not None:
            self.L.refdrv_finish()
            self.L = None

    def _arr(self, ptr, shape):
        return np.ctypeslib.as_array(ptr, shape=shape)

    # Views of the reference's own host arrays. NOTE: f's address changes at
    # every lb_propagation (pointer swap, src/propagation.c:223), so call f()
    # again after a propagation.
    def f(self):
        return self._arr(self.L.refdrv_f(), (self.nsite, self.ndist, self.nvel))

    def rho(self):
        return self._arr(self.L.refdrv_rho(), (self.nsite,))

    def u(self):
        return self._arr(self.L.refdrv_u(), (self.nsite, 3))

    def force(self):
        return self._arr(self.L.refdrv_force(), (self.nsite, 3))

    def phi(self):
        return self._arr(self.L.refdrv_phi(), (self.nsite,))

    def phi_all(self):
        """phi with the buffer planes of field_leesedwards behind the nall_x real ones."""
        return self._arr(self.L.refdrv_phi(), (self.L.refdrv_phi_nsites(),))

    def grad(self):
        return self._arr(self.L.refdrv_grad(), (self.nsite, 3))

    def delsq(self):
        return self._arr(self.L.refdrv_delsq(), (self.nsite,))

    def index(self, ic, jc, kc):
        return self.L.refdrv_index(ic, jc, kc)

    def addr_f(self, index, n, p):
        return self.L.refdrv_addr_f(index, n, p)

    def model(self):
        q = self.nvel
        wv = np.zeros(q)
        ma = np.zeros((q, q))
        mi = np.zeros((q, q))
        cv = np.zeros((q, 3), dtype=np.int32)
        self.L.refdrv_model(wv.ctypes.data_as(c_double_p),
                            ma.ctypes.data_as(c_double_p),
                            mi.ctypes.data_as(c_double_p),
                            cv.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
        return wv, ma, mi, cv

    def physics(self, rho0=1.0, eta_shear=0.1, eta_bulk=0.1,
                force=(0.0, 0.0, 0.0), mobility=0.0):
        self.L.refdrv_physics(ctypes.c_double(rho0), ctypes.c_double(eta_shear),
                              ctypes.c_double(eta_bulk),
                              (ctypes.c_double * 3)(*force),
                              ctypes.c_double(mobility))

    def relaxation(self, nrelax):
        self.L.refdrv_relaxation(nrelax)

    def fe_symm(self, a, b, kappa):
        self.L.refdrv_fe_symm(ctypes.c_double(a), ctypes.c_double(b),
                              ctypes.c_double(kappa))

    def rtau(self):
        out = np.zeros(self.nvel)
        self.L.refdrv_rtau(out.ctypes.data_as(c_double_p))
        return out

    def equilib_set(self, index, rho, u):
        self.L.refdrv_equilib_set(index, ctypes.c_double(rho),
                                  (ctypes.c_double * 3)(*u))

    def collide(self):
        self.L.refdrv_collide()

    def halo(self, scheme=-1):
        self.L.refdrv_halo(scheme)

    def propagate(self):
        self.L.refdrv_propagate()

    def phi_from_lb(self):
        self.L.refdrv_phi_from_lb()

    def phi_halo(self):
        self.L.refdrv_phi_halo()

    def grad_compute(self):
        self.L.refdrv_grad_compute()

    def noise(self, on, kt, reseed=0):
        """isothermal_fluctuations on/off at temperature kt (reseed > 0: noise_init
        again with that master seed)."""
        assert self.L.refdrv_noise(int(on), float(kt), int(reseed)) == 0

    def le_apply(self, t):
        """lb_le_apply_boundary_conditions at time step t."""
        assert self.L.refdrv_le_apply(int(t)) == 0

    def le_plane_location(self, n):
        return self.L.refdrv_le_plane_location(int(n))

    def time_set(self, t):
        """The next refdrv_step starts with physics_control_next_step from here."""
        self.L.refdrv_time_set(int(t))

    def open_bc(self, flow_dir, u0=(0.0, 0.0, 0.0), rho0=1.0):
        """lb_bc_open: inflow and outflow of type rhou along flow_dir (walls across it first)."""
        self.L.refdrv_open_bc.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.c_double]
        assert self.L.refdrv_open_bc(int(flow_dir), (ctypes.c_double * 3)(*u0), float(rho0)) == 0

    def open_apply(self):
        """inflow/outflow update + impose (src/ludwig.c:820-831)."""
        assert self.L.refdrv_open_apply() == 0

    def time(self):
        """physics_control_time: what the displacement of the order parameter is taken at."""
        self.L.refdrv_time.restype = ctypes.c_double
        return self.L.refdrv_time()

    def wall(self, isboundary, ubot=(0.0, 0.0, 0.0), utop=(0.0, 0.0, 0.0)):
        """boundary_walls / boundary_speed_bottom,_top: wall_create + wall_commit."""
        rc = self.L.refdrv_wall((ctypes.c_int * 3)(*isboundary), (ctypes.c_double * 3)(*ubot),
                                (ctypes.c_double * 3)(*utop))
        assert rc == 0
        return self.L.refdrv_wall_nlink()

    def wall_slip(self, isboundary, sbot, stop, force_active=0):
        """boundary_walls with boundary_walls_slip_fraction_bot/_top (/_active)."""
        rc = self.L.refdrv_wall_slip((ctypes.c_int * 3)(*isboundary), (ctypes.c_double * 3)(*sbot),
                                     (ctypes.c_double * 3)(*stop), int(force_active))
        assert rc == 0
        return self.L.refdrv_wall_nlink()

    def wall_bbl(self):
        assert self.L.refdrv_wall_bbl() == 0

    def wall_momentum(self):
        out = (ctypes.c_double * 3)()
        self.L.refdrv_wall_momentum(out)
        return np.array(out[:])

    def ghost_modes(self, on):
        assert self.L.refdrv_ghost_modes(int(on)) == 0

    def noise_state(self):
        return self._arr(self.L.refdrv_noise_state(), (self.nsite, 4))

    def noise_reap_n(self, index, nmax):
        out = np.zeros(nmax)
        self.L.refdrv_noise_reap_n(int(index), int(nmax),
                                   out.ctypes.data_as(c_double_p))
        return out

    def interior(self, a):
        """a[nsite, ...] of the reference -> the interior block [nx, ny, nz, ...]."""
        h = self.nhalo
        b = a.reshape(self.nall + a.shape[1:])
        return b[h:-h, h:-h, h:-h]

    def inner(self, a, width=1):
        """the interior plus `width` halo layers: what a halo-1 array holds."""
        h = self.nhalo - width
        b = a.reshape(self.nall + a.shape[1:])
        if h == 0:
            return b
        return b[h:-h, h:-h, h:-h]

    def phi_force(self):
        assert self.L.refdrv_phi_force() == 0

    def cahn_hilliard(self):
        assert self.L.refdrv_cahn_hilliard() == 0

    def u_halo(self):
        self.L.refdrv_u_halo()

    def grad_mu(self, gm):
        self.L.refdrv_grad_mu((ctypes.c_double * 3)(*gm))

    def visc_arrhenius(self, eta_plus, eta_minus, phistar):
        """viscosity_model arrhenius (needs the order parameter: fd=... or ndist=2)."""
        self.L.refdrv_visc_arrhenius.argtypes = [ctypes.c_double] * 3
        assert self.L.refdrv_visc_arrhenius(eta_plus, eta_minus, phistar) == 0

    def visc_update(self):
        self.L.refdrv_visc_update()

    def eta(self):
        return self._arr(self.L.refdrv_eta(), (self.nsite,))

    def grad_select(self, npt):
        assert self.L.refdrv_grad_select(int(npt)) == 0

    def hydro_f_zero(self, f=(0.0, 0.0, 0.0)):
        self.L.refdrv_hydro_f_zero((ctypes.c_double * 3)(*f))

    def hydro_u_zero(self, u=(0.0, 0.0, 0.0)):
        self.L.refdrv_hydro_u_zero((ctypes.c_double * 3)(*u))

    def step(self, nsteps):
        """Runs nsteps of the src/ludwig.c:530-862 loop; returns seconds."""
        return self.L.refdrv_step(nsteps)
