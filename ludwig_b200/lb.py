"""ctypes binding of libb200lb.so and a thin mirror of the reference call surface.

Method names follow the reference (zazu29/ludwig v0.20.1) so tests read like its own:
lb_collide (src/collision.c:144), lb_halo / lb_halo_swap (src/model.c:553,570),
lb_propagation (src/propagation.c:43), lb_memcpy (src/model.c:228),
hydro_f_zero / hydro_u_zero (src/hydro.c:306,279), phi_lb_to_field
(src/phi_lb_coupler.c:39), field_halo (src/field.c:434), field_grad_compute ->
grad_3d_7pt_fluid_d2 (src/gradient_3d_7pt_fluid.c:75), phi_force_calculation
(src/phi_force.c:74), phi_cahn_hilliard (src/phi_cahn_hilliard.c:206).
"""

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBPATH = os.path.join(_HERE, "libb200lb.so")

HOST_AOS, HOST_SOA = 0, 1


def HOST_AOSOA(vvl):
    """-DADDR_AOSOA -DVVL=vvl builds of the reference (src/memory.h:260-284)."""
    return 256 + int(vvl)


RELAX_M10, RELAX_BGK, RELAX_TRT = 0, 1, 2
HALO_FULL, HALO_REDUCED = 0, 2
FIELD_RHO, FIELD_U, FIELD_FORCE, FIELD_PHI, FIELD_GRAD, FIELD_DELSQ, FIELD_ETA = range(7)
FORCE_NONE, FORCE_STRESS_DIVERGENCE, FORCE_PHI_GRADMU = 1, 2, 3      # fe_force_method_enum_t
_FIELD_NF = {FIELD_RHO: 1, FIELD_U: 3, FIELD_FORCE: 3, FIELD_PHI: 1, FIELD_GRAD: 3,
             FIELD_DELSQ: 1, FIELD_ETA: 1}

# every symbol include/b200lb.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "b200lb_last_error", "b200lb_device_count", "b200lb_version", "b200lb_create",
    "b200lb_destroy", "b200lb_nsite", "b200lb_set_option", "b200lb_get_option",
    "b200lb_f_upload", "b200lb_f_upload_range", "b200lb_f_download", "b200lb_field_upload",
    "b200lb_field_download", "b200lb_set_param", "b200lb_hydro_f_zero",
    "b200lb_hydro_u_zero", "b200lb_hydro_rho0", "b200lb_hydro_correct_momentum", "b200lb_collide", "b200lb_halo", "b200lb_halo_swap",
    "b200lb_propagate", "b200lb_phi_from_g", "b200lb_phi_halo", "b200lb_grad7", "b200lb_grad27", "b200lb_grad27_solid",
    "b200lb_noise_init", "b200lb_noise_present_set", "b200lb_noise_state_upload",
    "b200lb_noise_state_download",
    "b200lb_wall_commit", "b200lb_wall_slip_set", "b200lb_wall_bbl", "b200lb_wall_momentum", "b200lb_wall_momentum_add",
    "b200lb_le_commit", "b200lb_le_apply", "b200lb_le_plane_location", "b200lb_le_time_set",
    "b200lb_le_field_time_set", "b200lb_open_bc_commit", "b200lb_open_bc_apply",
    "b200lb_ch_create", "b200lb_ch_grad_mu_set", "b200lb_phi_force", "b200lb_cahn_hilliard",
    "b200lb_hydro_u_halo", "b200lb_visc_arrhenius_commit", "b200lb_visc_update", "b200lb_visc_model_set",
    "b200lb_field_stats", "b200lb_field_stats_noalloc", "b200lb_io_record_doubles", "b200lb_io_pack", "b200lb_io_unpack",
    "b200lb_step", "b200lb_stats", "b200lb_step_diag", "b200lb_sync", "b200lb_nccl_unique_id",
    "b200lb_comm_init", "b200lb_p2p_export", "b200lb_p2p_import", "b200lb_timer_start", "b200lb_timer_stop",
    "b200lb_launch_count", "b200lb_kernel_timer_start", "b200lb_kernel_timer_stop", "b200lb_host_alloc", "b200lb_host_free",
]


class B200LBError(RuntimeError):
    pass


class Desc(ctypes.Structure):
    _fields_ = [
        ("ndim", ctypes.c_int), ("nvel", ctypes.c_int), ("ndist", ctypes.c_int),
        ("nlocal", ctypes.c_int * 3), ("nhalo", ctypes.c_int),
        ("cartsz", ctypes.c_int * 3), ("cartcoords", ctypes.c_int * 3),
        ("periodic", ctypes.c_int * 3), ("nbr_lo", ctypes.c_int * 3),
        ("nbr_hi", ctypes.c_int * 3), ("rank", ctypes.c_int),
        ("nranks", ctypes.c_int), ("device", ctypes.c_int), ("zalign", ctypes.c_int),
        ("lattices", ctypes.c_int),
    ]


class Param(ctypes.Structure):
    _fields_ = [
        ("rho0", ctypes.c_double), ("eta_shear", ctypes.c_double),
        ("eta_bulk", ctypes.c_double), ("force_global", ctypes.c_double * 3),
        ("nrelax", ctypes.c_int), ("mobility", ctypes.c_double),
        ("a", ctypes.c_double), ("b", ctypes.c_double), ("kappa", ctypes.c_double),
        ("kt", ctypes.c_double), ("ghost_off", ctypes.c_int),
    ]


def library_path():
    return _LIBPATH


def build(verbose=False):
    """Compile libb200lb.so for sm_100a (nvcc cross-compiles without a GPU)."""
    ncpu = str(min(os.cpu_count() or 4, 8))
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-s", "-j", ncpu, "-C", os.path.join(_HERE, "csrc")],
                          stdout=out)


_lib = None


def load():
    """Load the C-ABI library. Fails loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIBPATH):
        raise B200LBError("%s is missing: run `python -c 'import __graft_entry__ as g; "
                          "g.build()'` (there is no CPU fallback)" % _LIBPATH)
    L = ctypes.CDLL(_LIBPATH)
    L.b200lb_last_error.restype = ctypes.c_char_p
    L.b200lb_version.restype = ctypes.c_char_p
    L.b200lb_nsite.restype = ctypes.c_long
    L.b200lb_get_option.restype = ctypes.c_long
    L.b200lb_launch_count.restype = ctypes.c_long
    vp = ctypes.c_void_p
    L.b200lb_create.argtypes = [ctypes.POINTER(Desc), ctypes.POINTER(vp)]
    L.b200lb_set_option.argtypes = [vp, ctypes.c_char_p, ctypes.c_long]
    L.b200lb_get_option.argtypes = [vp, ctypes.c_char_p]
    L.b200lb_set_param.argtypes = [vp, ctypes.POINTER(Param)]
    for name in ("b200lb_f_upload", "b200lb_f_download"):
        getattr(L, name).argtypes = [vp, vp, ctypes.c_int]
    for name in ("b200lb_field_upload", "b200lb_field_download"):
        getattr(L, name).argtypes = [vp, ctypes.c_int, vp, ctypes.c_int]
    for name in ("b200lb_destroy", "b200lb_collide", "b200lb_halo", "b200lb_propagate",
                 "b200lb_phi_from_g", "b200lb_phi_halo", "b200lb_grad7", "b200lb_grad27", "b200lb_grad27_solid", "b200lb_sync",
                 "b200lb_hydro_correct_momentum", "b200lb_timer_start", "b200lb_nsite", "b200lb_launch_count"):
        getattr(L, name).argtypes = [vp]
    L.b200lb_f_upload_range.argtypes = [vp, vp, ctypes.c_long, ctypes.c_long]
    L.b200lb_halo_swap.argtypes = [vp, ctypes.c_int]
    L.b200lb_wall_commit.argtypes = [vp, vp, vp, vp]
    L.b200lb_wall_slip_set.argtypes = [vp, vp, vp, ctypes.c_int]
    L.b200lb_wall_bbl.argtypes = [vp]
    L.b200lb_wall_momentum.argtypes = [vp, vp]
    L.b200lb_wall_momentum_add.argtypes = [vp, vp]
    L.b200lb_le_commit.argtypes = [vp, ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    L.b200lb_le_apply.argtypes = [vp, ctypes.c_double]
    L.b200lb_le_plane_location.argtypes = [vp, ctypes.c_int]
    L.b200lb_le_time_set.argtypes = [vp, ctypes.c_int]
    L.b200lb_le_field_time_set.argtypes = [vp, ctypes.c_double]
    L.b200lb_open_bc_commit.argtypes = [vp, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.c_double]
    L.b200lb_open_bc_apply.argtypes = [vp]
    L.b200lb_ch_create.argtypes = [vp, ctypes.c_int, ctypes.c_int]
    L.b200lb_ch_grad_mu_set.argtypes = [vp, vp]
    for name in ("b200lb_phi_force", "b200lb_cahn_hilliard", "b200lb_hydro_u_halo"):
        getattr(L, name).argtypes = [vp]
    L.b200lb_visc_arrhenius_commit.argtypes = [vp, ctypes.c_double, ctypes.c_double, ctypes.c_double]
    L.b200lb_visc_update.argtypes = [vp]
    L.b200lb_visc_model_set.argtypes = [vp, ctypes.c_int]
    L.b200lb_field_stats.argtypes = [vp, ctypes.c_int, vp]
    L.b200lb_field_stats_noalloc.argtypes = [vp, ctypes.c_int, vp]
    L.b200lb_io_record_doubles.argtypes = [vp]
    L.b200lb_io_record_doubles.restype = ctypes.c_long
    L.b200lb_io_pack.argtypes = [vp, vp]
    L.b200lb_io_unpack.argtypes = [vp, vp]
    L.b200lb_noise_init.argtypes = [vp, ctypes.c_int]
    L.b200lb_noise_present_set.argtypes = [vp, ctypes.c_int]
    L.b200lb_noise_state_upload.argtypes = [vp, vp, ctypes.c_int]
    L.b200lb_noise_state_download.argtypes = [vp, vp, ctypes.c_int]
    L.b200lb_hydro_rho0.argtypes = [vp, ctypes.c_double]
    L.b200lb_step.argtypes = [vp, ctypes.c_int]
    L.b200lb_hydro_f_zero.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
    L.b200lb_hydro_u_zero.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
    L.b200lb_stats.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
    L.b200lb_step_diag.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
    L.b200lb_timer_stop.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    L.b200lb_kernel_timer_start.argtypes = [vp, ctypes.c_int]
    L.b200lb_kernel_timer_stop.argtypes = [vp, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_long)]
    L.b200lb_nccl_unique_id.argtypes = [ctypes.c_char_p]
    L.b200lb_comm_init.argtypes = [vp, ctypes.c_char_p]
    L.b200lb_p2p_export.argtypes = [vp, ctypes.c_char_p]
    L.b200lb_p2p_import.argtypes = [vp, ctypes.c_char_p, ctypes.c_int]
    L.b200lb_host_alloc.argtypes = [ctypes.POINTER(vp), ctypes.c_long]
    L.b200lb_host_free.argtypes = [vp]
    _lib = L
    return L


def _ptr(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return ctypes.c_void_p(a.ctypes.data)


def pinned_empty(shape):
    """numpy array over page-locked host memory (b200lb_host_alloc = cudaMallocHost),
    e.g. for lb->f: lb_memcpy then runs at full PCIe rate. The block lives until
    pinned_free(array) or process exit."""
    L = load()
    n = int(np.prod(shape))
    p = ctypes.c_void_p()
    if L.b200lb_host_alloc(ctypes.byref(p), n * 8) != 0:
        raise B200LBError(L.b200lb_last_error().decode())
    buf = (ctypes.c_double * n).from_address(p.value)
    arr = np.frombuffer(buf, dtype=np.float64).reshape(shape)
    _PINNED[arr.ctypes.data] = p
    return arr


def pinned_free(arr):
    p = _PINNED.pop(arr.ctypes.data, None)
    if p is not None:
        load().b200lb_host_free(p)


_PINNED = {}


class Lattice:
    """One rank's lattice on one B200: the lb_t + hydro_t (+ phi, grad) of the
    reference, device-resident in the library's private layout."""

    def __init__(self, nvel, nlocal, ndist=1, decomp=None, device=-1, zalign=0,
                 periodic=(1, 1, 1), lattices=2):
        self.L = load()
        self.nvel, self.ndist = nvel, ndist
        self.ndim = 2 if nvel == 9 else 3
        self.nlocal = tuple(int(x) for x in nlocal)
        d = Desc()
        d.ndim, d.nvel, d.ndist, d.nhalo = self.ndim, nvel, ndist, 1
        d.nlocal[:] = self.nlocal
        if decomp is None:
            d.cartsz[:] = (1, 1, 1)
            d.cartcoords[:] = (0, 0, 0)
            d.periodic[:] = tuple(int(bool(x)) for x in periodic)
            d.nbr_lo[:] = tuple(0 if x else -1 for x in periodic)
            d.nbr_hi[:] = tuple(0 if x else -1 for x in periodic)
            d.rank, d.nranks = 0, 1
        else:
            d.cartsz[:] = decomp.cartsz
            d.cartcoords[:] = decomp.coords
            d.periodic[:] = decomp.periodic
            d.nbr_lo[:] = decomp.nbr_lo
            d.nbr_hi[:] = decomp.nbr_hi
            d.rank, d.nranks = decomp.rank, decomp.nranks
        d.device, d.zalign = device, zalign
        d.lattices = int(lattices)
        self.h = ctypes.c_void_p()
        self._ck(self.L.b200lb_create(ctypes.byref(d), ctypes.byref(self.h)))
        self.nall = tuple(n + 2 for n in self.nlocal)
        self.nsite = int(self.L.b200lb_nsite(self.h))

    def _ck(self, rc):
        if rc != 0:
            raise B200LBError(self.L.b200lb_last_error().decode())

    def close(self):
        if self.h:
            self.L.b200lb_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- options / parameters -------------------------------------------------
    def set_option(self, name, value):
        self._ck(self.L.b200lb_set_option(self.h, name.encode(), int(value)))

    def get_option(self, name):
        return int(self.L.b200lb_get_option(self.h, name.encode()))

    def set_param(self, rho0=1.0, eta_shear=0.1, eta_bulk=0.1, force=(0.0, 0.0, 0.0),
                  nrelax=RELAX_M10, mobility=0.0, a=0.0, b=0.0, kappa=0.0, kt=0.0,
                  ghost_off=0):
        p = Param()
        p.kt = kt
        p.ghost_off = ghost_off
        p.rho0, p.eta_shear, p.eta_bulk = rho0, eta_shear, eta_bulk
        p.force_global[:] = force
        p.nrelax, p.mobility, p.a, p.b, p.kappa = nrelax, mobility, a, b, kappa
        self._ck(self.L.b200lb_set_param(self.h, ctypes.byref(p)))

    # -- lb_memcpy and friends --------------------------------------------------
    def alloc_f(self):
        return np.zeros((self.nsite, self.ndist, self.nvel))

    def lb_memcpy_h2d(self, f, layout=HOST_AOS):
        self._ck(self.L.b200lb_f_upload(self.h, _ptr(f), layout))

    def lb_memcpy_h2d_range(self, f_part, first):
        """lb_io_read piecewise: f_part [count, ndist, nvel] holds host sites first..first+count."""
        self._ck(self.L.b200lb_f_upload_range(self.h, _ptr(f_part), int(first), int(f_part.shape[0])))

    def lb_memcpy_d2h(self, f=None, layout=HOST_AOS):
        if f is None:
            f = (self.alloc_f() if layout == HOST_AOS
                 else np.zeros((self.ndist, self.nvel, self.nsite)))
        self._ck(self.L.b200lb_f_download(self.h, _ptr(f), layout))
        return f

    def field_h2d(self, field, a, layout=HOST_AOS):
        self._ck(self.L.b200lb_field_upload(self.h, field, _ptr(a), layout))

    def field_d2h(self, field, layout=HOST_AOS):
        nf = _FIELD_NF[field]
        shape = (self.nsite,) if nf == 1 else ((self.nsite, nf) if layout == HOST_AOS
                                              else (nf, self.nsite))
        a = np.zeros(shape)
        self._ck(self.L.b200lb_field_download(self.h, field, _ptr(a), layout))
        return a

    # -- the reference call surface ---------------------------------------------
    def hydro_f_zero(self, f=(0.0, 0.0, 0.0)):
        self._ck(self.L.b200lb_hydro_f_zero(self.h, (ctypes.c_double * 3)(*f)))

    def hydro_u_zero(self, u=(0.0, 0.0, 0.0)):
        self._ck(self.L.b200lb_hydro_u_zero(self.h, (ctypes.c_double * 3)(*u)))

    def hydro_rho0(self, rho0):
        self._ck(self.L.b200lb_hydro_rho0(self.h, rho0))

    def hydro_correct_momentum(self):
        self._ck(self.L.b200lb_hydro_correct_momentum(self.h))

    def lb_collide(self):
        self._ck(self.L.b200lb_collide(self.h))

    def lb_halo(self):
        self._ck(self.L.b200lb_halo(self.h))

    def lb_halo_swap(self, scheme):
        self._ck(self.L.b200lb_halo_swap(self.h, scheme))

    def lb_propagation(self):
        self._ck(self.L.b200lb_propagate(self.h))

    def phi_lb_to_field(self):
        self._ck(self.L.b200lb_phi_from_g(self.h))

    def field_halo(self):
        self._ck(self.L.b200lb_phi_halo(self.h))

    def field_grad_set(self, npt):
        """Which d2 stencil field_grad_compute() and step() run: 7 (grad_3d_7pt_fluid_d2)
        or 27 (grad_3d_27pt_fluid_d2); the reference's field_grad_set plug point."""
        self.set_option("gradient", int(npt))

    def field_grad_compute(self):
        if self.get_option("gradient") == 28:
            self._ck(self.L.b200lb_grad27_solid(self.h))
        elif self.get_option("gradient") == 27:
            self._ck(self.L.b200lb_grad27(self.h))
        else:
            self._ck(self.L.b200lb_grad7(self.h))

    # -- the order parameter as a finite-difference field ("free_energy symmetric") ----
    def phi_ch_create(self, order=1, method=FORCE_STRESS_DIVERGENCE, grad_mu=None):
        """phi_ch_create + pth_create + advection_order_set: order = fd_advection_scheme_order,
        method = fe_force_method; grad_mu = the external chemical potential gradient."""
        self._ck(self.L.b200lb_ch_create(self.h, int(order), int(method)))
        if grad_mu is not None:
            self._ck(self.L.b200lb_ch_grad_mu_set(self.h, (ctypes.c_double * 3)(*grad_mu)))

    def phi_force_calculation(self):
        self._ck(self.L.b200lb_phi_force(self.h))

    def phi_cahn_hilliard(self):
        self._ck(self.L.b200lb_cahn_hilliard(self.h))

    def hydro_u_halo(self):
        self._ck(self.L.b200lb_hydro_u_halo(self.h))

    # -- viscosity model (visc_t, src/visc.h; src/visc_arrhenius.c) -------------------
    def visc_arrhenius_commit(self, eta_plus, eta_minus, phistar=1.0):
        """viscosity_model arrhenius: lb_collide reads a per-site viscosity from now on."""
        self._ck(self.L.b200lb_visc_arrhenius_commit(self.h, float(eta_plus), float(eta_minus),
                                                     float(phistar)))

    def visc_update(self):
        self._ck(self.L.b200lb_visc_update(self.h))

    def visc_model_set(self, on=1):
        """lb_collide(..., visc != NULL): take eta from FIELD_ETA (uploaded by the caller)."""
        self._ck(self.L.b200lb_visc_model_set(self.h, int(on)))

    # -- flat walls (wall_t, src/wall.h) ---------------------------------------------
    def wall_commit(self, isboundary, ubot=(0.0, 0.0, 0.0), utop=(0.0, 0.0, 0.0)):
        self._ck(self.L.b200lb_wall_commit(self.h, (ctypes.c_int * 3)(*isboundary),
                                           (ctypes.c_double * 3)(*ubot),
                                           (ctypes.c_double * 3)(*utop)))

    def wall_slip_set(self, sbot, stop, force_active=0):
        self._ck(self.L.b200lb_wall_slip_set(self.h, (ctypes.c_double * 3)(*sbot),
                                             (ctypes.c_double * 3)(*stop), int(force_active)))

    def open_bc_commit(self, flow_dir, u0=(0.0, 0.0, 0.0), rho0=1.0):
        """lb_bc_open: inflow (rhou, velocity u0) and outflow (rhou, density rho0) along flow_dir."""
        self._ck(self.L.b200lb_open_bc_commit(self.h, int(flow_dir), (ctypes.c_double * 3)(*u0), float(rho0)))

    def open_bc_apply(self):
        """inflow/outflow update + impose: between lb_halo() and lb_propagation()."""
        self._ck(self.L.b200lb_open_bc_apply(self.h))

    def wall_bbl(self):
        self._ck(self.L.b200lb_wall_bbl(self.h))

    def wall_momentum(self):
        out = (ctypes.c_double * 3)()
        self._ck(self.L.b200lb_wall_momentum(self.h, out))
        return np.array(out[:])

    # -- Lees-Edwards planes (lees_edw_create src/leesedwards.c:80,
    #    lb_le_apply_boundary_conditions src/model_le.c:57) --------------------------

    def le_commit(self, nplanes, uy, oscillatory=False, period=0, nt0=0):
        """lees_edw_options_t {nplanes, type, period, nt0, uy} (src/lees_edwards_options.h)."""
        self._ck(self.L.b200lb_le_commit(self.h, int(nplanes), float(uy), int(bool(oscillatory)),
                                         int(period), int(nt0)))

    def le_apply_boundary_conditions(self, t):
        """Between collide() and halo(), t = physics time (the step number)."""
        self._ck(self.L.b200lb_le_apply(self.h, float(t)))

    def le_plane_location(self, n):
        ic = self.L.b200lb_le_plane_location(self.h, int(n))
        if ic < 0:
            raise RuntimeError(self.L.b200lb_last_error().decode())
        return ic

    def le_time_set(self, nstep):
        self._ck(self.L.b200lb_le_time_set(self.h, int(nstep)))

    def le_field_time_set(self, t):
        """physics_control_time for the field_leesedwards of the next field_grad_compute."""
        self._ck(self.L.b200lb_le_field_time_set(self.h, float(t)))

    # -- statistics and file records without the full-field D2H ---------------------
    def field_stats(self, field):
        """[nf, 4]: sum, sum of squares, min, max of each component over interior sites."""
        nf = _FIELD_NF[field]
        out = np.zeros((nf, 4))
        self._ck(self.L.b200lb_field_stats(self.h, field, _ptr(out)))
        return out

    def field_stats_noalloc(self, field):
        """field_stats through the entry point that allocates nothing (static device scratch,
        at most 65536 blocks of 128 x 2 sites; no counterpart in the reference)."""
        nf = _FIELD_NF[field]
        out = np.zeros((nf, 4))
        self._ck(self.L.b200lb_field_stats_noalloc(self.h, field, _ptr(out)))
        return out

    def lb_io_aggr_pack(self, buf=None):
        """Binary records of lb_io_write: [interior site (kc fastest), n, p]."""
        if buf is None:
            buf = np.zeros((int(np.prod(self.nlocal)), self.ndist, self.nvel))
        assert buf.size == int(np.prod(self.nlocal)) * self.L.b200lb_io_record_doubles(self.h)
        self._ck(self.L.b200lb_io_pack(self.h, _ptr(buf)))
        return buf

    def lb_io_aggr_unpack(self, buf):
        assert buf.size == int(np.prod(self.nlocal)) * self.L.b200lb_io_record_doubles(self.h)
        self._ck(self.L.b200lb_io_unpack(self.h, _ptr(buf)))

    # -- isothermal fluctuations (noise_t, src/noise.h) ------------------------------
    def noise_init(self, master_seed=0):
        self._ck(self.L.b200lb_noise_init(self.h, int(master_seed)))

    def noise_present_set(self, on=1):
        self._ck(self.L.b200lb_noise_present_set(self.h, int(on)))

    def noise_memcpy_h2d(self, state, layout=HOST_AOS):
        assert state.dtype == np.uint32 and state.flags["C_CONTIGUOUS"]
        self._ck(self.L.b200lb_noise_state_upload(self.h, state.ctypes.data_as(ctypes.c_void_p),
                                                  layout))

    def noise_memcpy_d2h(self, layout=HOST_AOS):
        shape = (self.nsite, 4) if layout == HOST_AOS else (4, self.nsite)
        state = np.zeros(shape, dtype=np.uint32)
        self._ck(self.L.b200lb_noise_state_download(self.h, state.ctypes.data_as(ctypes.c_void_p),
                                                    layout))
        return state

    def step(self, nsteps=1):
        self._ck(self.L.b200lb_step(self.h, nsteps))

    def stats(self):
        out = (ctypes.c_double * 8)()
        self._ck(self.L.b200lb_stats(self.h, out))
        return np.array(out[:])

    def step_diag(self):
        out = (ctypes.c_double * 8)()
        self._ck(self.L.b200lb_step_diag(self.h, out))
        return np.array(out[:])

    def sync(self):
        self._ck(self.L.b200lb_sync(self.h))

    # -- timing / comm ------------------------------------------------------------
    def timer_start(self):
        self._ck(self.L.b200lb_timer_start(self.h))

    def timer_stop(self):
        ms = ctypes.c_float()
        self._ck(self.L.b200lb_timer_stop(self.h, ctypes.byref(ms)))
        return float(ms.value)

    def launch_count(self):
        return int(self.L.b200lb_launch_count(self.h))

    def kernel_timer_start(self, max_launches=4096):
        """TIMER_COLLIDE_KERNEL: CUDA events around every sweep launch from here on."""
        self._ck(self.L.b200lb_kernel_timer_start(self.h, int(max_launches)))

    def kernel_timer_stop(self):
        """(ms summed over the sweep launches since kernel_timer_start, how many)."""
        ms, n = ctypes.c_double(), ctypes.c_long()
        self._ck(self.L.b200lb_kernel_timer_stop(self.h, ctypes.byref(ms), ctypes.byref(n)))
        return float(ms.value), int(n.value)

    @staticmethod
    def nccl_unique_id():
        L = load()
        buf = ctypes.create_string_buffer(128)
        if L.b200lb_nccl_unique_id(buf) != 0:
            raise B200LBError(L.b200lb_last_error().decode())
        return buf.raw

    def p2p_export(self):
        buf = ctypes.create_string_buffer(192)
        self._ck(self.L.b200lb_p2p_export(self.h, buf))
        return buf.raw

    def p2p_import(self, blobs, nranks):
        """blobs: the concatenated p2p_export() of all ranks in rank order (or None for
        a single rank in loopback mode)."""
        arg = ctypes.create_string_buffer(blobs, len(blobs)) if blobs else None
        self._ck(self.L.b200lb_p2p_import(self.h, arg, nranks))

    def comm_init(self, uid):
        self._ck(self.L.b200lb_comm_init(self.h, ctypes.create_string_buffer(uid, 128)))
