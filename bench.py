#!/usr/bin/env python
"""bench.py -- D3Q19 fp64 MLUPS of the Ludwig LB time step on B200, with the HBM
bytes-per-site roofline and the reference's CPU build timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W]           this implementation
    python bench.py --impl reference [--steps K] [--warmup W]     Ludwig's own CPU code

One "step" is one pass of the reference's time-step hot path (src/ludwig.c:530-862
restricted to it): hydro_f_zero, hydro_u_zero, lb_collide, lb_halo, lb_propagation
over one synthetic lattice. N = 1 runs BASELINE.json configs[1] (D3Q19 single-fluid MRT,
256^3 periodic) and, in "configs", one short measurement of every other named shape
(64^3, binary 128^3, D3Q27 256^3, D2Q9 4096^2). N > 1 is weak scaling: every GPU keeps a
256^3 block of a lattice decomposed in 3-d over the GPUs (2x1x1, 2x2x1, 2x2x2, the
reference's MPI_Dims_create grids, src/coords.c:173-194), halo planes pushed into the
neighbours' memory over NVLink from inside the sweep.

Prints ONE JSON line (rank 0). Keys are documented in DESIGN.md section "Measurement".
"""

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_ALG = {19: 304, 27: 432, 9: 144}      # bytes per lattice-site update, SURVEY.md 8(d): 2*Q*8


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nvel", type=int, default=19)
    ap.add_argument("--local", "--lsize", dest="local", type=int, nargs=3, default=None,
                    help="local lattice per GPU (default 256 256 256; D2Q9 4096 4096 1)")
    ap.add_argument("--global-size", type=int, nargs=3, default=None,
                    help="strong scaling: this global lattice divided over the grid")
    ap.add_argument("--ndist", type=int, default=1)
    ap.add_argument("--fused", type=int, default=1)
    ap.add_argument("--vec", type=int, default=0, help="sites per thread (0 = library default)")
    ap.add_argument("--block", type=int, default=128)
    ap.add_argument("--zalign", type=int, default=0)
    ap.add_argument("--halo", default="reduced", choices=["full", "reduced"])
    ap.add_argument("--hydro-write", type=int, default=2, choices=[0, 1, 2],
                    help="2 (library default): rho,u stored by the collision only when asked "
                         "for since the previous one; 1: every step, as the reference")
    ap.add_argument("--lattices", type=int, default=2, choices=[1, 2],
                    help="1: single-lattice streaming (AA pattern), half the memory")
    ap.add_argument("--overlap", type=int, default=1)
    ap.add_argument("--direct", type=int, default=1)
    ap.add_argument("--p2p", type=int, default=1, help="peer-memory halo (CUDA IPC) instead of NCCL messages")
    ap.add_argument("--threads", type=int, default=0, help="threads per block of the sweep (0 = library default)")
    ap.add_argument("--zpack", type=int, default=1,
                    help="peer-memory halo: z faces through packed planes (1, library default) or row-end halo words (0)")
    ap.add_argument("--loopback", type=int, default=0, help="single GPU: run the halo exchange machinery with the GPU as its own neighbour")
    ap.add_argument("--grid", type=int, nargs=3, default=None, help="process grid (default 2x1x1, 2x2x1, 2x2x2)")
    ap.add_argument("--gradient", type=int, default=7, choices=[7, 27],
                    help="binary fluid: order-parameter gradient stencil")
    ap.add_argument("--kt", type=float, default=0.0,
                    help="> 0: isothermal fluctuations at this temperature (not the headline workload)")
    ap.add_argument("--le", type=int, default=0,
                    help="Lees-Edwards planes in total, uy 0.01 (not the headline workload)")
    ap.add_argument("--fd", type=int, default=0, choices=[0, 1, 2, 3],
                    help="free_energy symmetric: the order parameter as a finite-difference field, "
                         "advection scheme of this order (phi_force_calculation + phi_cahn_hilliard)")
    ap.add_argument("--walls", type=int, nargs=3, default=None,
                    help="flat no-slip walls in these directions, e.g. 0 1 1 (not the headline workload)")
    ap.add_argument("--init", default="auto", choices=["auto", "seeded", "tiled"],
                    help="seeded: every site from the global-id generator; tiled: a seeded block repeated")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the other named shapes (N = 1)")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-variants", action="store_true",
                    help="skip the hydro_every_step variant (profiling: one kernel variant per run)")
    ap.add_argument("--ref-size", type=int, nargs=3, default=None)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML during the timed region."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.stop_flag = threading.Event()
        self.sm, self.reasons, self.sm_max = [], set(), None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self.stop_flag.is_set():
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.dev, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def result(self):
        self.stop_flag.set()
        if self.ok and self.is_alive():
            self.join(timeout=2)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def workload_name(nvel, ndist, local, grid, extra=""):
    ndim = 2 if nvel == 9 else 3
    return ("D%dQ%d %s MRT %dx%dx%d periodic per GPU, grid %s%s"
            % (ndim, nvel, "binary symmetric_lb" if ndist == 2 else "single-fluid",
               local[0], local[1], local[2], "x".join(str(g) for g in grid), extra))


def traffic_key(nvel, ndist, local, hydro_store):
    return "q%d-n%d-%dx%dx%d-%s" % (nvel, ndist, local[0], local[1], local[2],
                                    "hydro_store" if hydro_store else "hydro_on_demand")


def ncu_traffic(nvel, ndist, local, hydro_store):
    """DRAM bytes per launch of the dominant kernel from the committed ncu captures
    (profiles/r2_ncu_traffic.json, one entry per configuration), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")) as fh:
            return json.load(fh).get(traffic_key(nvel, ndist, local, hydro_store), {}).get(
                "dram_bytes_per_launch")
    except Exception:
        return None


def grid_for(n):
    # MPI_Dims_create (src/coords.c:173-194): 2 -> 2x1x1, 4 -> 2x2x1, 8 -> 2x2x2
    return {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}.get(n)


def host_threads():
    """All host cores for the reference's OpenMP kernels, whatever the launcher put in
    OMP_NUM_THREADS (torch.distributed.run sets it to 1). libgomp reads the variable when it
    is loaded, which may have happened already: also set it through the runtime."""
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0)) or n
    except Exception:
        pass
    os.environ["OMP_NUM_THREADS"] = str(n)
    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(int(n))
    except Exception:
        pass
    return n


BINARY_KW = dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75)
BINARY_FE = dict(a=-0.00625, b=0.00625, kappa=0.004)


def cpu_reference(nvel, size, seconds, nsteps_fixed=None, warmup=1, ndist=1, gradient=7):
    """Times the UNMODIFIED reference (oracle/_ref, OpenMP over the host cores) on a
    bounded sample of the named workload -- single fluid, or the binary fluid with its
    gradient stencil; falls back to the oracle port when the prebuilt reference is absent."""
    from oracle import pyref, pyoracle
    from ludwig_b200 import models_host, synthetic
    cores = host_threads()
    cv, wv = models_host.cv(nvel), models_host.wv(nvel)
    ndim = 2 if nvel == 9 else 3
    f0 = synthetic.tiled_state(cv, wv, size, ndist=ndist, ndim=ndim)
    sites = float(size[0]) * size[1] * size[2]
    kw = dict(BINARY_KW) if ndist == 2 else dict(eta_shear=0.1, eta_bulk=0.1)
    if pyref.available(nvel):
        kind = "reference"
        with pyref.Ref(nvel, size, ndist=ndist) as ref:
            ref.physics(**kw)
            if ndist == 2:
                ref.fe_symm(BINARY_FE["a"], BINARY_FE["b"], BINARY_FE["kappa"])
                ref.grad_select(gradient)
            ref.f()[...] = f0
            ref.halo()
            t1 = ref.step(max(warmup, 1)) / max(warmup, 1)
            settle = 0.0
            while settle < 1.0:                 # first touch, thread pool: let the step time settle
                t1 = ref.step(1)
                settle += max(t1, 0.05)
            n = nsteps_fixed or max(2, min(200, int(seconds / max(t1, 1e-6))))
            t = ref.step(n)
    else:
        kind = "port"
        orc = pyoracle.Oracle(nvel, size, ndist=ndist)
        prm = pyoracle.make_param(**kw, **(BINARY_FE if ndist == 2 else {}))
        orc.gradient_select(gradient)
        f, fp = f0.copy(), np.zeros_like(f0)
        ns = orc.nsite
        rho, u, frc = np.zeros(ns), np.zeros((ns, 3)), np.zeros((ns, 3))
        phi, grad, delsq = np.zeros(ns), np.zeros((ns, 3)), np.zeros(ns)
        t0 = time.perf_counter()
        orc.step(prm, max(warmup, 1), f, fp, rho, u, frc, phi, grad, delsq)
        t1 = (time.perf_counter() - t0) / max(warmup, 1)
        n = nsteps_fixed or max(2, min(200, int(seconds / max(t1, 1e-6))))
        t0 = time.perf_counter()
        orc.step(prm, n, f, fp, rho, u, frc, phi, grad, delsq)
        t = time.perf_counter() - t0
        orc.gradient_select(7)
    return {"value": sites * n / t / 1e6, "unit": "MLUPS", "cores": cores, "kind": kind,
            "sample": "D%dQ%d %s%dx%dx%d periodic, %d steps in %.1f s, %d OpenMP threads"
                      % (ndim, nvel, "binary symmetric_lb %d-point gradient " % gradient if ndist == 2 else "",
                         size[0], size[1], size[2], n, t, cores)}, n, t


def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path (oracle/_ref =
    the unmodified Ludwig sources, OpenMP over ALL host cores; the oracle port if that
    build is absent) on this arm's workload. Each step is a bounded sample: one rank's
    block of the lattice when K + W steps of it fit in about two minutes, else a 128^3
    block of it. MLUPS of a CPU is an intensive quantity, and the reference's int indexing
    (src/memory.h:187-188) cannot hold the global lattice of the multi-GPU runs: at N > 1
    the value is still what this box's host cores sustain, to set against the box's N GPUs.
    Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ncores = os.cpu_count() or 1
    try:
        ncores = len(os.sched_getaffinity(0)) or ncores
    except Exception:
        pass
    if os.environ.get("OMP_NUM_THREADS") != str(ncores) and not os.environ.get("B200LB_BENCH_REEXEC"):
        # torch.distributed.run exports OMP_NUM_THREADS=1. An OpenMP runtime that starts like
        # that stays slow for its first second of parallel regions even after the thread count
        # is raised: start this (rank 0, CPU-only) process again with the right environment.
        os.environ["OMP_NUM_THREADS"] = str(ncores)
        os.environ["B200LB_BENCH_REEXEC"] = "1"
        sys.stdout.flush()
        os.execv(sys.executable, [sys.executable] + sys.argv)
    nvel, ndist = args.nvel, args.ndist
    ndim = 2 if nvel == 9 else 3
    grid = tuple(args.grid) if args.grid else grid_for(args.gpus)
    local = local_size(args, grid)
    cannot = [name for name, on in (("walls", args.walls), ("isothermal fluctuations", args.kt > 0.0),
                                    ("Lees-Edwards planes", args.le),
                                    ("finite-difference order parameter", args.fd)) if on]
    if cannot:
        print(json.dumps({"impl": "reference", "unavailable":
                          "the reference arm times the periodic single-fluid and binary workloads; "
                          "not run with: " + ", ".join(cannot)}), flush=True)
        return
    small = (1024, 1024, 1) if nvel == 9 else (128, 128, 128)
    if args.ref_size:
        size = tuple(args.ref_size)
    else:
        probe, _, _ = cpu_reference(nvel, small, 0.0, nsteps_fixed=2, warmup=1, ndist=ndist,
                                    gradient=args.gradient)
        t_full = float(np.prod(local)) / (probe["value"] * 1e6)
        fits_int = float(np.prod([n + 2 for n in local])) * ndist * nvel < 2.0**31
        size = local if (fits_int and t_full * (args.steps + args.warmup) < 120.0) else small
    base, n, t = cpu_reference(nvel, size, 0.0, nsteps_fixed=args.steps, warmup=args.warmup,
                               ndist=ndist, gradient=args.gradient)
    line = {
        "impl": "reference",
        "metric": "D3Q19 fp64 MLUPS" if nvel == 19 else "D%dQ%d fp64 MLUPS" % (ndim, nvel),
        "value": base["value"], "unit": "MLUPS", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / n, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(nvel, ndist, local, grid),
                   "sample": base["sample"],
                   "note": "one block of the lattice on all host cores of the box"
                           if args.gpus > 1 else None},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": "MLUPS", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def local_size(args, grid):
    if args.global_size:
        g = tuple(args.global_size)
        for a in range(3):
            if g[a] % grid[a]:
                raise SystemExit("--global-size %s is not divisible by the grid %s" % (g, grid))
        return tuple(g[a] // grid[a] for a in range(3))
    return tuple(args.local or ((4096, 4096, 1) if args.nvel == 9 else (256, 256, 256)))


class Comm:
    """torch.distributed as plumbing: ids and IPC handles travel through it, timings are
    maximised over ranks; the data path itself never touches it."""

    def __init__(self, torch, dist, rank, world):
        self.torch, self.dist, self.rank, self.world = torch, dist, rank, world

    def barrier(self, lat=None):
        if lat is not None:
            lat.sync()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, xs):
        if self.world == 1:
            return [float(x) for x in xs]
        t = self.torch.tensor(list(xs), dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return [float(x) for x in t.tolist()]

    def connect(self, lb, lat, p2p=True):
        """NCCL communicator and peer-memory mappings of one lattice."""
        torch, dist = self.torch, self.dist
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if self.rank == 0:
            uid = torch.tensor(list(lb.Lattice.nccl_unique_id()), dtype=torch.uint8, device="cuda")
        dist.broadcast(uid, 0)
        lat.comm_init(bytes(uid.cpu().tolist()))
        if p2p:
            mine = torch.tensor(list(lat.p2p_export()), dtype=torch.uint8, device="cuda")
            allh = [torch.empty_like(mine) for _ in range(self.world)]
            dist.all_gather(allh, mine)
            lat.p2p_import(b"".join(bytes(t.cpu().tolist()) for t in allh), self.world)


def make_input(nvel, ndist, local, ntotal, offset, mode):
    """Synthetic lattice, host side in the reference's AOS layout. seeded: every site from
    the counter-based generator keyed on the GLOBAL site id (any decomposition sees the
    same data); tiled: a seeded 37 x 41 x 43 block repeated (seconds instead of minutes)."""
    from ludwig_b200 import models_host, synthetic
    cv, wv = models_host.cv(nvel), models_host.wv(nvel)
    ndim = 2 if nvel == 9 else 3
    if mode == "seeded":
        return synthetic.make_state(cv, wv, local, ndist=ndist, ntotal=ntotal, offset=offset, ndim=ndim)
    return synthetic.tiled_state(cv, wv, local, ndist=ndist, ndim=ndim)


def timed_steps(lat, comm, steps, warmup):
    """(ms per step maximised over ranks, launches per step) of the whole step; and the sweep
    launches ALONE, bracketed by CUDA events inside the library: (ms per step, launches per
    step) -- what the roofline of the dominant kernel is quoted on."""
    lat.step(max(warmup, 3))
    comm.barrier(lat)
    l0 = lat.launch_count()
    lat.timer_start()
    lat.step(steps)
    ms = lat.timer_stop()
    comm.barrier(lat)
    nl = (lat.launch_count() - l0) / float(steps)
    ms = comm.max(ms) / steps
    lat.kernel_timer_start(8 * steps + 64)
    lat.step(steps)
    kms, nk = lat.kernel_timer_stop()
    comm.barrier(lat)
    return ms, nl, comm.max(kms) / steps, nk / float(steps)


def roofline(nvel, ndist, local, sites_local, step_ms, sweep_ms, sweeps_per_step, hydro_store):
    """Dominant kernel = the fused stream-and-collide sweep. Algorithmic bytes per site
    (SURVEY 8(d)): 2 Q 8 per distribution. The binary step moves its 608 B in two kernels
    (phi_from_g reads g, the sweep reads f and writes f and g): its kernel figure is quoted
    on the sweep's own share, 3 Q 8 = 456 B; step_frac always on the full figure."""
    hbm, src = peaks()
    balg = B_ALG[nvel] * ndist
    kalg = balg if ndist == 1 else 3 * B_ALG[nvel] // 2
    ach = kalg * sites_local / (sweep_ms * 1e-3) / 1e9
    return {"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm,
            "traffic": ncu_traffic(nvel, ndist, local, hydro_store),
            "kernel": "lb_sweep_kernel<pull,collide>", "kernel_ms": sweep_ms,
            "sweep_launches_per_step": sweeps_per_step,
            "bytes_per_site": kalg, "step_bytes_per_site": balg, "peak_source": src,
            "step_frac": balg * sites_local / (step_ms * 1e-3) / 1e9 / hbm,
            "timing": "CUDA events around the sweep launches only (b200lb_kernel_timer_*)"}


def configure(lb, lat, args, nvel, ndist, gradient=7):
    lat.set_option("fused", args.fused)
    lat.set_option("block", args.block)
    lat.set_option("halo", lb.HALO_REDUCED if args.halo == "reduced" else lb.HALO_FULL)
    lat.set_option("hydro_write", args.hydro_write)
    lat.set_option("overlap", args.overlap)
    lat.set_option("direct", args.direct)
    lat.set_option("zpack", args.zpack)
    lat.set_option("threads", args.threads)
    if ndist == 2:
        lat.set_param(kt=args.kt, **BINARY_KW, **BINARY_FE)
        lat.field_grad_set(gradient)
    else:
        lat.set_param(eta_shear=0.1, eta_bulk=0.1, kt=args.kt)


def small_config(lb, comm, args, nvel, ndist, local, steps, name):
    """One of the other named shapes, single GPU: whole-step MLUPS and the sweep's roofline."""
    lat = lb.Lattice(nvel, local, ndist=ndist, device=0)
    configure(lb, lat, args, nvel, ndist)
    f0 = make_input(nvel, ndist, local, local, (0, 0, 0), "tiled")
    lat.lb_memcpy_h2d(f0)
    sites = float(local[0]) * local[1] * local[2]
    m0 = lat.stats()[0]
    ms, nl, kms, nk = timed_steps(lat, comm, steps, 5)
    m1 = lat.stats()[0]
    lat.close()
    r = roofline(nvel, ndist, local, sites, ms, kms, nk, False)
    return {"workload": name, "value": sites / ms / 1e3, "unit": "MLUPS", "steps": steps,
            "ms_per_step": ms, "launches_per_step": nl, "roofline": r,
            "mass_drift_rel": abs(m1 - m0) / abs(m0)}


def decomposed_equals_single(lb, comm, grid, nvel=19, per_rank=(32, 32, 32), nsteps=10, lattices=2):
    """Parity inside the bench run: every rank steps its 32^3 block of a small lattice with
    the same options as the timed run, rank 0 also steps the whole lattice on its GPU alone;
    the interiors must be identical bit for bit (what the reference's pmpi08-* logs test
    against serial logs)."""
    torch, dist = comm.torch, comm.dist
    ntotal = tuple(per_rank[a] * grid[a] for a in range(3))
    dec = lb.Decomposition(ntotal, comm.rank, comm.world, grid=grid)
    lat = lb.Lattice(nvel, per_rank, decomp=dec, device=torch.cuda.current_device(), lattices=lattices)
    lat.set_option("halo", lb.HALO_REDUCED)
    lat.set_param(eta_shear=0.1, eta_bulk=0.1)
    comm.connect(lb, lat)
    lat.lb_memcpy_h2d(make_input(nvel, 1, per_rank, ntotal, dec.offset, "seeded"))
    comm.barrier(lat)
    lat.step(nsteps)
    nall = tuple(n + 2 for n in per_rank)
    mine = lat.lb_memcpy_d2h().reshape(nall + (nvel,))[1:-1, 1:-1, 1:-1]
    comm.barrier(lat)
    lat.close()
    t = torch.tensor(np.ascontiguousarray(mine), device="cuda")
    parts = [torch.empty_like(t) for _ in range(comm.world)]
    dist.all_gather(parts, t)
    ok = None
    if comm.rank == 0:
        one = lb.Lattice(nvel, ntotal, device=torch.cuda.current_device())
        one.set_option("halo", lb.HALO_REDUCED)
        one.set_param(eta_shear=0.1, eta_bulk=0.1)
        one.lb_memcpy_h2d(make_input(nvel, 1, ntotal, ntotal, (0, 0, 0), "seeded"))
        one.step(nsteps)
        gall = tuple(n + 2 for n in ntotal)
        whole = one.lb_memcpy_d2h().reshape(gall + (nvel,))[1:-1, 1:-1, 1:-1]
        one.close()
        ok = True
        for r in range(comm.world):
            d = lb.Decomposition(ntotal, r, comm.world, grid=grid)
            o = d.offset
            blk = whole[o[0]:o[0] + per_rank[0], o[1]:o[1] + per_rank[1], o[2]:o[2] + per_rank[2]]
            ok = ok and bool(np.array_equal(blk, parts[r].cpu().numpy()))
    return ok


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import ludwig_b200 as lb

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # convenience: re-launch under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                   "--nproc-per-node", str(args.gpus), "--master-addr", "127.0.0.1",
                   "--master-port", "29517"] + sys.argv
            sys.exit(subprocess.call(cmd))
        raise SystemExit("WORLD_SIZE %d != --gpus %d" % (world, args.gpus))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    comm = Comm(torch, dist, rank, world)

    nvel, ndist = args.nvel, args.ndist
    ndim = 2 if nvel == 9 else 3
    grid = tuple(args.grid) if args.grid else grid_for(world)
    if grid is None:
        raise SystemExit("--gpus must be 1, 2, 4 or 8")
    if ndim == 2 and grid[2] != 1:
        grid = (grid[0], grid[1] * grid[2], 1)
    local = local_size(args, grid)
    ntotal = tuple(local[a] * grid[a] for a in range(3))
    periodic = tuple(0 if w else 1 for w in args.walls) if args.walls else (1, 1, 1)
    dec = lb.Decomposition(ntotal, rank, world, grid=grid, periodic=periodic)
    sites_local = float(local[0]) * local[1] * local[2]
    sites_total = sites_local * world

    # ---- parity inside the run (N > 1): decomposed == single GPU, bit for bit -------
    parity = {}
    if world > 1 and not args.no_parity:
        parity["decomposed_equals_single"] = decomposed_equals_single(lb, comm, grid, lattices=args.lattices)
        parity["what"] = ("D3Q19 %s per rank on the %s grid, 10 steps, peer-memory halo, lattices = %d: "
                          "interiors compared bit for bit on rank 0 with the whole (two-lattice) "
                          "lattice on one GPU" % ("32^3", "x".join(str(g) for g in grid), args.lattices))

    lat = lb.Lattice(nvel, local, ndist=ndist, decomp=dec, device=local_rank, zalign=args.zalign,
                     lattices=args.lattices)
    configure(lb, lat, args, nvel, ndist, args.gradient)
    if args.vec:
        lat.set_option("vec", args.vec)
    args.vec = lat.get_option("vec")
    if args.loopback:
        lat.set_option("loopback", args.loopback)
    if args.walls:
        lat.wall_commit(args.walls)
    if args.kt > 0.0:
        lat.noise_init(0)
        lat.noise_present_set(1)
    if args.le:
        lat.le_commit(args.le, 0.01)
    if args.fd:
        lat.set_param(eta_shear=0.00625, eta_bulk=0.00625, mobility=1.25, kt=args.kt, **BINARY_FE)
        lat.phi_ch_create(args.fd, lb.FORCE_STRESS_DIVERGENCE)
        lat.field_grad_set(args.gradient)
    if world > 1:
        comm.connect(lb, lat, p2p=bool(args.p2p))
    elif args.loopback and args.p2p:
        lat.p2p_import(None, 1)

    # synthetic input, host side in the reference's AOS layout
    mode = args.init
    if mode == "auto":
        mode = "seeded" if sites_local <= 2.0**24 else "tiled"
    lattice_bytes = float(np.prod([n + 2 for n in local])) * ndist * nvel * 8
    if lattice_bytes > 12.0e9:
        # a host copy of the whole lattice (plus what numpy needs to build it) times the ranks of
        # the box would not fit: x-slabs of the tiled input, uploaded piece by piece
        from ludwig_b200 import models_host, synthetic
        mode, f0 = "tiled, uploaded in x-slabs", None
        for first, part in synthetic.tiled_slabs(models_host.cv(nvel), models_host.wv(nvel), local,
                                                 ndist=ndist, ndim=ndim):
            lat.lb_memcpy_h2d_range(part, first)
    else:
        f0 = make_input(nvel, ndist, local, ntotal, dec.offset, mode)
        lat.lb_memcpy_h2d(f0)
    if args.fd:
        rng = np.random.default_rng(8361235 + rank)      # spinodal start: phi = 0.1 (r - 1/2)
        phi0 = np.zeros(tuple(n + 2 for n in local))
        phi0[1:-1, 1:-1, 1:-1] = 0.1 * (rng.random(local) - 0.5)
        lat.field_h2d(lb.FIELD_PHI, np.ascontiguousarray(phi0.reshape(-1)))
        del phi0
    sums0 = comm.sum(lat.stats()[:7])                    # mass, ., ., ., momentum before

    # ---- device-resident throughput: the whole step, and the sweep alone -------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = lat.launch_count()
    ms, nl, kms, nk = timed_steps(lat, comm, args.steps, args.warmup)
    launches = int(round(nl * args.steps))
    z_faces_packed = bool(lat.get_option("zpacked"))     # the state the timed sweeps left behind
    clocks = sampler.result()
    mlups = sites_total / ms / 1e3
    stores_hydro = (args.hydro_write == 1) or bool(args.fd) or bool(args.le)
    roof = roofline(nvel, ndist, local, sites_local, ms, kms, nk, stores_hydro) if args.fused else None
    sums1 = comm.sum(lat.stats()[:7])
    parity["mass_drift_rel"] = abs(sums1[0] - sums0[0]) / abs(sums0[0])
    parity["momentum_drift_per_site"] = [abs(sums1[4 + a] - sums0[4 + a]) / sites_total for a in range(3)]
    parity["drift_over_steps"] = 2 * args.steps + 2 * max(args.warmup, 3)

    # ---- the reference's behaviour: rho,u stored by every collision ------------------
    every = None
    if args.hydro_write != 1 and args.fused and not args.no_variants:
        lat.set_option("hydro_write", 1)
        ems, enl, ekms, enk = timed_steps(lat, comm, args.steps, 3)
        lat.set_option("hydro_write", args.hydro_write)
        r1 = roofline(nvel, ndist, local, sites_local, ems, ekms, enk, True)
        every = {"value": sites_total / ems / 1e3, "unit": "MLUPS", "ms_per_step": ems,
                 "frac": r1["frac"], "step_frac": r1["step_frac"], "traffic": r1["traffic"],
                 "what": "option hydro_write = 1: the collision stores rho,u (32 B/site) every step "
                         "as src/collision.c:570-579 does; `value` is the default, which stores them "
                         "only when they were asked for since the previous collision"}

    # ---- single-lattice streaming (AA pattern): the same steps in half the memory --------
    single = None
    if (args.lattices == 2 and args.fused and not args.no_variants and world == 1 and ndist == 1
            and not (args.walls or args.le or args.fd or args.kt > 0.0 or args.loopback)):
        lat.set_option("lattices", 1)
        sms, snl, skms, snk = timed_steps(lat, comm, args.steps + (args.steps % 2), 4)
        ms1 = lat.stats()[0]
        lat.set_option("lattices", 2)
        r2 = roofline(nvel, ndist, local, sites_local, sms, skms, snk, False)
        single = {"value": sites_total / sms / 1e3, "unit": "MLUPS", "ms_per_step": sms,
                  "frac": r2["frac"], "step_frac": r2["step_frac"],
                  "lattice_bytes_on_device": lattice_bytes, "mass_drift_rel": abs(ms1 - sums0[0]) / abs(sums0[0]),
                  "what": "option lattices = 1: one array instead of f and fprime; an in-place sweep "
                          "and a pull-collide-push sweep alternate (bit-identical lattices, "
                          "tests/test_zzz_gpu_single_lattice.py); timed over an even number of steps"}

    # ---- finite-difference order parameter: its passes alone --------------------------
    fd = None
    if args.fd:
        lat.step(1)
        comm.barrier(lat)
        k0 = lat.launch_count()
        lat.timer_start()
        for _ in range(args.steps):
            lat.field_halo()
            lat.field_grad_compute()
            lat.phi_force_calculation()
            lat.phi_cahn_hilliard()
        fms = lat.timer_stop() / args.steps
        fnk = (lat.launch_count() - k0) / args.steps
        fd_bytes = 40.0 + 96.0      # gradient: phi in, grad + delsq out; update: 8 doubles in, force + phi out
        fd = {"advection_order": args.fd, "ms_per_step": fms, "launches_per_step": fnk,
              "bytes_per_site": fd_bytes,
              "achieved_gbs": fd_bytes * sites_local / (fms * 1e-3) / 1e9,
              "frac": fd_bytes * sites_local / (fms * 1e-3) / 1e9 / peaks()[0]}

    # ---- end to end through the host call surface with HOST buffers ---------------
    # The user-visible job: the lattice lives in a page-locked host array in the reference's
    # AOS layout; lb_memcpy H2D, then every step = the hot path + the conserved sums of that
    # step read back to the host (gathered inside the collision sweep), then lb_memcpy D2H.
    e2e = None
    if not args.no_e2e and f0 is not None:
        hbuf = lb.pinned_empty(f0.shape)
        hbuf[...] = f0
        lat.set_option("diag", 1)
        comm.barrier(lat)
        t0 = time.perf_counter()
        lat.lb_memcpy_h2d(hbuf)
        lat.sync()
        t1 = time.perf_counter()
        mass = 0.0
        for _ in range(args.steps):
            lat.step(1)
            mass = lat.step_diag()[0]           # 64 B device -> host, every step
        lat.sync()
        t2 = time.perf_counter()
        lat.lb_memcpy_d2h(hbuf)
        lat.sync()
        t3 = time.perf_counter()
        te = comm.max(t3 - t0)
        lat.set_option("diag", 0)
        copy_s, loop_s = (t1 - t0) + (t3 - t2), t2 - t1
        e2e = {"value": sites_total * args.steps / te / 1e6, "unit": "MLUPS",
               "h2d_bytes_per_step": lattice_bytes / args.steps,
               "d2h_bytes_per_step": lattice_bytes / args.steps + 64,
               "seconds": te, "copy_seconds": copy_s, "loop_seconds": loop_s,
               "steps_at_which_copies_equal_compute": copy_s / (loop_s / args.steps),
               "what": "pinned host lattice (reference AOS layout): lb_memcpy H2D, %d x "
                       "[step + per-step conserved sums D2H], lb_memcpy D2H; wall clock. The two "
                       "copies are paid once per job, so the figure grows with --steps until the "
                       "loop dominates" % args.steps,
               "mass": float(mass)}
        lb.pinned_free(hbuf)
    del f0
    lat.close()

    # ---- the other named shapes (BASELINE configs 0, 2, 5), one GPU ------------------
    configs = None
    if world == 1 and not args.no_configs and args.local is None and nvel == 19 and ndist == 1 \
            and not (args.walls or args.le or args.fd or args.kt > 0.0 or args.loopback):
        configs = [
            small_config(lb, comm, args, 19, 1, (64, 64, 64), 1000,
                         "configs[0]: D3Q19 single-fluid MRT 64^3 periodic, 1000 steps"),
            small_config(lb, comm, args, 19, 2, (128, 128, 128), 200,
                         "configs[2]: binary symmetric_lb D3Q19 + 3d_7pt_fluid gradient, 128^3"),
            small_config(lb, comm, args, 27, 1, (256, 256, 256), 100,
                         "configs[4]: D3Q27 256^3 collide+stream"),
            small_config(lb, comm, args, 9, 1, (4096, 4096, 1), 100,
                         "configs[4]: D2Q9 4096^2 collide+stream"),
        ]

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            size = tuple(args.ref_size or ((1024, 1024, 1) if nvel == 9 else (128, 128, 128)))
            cpu, _, _ = cpu_reference(nvel, size, args.cpu_seconds, ndist=ndist, gradient=args.gradient)
        except Exception as exc:       # the checker failing must not hide the GPU number
            cpu = {"value": None, "error": str(exc)}

    if rank == 0:
        par = "x".join(str(g) for g in grid)
        line = {
            "metric": "D3Q19 fp64 MLUPS" if nvel == 19 else "D%dQ%d fp64 MLUPS" % (ndim, nvel),
            "value": mlups, "unit": "MLUPS", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms,
            "higher_is_better": True, "scaling": "strong" if args.global_size else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {
                "workload": workload_name(nvel, ndist, local, grid),
                "global_lattice": list(ntotal), "decomposition": par, "input": mode,
                "fused": bool(args.fused), "vec": args.vec, "halo": args.halo, "lattices": args.lattices,
                "hydro_write": args.hydro_write, "overlap": bool(args.overlap),
                "direct_halo": bool(args.direct),
                "p2p_halo": bool(args.p2p and (world > 1 or args.loopback)),
                "z_faces_packed": z_faces_packed,
                "loopback": args.loopback,
                "gradient_points": args.gradient if ndist == 2 else None,
                "walls": args.walls, "lees_edwards_planes": args.le or None,
                "fd_order_parameter": fd,
                "isothermal_fluctuations_kt": args.kt if args.kt > 0.0 else None,
                "l2": "lattice %.2f GB per GPU >> 126 MB L2 (no flush needed)" % (2 * lattice_bytes / 1e9),
            },
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "hydro_every_step": every,
            "single_lattice": single,
            "configs": configs, "parity": parity,
            "gpu_launches": launches, "launches_per_step": nl, "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
