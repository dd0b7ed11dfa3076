#!/usr/bin/env python
"""bench.py -- D3Q19 fp64 MLUPS of the Ludwig LB time step on B200, with the HBM
bytes-per-site roofline and the reference's CPU build timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W]           this implementation
    python bench.py --impl reference [--steps K] [--warmup W]     Ludwig's own CPU code

One "step" is one pass of the reference's time-step hot path (src/ludwig.c:530-862
restricted to it): hydro_f_zero, hydro_u_zero, lb_collide, lb_halo, lb_propagation
over one synthetic lattice. N = 1 runs BASELINE.json configs[1] (D3Q19 single-fluid MRT,
256^3 periodic); N > 1 is weak scaling: every GPU keeps a 256^3 block of a lattice
decomposed over the GPUs (2x1x1, 2x2x1, 4x2x1; --grid 2 2 2 for the cubic grid) with
NCCL halo exchange overlapped with the interior sweep.

Prints ONE JSON line (rank 0). Keys are documented in DESIGN.md section "Measurement".
"""

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

B_ALG = {19: 304, 27: 432, 9: 144}      # bytes per lattice-site update, SURVEY.md 8(d): 2*Q*8


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nvel", type=int, default=19)
    ap.add_argument("--local", type=int, nargs=3, default=None,
                    help="local lattice per GPU (default 256 256 256; D2Q9 4096 4096 1)")
    ap.add_argument("--ndist", type=int, default=1)
    ap.add_argument("--fused", type=int, default=1)
    ap.add_argument("--vec", type=int, default=0, help="sites per thread (0 = library default)")
    ap.add_argument("--block", type=int, default=128)
    ap.add_argument("--zalign", type=int, default=0)
    ap.add_argument("--halo", default="reduced", choices=["full", "reduced"])
    ap.add_argument("--hydro-write", type=int, default=1)
    ap.add_argument("--overlap", type=int, default=1)
    ap.add_argument("--direct", type=int, default=1)
    ap.add_argument("--p2p", type=int, default=1, help="peer-memory halo (CUDA IPC) instead of NCCL messages")
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
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
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
            time.sleep(0.05)

    def result(self):
        self.stop_flag.set()
        if self.ok and self.is_alive():
            self.join(timeout=2)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def workload_name(nvel, ndist, local, grid):
    ndim = 2 if nvel == 9 else 3
    return ("D%dQ%d %s MRT %dx%dx%d periodic per GPU, grid %s"
            % (ndim, nvel, "binary symmetric_lb" if ndist == 2 else "single-fluid",
               local[0], local[1], local[2], "x".join(str(g) for g in grid)))


def ncu_traffic(nvel, ndist, local, hydro_write, vec):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture, when
    it was taken on this very configuration; else None."""
    path = os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")
    try:
        with open(path) as fh:
            t = json.load(fh)
        if (t["nvel"], t["ndist"], tuple(t["local"]), t["hydro_write"], t["vec"]) == \
                (nvel, ndist, tuple(local), int(hydro_write), vec):
            return t["dram_bytes_per_launch"]
    except Exception:
        pass
    return None


def grid_for(n):
    # The contiguous (z) direction is left undivided: its faces are strided in memory
    # and NVSwitch gives every pair of GPUs the same bandwidth, so a 4x2x1 grid beats
    # 2x2x2 (measured, 8 x 256^3: 140.4 vs 123.4 GLUPS). --grid 2 2 2 selects the latter.
    return {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (4, 2, 1)}.get(n)


def synthetic_state(nvel, nlocal, ntotal, offset, ndist):
    """Seeded equilibrium + perturbation keyed on the global site id (tests/lbinit.py),
    identical for any decomposition. Built plane by plane to bound host memory."""
    import lbinit
    from ludwig_b200 import models_host
    cv, wv = models_host.cv(nvel), models_host.wv(nvel)
    ndim = 2 if nvel == 9 else 3
    return lbinit.make_state(cv, wv, nlocal, ndist=ndist, ntotal=ntotal, offset=offset,
                             ndim=ndim)


def cpu_reference(nvel, size, seconds, nsteps_fixed=None, warmup=1):
    """Times the UNMODIFIED reference (oracle/_ref, OpenMP over the host cores) on a
    bounded sample; falls back to the oracle port when the prebuilt reference is absent."""
    from oracle import pyref, pyoracle
    import lbinit
    from ludwig_b200 import models_host
    cores = os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    cv, wv = models_host.cv(nvel), models_host.wv(nvel)
    ndim = 2 if nvel == 9 else 3
    f0 = lbinit.make_state(cv, wv, size, ndim=ndim)
    sites = float(size[0]) * size[1] * size[2]
    if pyref.available(nvel):
        kind = "reference"
        with pyref.Ref(nvel, size) as ref:
            ref.physics(eta_shear=0.1, eta_bulk=0.1)
            ref.f()[...] = f0
            t1 = ref.step(max(warmup, 1)) / max(warmup, 1)
            n = nsteps_fixed or max(2, min(200, int(seconds / max(t1, 1e-6))))
            t = ref.step(n)
    else:
        kind = "port"
        orc = pyoracle.Oracle(nvel, size)
        prm = pyoracle.make_param(eta_shear=0.1, eta_bulk=0.1)
        f, fp = f0.copy(), np.zeros_like(f0)
        ns = orc.nsite
        rho, u, frc = np.zeros(ns), np.zeros((ns, 3)), np.zeros((ns, 3))
        t0 = time.perf_counter()
        orc.step(prm, max(warmup, 1), f, fp, rho, u, frc)
        t1 = (time.perf_counter() - t0) / max(warmup, 1)
        n = nsteps_fixed or max(2, min(200, int(seconds / max(t1, 1e-6))))
        t0 = time.perf_counter()
        orc.step(prm, n, f, fp, rho, u, frc)
        t = time.perf_counter() - t0
    return {"value": sites * n / t / 1e6, "unit": "MLUPS", "cores": cores, "kind": kind,
            "sample": "D%dQ%d %dx%dx%d periodic, %d steps in %.1f s, OMP_NUM_THREADS=%s"
                      % (ndim, nvel, size[0], size[1], size[2], n, t,
                         os.environ.get("OMP_NUM_THREADS"))}, n, t


def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path (oracle/_ref =
    the unmodified Ludwig sources, OpenMP over all host cores; the oracle port if that
    build is absent) on this arm's workload. Each step is a bounded sample: the full
    lattice when K + W steps of it fit in about two minutes, else a 128^3 block of it
    (MLUPS is an intensive quantity). Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nvel, ndist = args.nvel, args.ndist
    ndim = 2 if nvel == 9 else 3
    local = tuple(args.local or ((4096, 4096, 1) if nvel == 9 else (256, 256, 256)))
    grid = tuple(args.grid) if args.grid else grid_for(args.gpus)
    small = (1024, 1024, 1) if nvel == 9 else (128, 128, 128)
    if args.ref_size:
        size = tuple(args.ref_size)
    else:
        probe, _, _ = cpu_reference(nvel, small, 0.0, nsteps_fixed=2, warmup=1)
        t_full = float(np.prod(local)) / (probe["value"] * 1e6)
        size = local if t_full * (args.steps + args.warmup) < 120.0 else small
    base, n, t = cpu_reference(nvel, size, 0.0, nsteps_fixed=args.steps, warmup=args.warmup)
    line = {
        "impl": "reference",
        "metric": "D3Q19 fp64 MLUPS" if nvel == 19 else "D%dQ%d fp64 MLUPS" % (ndim, nvel),
        "value": base["value"], "unit": "MLUPS", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / n, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(nvel, ndist, local, grid),
                   "sample": base["sample"]},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": "MLUPS", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


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

    nvel, ndist = args.nvel, args.ndist
    ndim = 2 if nvel == 9 else 3
    local = tuple(args.local or ((4096, 4096, 1) if nvel == 9 else (256, 256, 256)))
    grid = tuple(args.grid) if args.grid else grid_for(world)
    if grid is None:
        raise SystemExit("--gpus must be 1, 2, 4 or 8")
    if ndim == 2 and grid[2] != 1:
        grid = (grid[0], grid[1] * grid[2], 1)
    ntotal = tuple(local[a] * grid[a] for a in range(3))
    periodic = tuple(0 if w else 1 for w in args.walls) if args.walls else (1, 1, 1)
    dec = lb.Decomposition(ntotal, rank, world, grid=grid, periodic=periodic)

    lat = lb.Lattice(nvel, local, ndist=ndist, decomp=dec, device=local_rank,
                     zalign=args.zalign)
    lat.set_option("fused", args.fused)
    if args.vec:
        lat.set_option("vec", args.vec)
    args.vec = lat.get_option("vec")
    lat.set_option("block", args.block)
    lat.set_option("halo", lb.HALO_REDUCED if args.halo == "reduced" else lb.HALO_FULL)
    lat.set_option("hydro_write", args.hydro_write)
    lat.set_option("overlap", args.overlap)
    lat.set_option("direct", args.direct)
    if args.loopback:
        lat.set_option("loopback", 1)
    if ndist == 2:
        lat.set_param(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75, a=-0.00625,
                      b=0.00625, kappa=0.004, kt=args.kt)
        lat.field_grad_set(args.gradient)
    else:
        lat.set_param(eta_shear=0.1, eta_bulk=0.1, kt=args.kt)
    if args.walls:
        lat.wall_commit(args.walls)
    if args.kt > 0.0:
        lat.noise_init(0)
        lat.noise_present_set(1)
    if args.le:
        lat.le_commit(args.le, 0.01)
    if args.fd:
        lat.set_param(eta_shear=0.00625, eta_bulk=0.00625, mobility=1.25, a=-0.00625,
                      b=0.00625, kappa=0.004, kt=args.kt)
        lat.phi_ch_create(args.fd, lb.FORCE_STRESS_DIVERGENCE)
        lat.field_grad_set(args.gradient)

    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid = torch.tensor(list(lb.Lattice.nccl_unique_id()), dtype=torch.uint8,
                               device="cuda")
        dist.broadcast(uid, 0)
        lat.comm_init(bytes(uid.cpu().tolist()))
        if args.p2p:
            mine = torch.tensor(list(lat.p2p_export()), dtype=torch.uint8, device="cuda")
            allh = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(allh, mine)
            lat.p2p_import(b"".join(bytes(t.cpu().tolist()) for t in allh), world)
    elif args.loopback and args.p2p:
        lat.p2p_import(None, 1)

    def barrier():
        lat.sync()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # synthetic input, host side in the reference's AOS layout
    f0 = synthetic_state(nvel, local, ntotal, dec.offset, ndist)
    lat.lb_memcpy_h2d(f0)
    if args.fd:
        # spinodal start: phi = 0.1 (r - 1/2), as phi_initialisation spinodal does
        import numpy as np
        rng = np.random.default_rng(8361235 + rank)
        phi0 = np.zeros(tuple(n + 2 for n in local))
        phi0[1:-1, 1:-1, 1:-1] = 0.1 * (rng.random(local) - 0.5)
        lat.field_h2d(lb.FIELD_PHI, np.ascontiguousarray(phi0.reshape(-1)))
        del phi0
    sites_local = float(local[0]) * local[1] * local[2]
    sites_total = sites_local * world
    lattice_bytes = f0.nbytes

    # ---- device-resident throughput: the whole step -----------------------------
    lat.step(max(args.warmup, 3))
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = lat.launch_count()
    barrier()
    lat.timer_start()
    lat.step(args.steps)
    ms = lat.timer_stop()
    barrier()
    launches = lat.launch_count() - l0
    ms = max_over_ranks(ms)
    mlups = sites_total * args.steps / (ms * 1e-3) / 1e6

    # ---- the dominant kernel alone (collide + deferred propagation = 1 launch) ----
    kernel_ms = None
    if args.fused:
        lat.lb_collide(); lat.lb_propagation()
        barrier()
        k0 = lat.launch_count()
        lat.timer_start()
        for _ in range(args.steps):
            lat.lb_collide()
            lat.lb_propagation()
        kms = lat.timer_stop()
        nk = lat.launch_count() - k0
        kernel_ms = kms / max(nk, 1)
    # ---- finite-difference order parameter: its passes alone ------------------------
    fd = None
    if args.fd:
        lat.step(1)
        barrier()
        k0 = lat.launch_count()
        lat.timer_start()
        for _ in range(args.steps):
            lat.field_halo()
            lat.field_grad_compute()
            lat.phi_force_calculation()
            lat.phi_cahn_hilliard()
        fms = lat.timer_stop() / args.steps
        nk = (lat.launch_count() - k0) / args.steps
        fd_bytes = 40.0 + 96.0      # gradient: phi in, grad + delsq out; update: 8 doubles in, force + phi out
        fd = {"advection_order": args.fd, "ms_per_step": fms, "launches_per_step": nk,
              "bytes_per_site": fd_bytes,
              "achieved_gbs": fd_bytes * sites_local / (fms * 1e-3) / 1e9,
              "frac": fd_bytes * sites_local / (fms * 1e-3) / 1e9 / peaks()[0],
              "what": "field_halo + field_grad_compute + phi_force_calculation + phi_cahn_hilliard "
                      "(gradient kernel + one fused force/update kernel), timed without the LB step"}
    # ---- the same step with rho,u produced on demand instead of every step --------
    lazy = None
    if args.hydro_write and args.fused and world == 1:
        lat.set_option("hydro_write", 0)
        lat.step(3)
        barrier()
        lat.timer_start()
        lat.step(args.steps)
        lms = lat.timer_stop()
        lat.set_option("hydro_write", 1)
        lazy = {"value": sites_total * args.steps / (lms * 1e-3) / 1e6, "unit": "MLUPS",
                "ms_per_step": lms / args.steps,
                "frac": B_ALG[nvel] * ndist * sites_local * args.steps / (lms * 1e-3) / 1e9 / peaks()[0],
                "what": "option hydro_write = 0: rho,u not stored by the collision but "
                        "recomputed from the post-collision lattice when asked for "
                        "(SURVEY 8(d): hydro I/O is outside the 304 B/site)"}
    clocks = sampler.result()

    hbm, hbm_src = peaks()
    balg = B_ALG[nvel] * ndist
    roof = None
    if kernel_ms:
        ach = balg * sites_local / (kernel_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s",
                "frac": ach / hbm,
                "traffic": ncu_traffic(nvel, ndist, local, args.hydro_write, args.vec),
                "kernel": "lb_sweep_kernel<pull,collide>",
                "kernel_ms": kernel_ms, "bytes_per_site": balg, "peak_source": hbm_src,
                "step_frac": balg * sites_local * args.steps / (ms * 1e-3) / 1e9 / hbm}

    # ---- end to end through the host call surface with HOST buffers ---------------
    # The user-visible job: the lattice lives in a (page-locked) host array in the
    # reference's AOS layout; lb_memcpy H2D, then every step = the hot path + the
    # conserved sums of that step read back to the host (gathered inside the collision
    # sweep), then lb_memcpy D2H. Wall clock around the lot.
    e2e = None
    if not args.no_e2e:
        hbuf = lb.pinned_empty(f0.shape)
        hbuf[...] = f0
        lat.set_option("diag", 1)
        barrier()
        t0 = time.perf_counter()
        lat.lb_memcpy_h2d(hbuf)
        mass = 0.0
        for _ in range(args.steps):
            lat.step(1)
            mass = lat.step_diag()[0]           # 64 B device -> host, every step
        lat.lb_memcpy_d2h(hbuf)
        lat.sync()
        te = time.perf_counter() - t0
        te = max_over_ranks(te)
        lat.set_option("diag", 0)
        e2e = {"value": sites_total * args.steps / te / 1e6, "unit": "MLUPS",
               "h2d_bytes_per_step": lattice_bytes / args.steps,
               "d2h_bytes_per_step": lattice_bytes / args.steps + 64,
               "seconds": te,
               "what": "pinned host lattice (reference AOS layout): lb_memcpy H2D, %d x "
                       "[step + per-step conserved sums D2H], lb_memcpy D2H; wall clock"
                       % args.steps,
               "mass": float(mass)}
        lb.pinned_free(hbuf)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            size = tuple(args.ref_size or ((1024, 1024, 1) if nvel == 9 else (128, 128, 128)))
            cpu, _, _ = cpu_reference(nvel, size, args.cpu_seconds)
        except Exception as exc:       # the checker failing must not hide the GPU number
            cpu = {"value": None, "error": str(exc)}

    if rank == 0:
        par = "x".join(str(g) for g in grid)
        line = {
            "metric": "D3Q19 fp64 MLUPS" if nvel == 19 else "D%dQ%d fp64 MLUPS" % (ndim, nvel),
            "value": mlups, "unit": "MLUPS", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {
                "workload": workload_name(nvel, ndist, local, grid),
                "global_lattice": list(ntotal), "decomposition": par,
                "fused": bool(args.fused), "vec": args.vec, "halo": args.halo,
                "hydro_write": bool(args.hydro_write), "overlap": bool(args.overlap), "direct_halo": bool(args.direct), "p2p_halo": bool(args.p2p and (world > 1 or args.loopback)), "loopback": bool(args.loopback),
                "gradient_points": args.gradient if ndist == 2 else None,
                "walls": args.walls,
                "lees_edwards_planes": args.le or None,
                "fd_order_parameter": fd,
                "isothermal_fluctuations_kt": args.kt if args.kt > 0.0 else None,
                "l2": "lattice %.2f GB per GPU >> 126 MB L2 (no flush needed)" % (2 * lattice_bytes / 1e9),
            },
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "hydro_on_demand": lazy,
            "gpu_launches": launches, "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    lat.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
