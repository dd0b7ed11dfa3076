"""Multi-GPU parity worker. Launched by tests/test_multi_gpu.py (or by hand) as

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N \
        --master-addr 127.0.0.1 --master-port P tests/mgpu_worker.py [--nvel 19] ...

Every rank owns one block of a 3-d decomposed periodic lattice on its own GPU, halo
planes travel with NCCL send/recv inside libb200lb. Rank 0 also runs the WHOLE lattice
on its GPU and on the CPU oracle:
  * decomposed result == single-GPU result, bit for bit (decomposition independence,
    the property the reference's pmpi08-* regression logs test against serial logs);
  * result vs oracle within 1e-12 of max|f|; halo planes after lb_halo bit-exact.
Exit code 0 = pass.
"""

import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

import lbinit                       # noqa: E402
import ludwig_b200 as lb            # noqa: E402
from ludwig_b200 import models_host  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nvel", type=int, default=19)
    ap.add_argument("--ndist", type=int, default=1)
    ap.add_argument("--lsize", type=int, nargs=3, default=[12, 10, 8])
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--halo", default="full")
    ap.add_argument("--fused", type=int, default=1)
    ap.add_argument("--overlap", type=int, default=1)
    ap.add_argument("--direct", type=int, default=1)
    ap.add_argument("--p2p", type=int, default=1)
    ap.add_argument("--zpack", type=int, default=1)
    ap.add_argument("--kt", type=float, default=0.0, help="> 0: isothermal fluctuations")
    ap.add_argument("--gradient", type=int, default=7)
    ap.add_argument("--walls", type=int, nargs=3, default=[0, 0, 0])
    ap.add_argument("--le", type=int, default=0, help="Lees-Edwards planes per rank in x (uy 0.03)")
    ap.add_argument("--fd", type=int, default=0,
                    help="free_energy symmetric: finite-difference order parameter, advection of this order")
    ap.add_argument("--visc", type=int, default=0, help="with --fd: viscosity_model arrhenius")
    ap.add_argument("--lattices", type=int, default=2,
                    help="1: single-lattice streaming from the start (desc.lattices = 1)")
    ap.add_argument("--grid", type=int, nargs=3, default=None,
                    help="process grid (default: 2x1x1, 2x2x1, 2x2x2 by number of ranks)")
    args = ap.parse_args()

    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    nvel, ndist = args.nvel, args.ndist
    ndim = 2 if nvel == 9 else 3
    grid = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}[world]
    if ndim == 2 and grid[2] != 1:
        grid = (grid[0] * grid[2], grid[1], 1)
    if args.le:
        # no decomposition in y with planes (src/model_le.c:399 is not on this path)
        grid = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 1, 2), 8: (4, 1, 2)}[world]
        if ndim == 2:
            grid = (world, 1, 1)
    if args.grid:
        grid = tuple(args.grid)
        assert grid[0]*grid[1]*grid[2] == world, "--grid does not match the number of ranks"
    local = tuple(args.lsize)
    ntotal = tuple(local[a] * grid[a] for a in range(3))
    walls = tuple(args.walls)
    periodic = tuple(0 if w else 1 for w in walls)
    dec = lb.Decomposition(ntotal, rank, world, grid=grid, periodic=periodic)
    cv, wv = models_host.cv(nvel), models_host.wv(nvel)

    lat = lb.Lattice(nvel, local, ndist=ndist, decomp=dec, device=local_rank, lattices=args.lattices)
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid = torch.tensor(list(lb.Lattice.nccl_unique_id()), dtype=torch.uint8, device="cuda")
    dist.broadcast(uid, 0)
    lat.comm_init(bytes(uid.cpu().tolist()))
    if args.p2p:
        mine = torch.tensor(list(lat.p2p_export()), dtype=torch.uint8, device="cuda")
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        lat.p2p_import(b"".join(bytes(t.cpu().tolist()) for t in allh), world)
    lat.set_option("fused", args.fused)
    lat.set_option("overlap", args.overlap)
    lat.set_option("direct", args.direct)
    lat.set_option("zpack", args.zpack)
    lat.set_option("halo", lb.HALO_REDUCED if args.halo == "reduced" else lb.HALO_FULL)
    kw = dict(eta_shear=0.1, eta_bulk=0.1)
    if ndist == 2:
        kw = dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75, a=-0.00625,
                  b=0.00625, kappa=0.004)
    if args.fd:
        kw = dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=1.25, a=-0.00625,
                  b=0.00625, kappa=0.004)
    fg = (1e-4, 0.0, 0.0) if (any(walls) and not walls[0]) else (0.0, 0.0, 0.0)

    def configure(l):
        l.set_param(kt=args.kt, force=fg, **kw)
        if ndist == 2:
            l.field_grad_set(args.gradient)
        if any(walls):
            l.wall_commit(walls)
        if args.kt > 0.0:
            l.noise_init(0)
            l.noise_present_set(1)
        if args.le:
            l.le_commit(args.le*grid[0], 0.03)
        if args.fd:
            l.phi_ch_create(args.fd, lb.FORCE_STRESS_DIVERGENCE)
            l.field_grad_set(args.gradient)
            if args.visc:
                l.visc_arrhenius_commit(0.625, 0.00625, 1.0)

    def fd_phi(nl, nt=None, offset=(0, 0, 0)):
        # the order parameter, keyed on the global site id: any decomposition sees the same
        return 0.8 * lbinit.uniform(lbinit.global_ids(nl, nt, offset).reshape(-1), 50)

    configure(lat)

    f0 = lbinit.make_state(cv, wv, local, ndist=ndist, ntotal=ntotal, offset=dec.offset,
                           ndim=ndim)
    # 1. halo planes: scramble the halo sites, exchange, compare with the periodic images
    nall = tuple(n + 2 for n in local)
    interior = np.zeros(nall, dtype=bool)
    interior[1:-1, 1:-1, 1:-1] = True
    interior = interior.reshape(-1)
    fs = f0.copy()
    fs[~interior] = -777.0
    lat.lb_memcpy_h2d(fs)
    lat.lb_halo_swap(lb.HALO_FULL)
    fh = lat.lb_memcpy_d2h()
    ok_halo = bool(np.array_equal(fh, f0))
    if any(walls):
        ok_halo = True          # wall layers are not exchanged (and never read)

    # 2. time loop. Ranks start it at different times on purpose: with the peer-memory
    # halo an early rank must not push into a neighbour that is still uploading.
    import time
    time.sleep(0.3 * (rank % 3))
    lat.lb_memcpy_h2d(f0)
    if args.fd:
        lat.field_h2d(lb.FIELD_PHI, fd_phi(local, ntotal, dec.offset))
    lat.step(args.steps)
    fd = lat.lb_memcpy_d2h()
    st = lat.stats()
    if args.fd:
        tp = torch.from_numpy(np.ascontiguousarray(
            lat.field_d2h(lb.FIELD_PHI).reshape(nall)[1:-1, 1:-1, 1:-1])).cuda()
        pparts = [torch.empty_like(tp) for _ in range(world)] if rank == 0 else None
        dist.gather(tp, pparts, dst=0)
    mine = fd.reshape(nall + (ndist, nvel))[1:-1, 1:-1, 1:-1].copy()
    gw = torch.tensor(lat.wall_momentum() if any(walls) else np.zeros(3), device="cuda")
    dist.all_reduce(gw, op=dist.ReduceOp.SUM)
    nstate = None
    if args.kt > 0.0:
        nstate = lat.noise_memcpy_d2h().reshape(nall + (4,))[1:-1, 1:-1, 1:-1].astype(np.int64)
        tn = torch.from_numpy(np.ascontiguousarray(nstate)).cuda()
        nparts = [torch.empty_like(tn) for _ in range(world)] if rank == 0 else None
        dist.gather(tn, nparts, dst=0)

    t = torch.from_numpy(mine).cuda()
    parts = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    dist.gather(t, parts, dst=0)
    sums = torch.tensor(st, device="cuda")
    dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    flags = torch.tensor([1.0 if ok_halo else 0.0], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    lat.close()

    rc = 0
    if rank == 0:
        glob = np.zeros(ntotal + (ndist, nvel))
        for r in range(world):
            d = lb.Decomposition(ntotal, r, world, grid=grid)
            o = d.offset
            glob[o[0]:o[0] + local[0], o[1]:o[1] + local[1], o[2]:o[2] + local[2]] = parts[r].cpu().numpy()
        # the same lattice on ONE GPU
        one = lb.Lattice(nvel, ntotal, ndist=ndist, device=local_rank, periodic=periodic)
        one.set_option("fused", args.fused)
        configure(one)
        g0 = lbinit.make_state(cv, wv, ntotal, ndist=ndist, ndim=ndim)
        one.lb_memcpy_h2d(g0)
        if args.fd:
            one.field_h2d(lb.FIELD_PHI, fd_phi(ntotal))
        one.step(args.steps)
        g1 = one.lb_memcpy_d2h().reshape(tuple(n + 2 for n in ntotal) + (ndist, nvel))[1:-1, 1:-1, 1:-1]
        s1 = one.stats()
        same = bool(np.array_equal(glob, g1))
        if args.fd:
            # the order parameter: decomposed == single GPU, bit for bit
            gphi = np.zeros(ntotal)
            for r in range(world):
                o = lb.Decomposition(ntotal, r, world, grid=grid).offset
                gphi[o[0]:o[0] + local[0], o[1]:o[1] + local[1], o[2]:o[2] + local[2]] = pparts[r].cpu().numpy()
            phi1 = one.field_d2h(lb.FIELD_PHI).reshape(tuple(n + 2 for n in ntotal))[1:-1, 1:-1, 1:-1]
            same = same and bool(np.array_equal(gphi, phi1))
        if any(walls):
            gw1 = one.wall_momentum()
            same = same and bool(np.allclose(gw.cpu().numpy(), gw1, rtol=1e-10, atol=1e-13))
        if args.kt > 0.0:
            # generator state: decomposition independent, word for word
            gs = np.zeros(ntotal + (4,), dtype=np.int64)
            for r in range(world):
                o = lb.Decomposition(ntotal, r, world, grid=grid, periodic=periodic).offset
                gs[o[0]:o[0] + local[0], o[1]:o[1] + local[1], o[2]:o[2] + local[2]] = nparts[r].cpu().numpy()
            s_one = one.noise_memcpy_d2h().reshape(tuple(n + 2 for n in ntotal) + (4,))[1:-1, 1:-1, 1:-1]
            same = same and bool(np.array_equal(gs, s_one.astype(np.int64)))
        one.close()
        # and on the CPU oracle
        from oracle import pyoracle
        orc = pyoracle.Oracle(nvel, ntotal, ndist=ndist)
        prm = pyoracle.make_param(force=fg, **kw)
        f, fp = g0.copy(), np.zeros_like(g0)
        n = orc.nsite
        rho, u, frc = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 3))
        phi, grad, delsq = np.zeros(n), np.zeros((n, 3)), np.zeros(n)
        if args.kt > 0.0:
            orc.noise_attach(orc.noise_init(0), args.kt)
        if any(walls):
            orc.wall_attach(orc.wall(walls))
        if args.le:
            orc.le_attach(orc.le(args.le*grid[0], 0.03))
        orc.gradient_select(args.gradient)
        if args.fd:
            phi[...] = fd_phi(ntotal)
            orc.ch_attach(orc.ch(args.fd, pyoracle.FORCE_STRESS_DIVERGENCE))
            if args.visc:
                orc.visc_attach(np.zeros(n), 0.625, 0.00625, 1.0)
        orc.step(prm, args.steps, f, fp, rho, u, frc, phi, grad, delsq)
        orc.ch_attach(None)
        orc.visc_attach(None)
        orc.le_attach(None)
        orc.noise_attach(None, 0.0)
        orc.wall_attach(None)
        orc.gradient_select(7)
        fo = f.reshape(tuple(n + 2 for n in ntotal) + (ndist, nvel))[1:-1, 1:-1, 1:-1]
        err = float(np.max(np.abs(glob - fo)) / np.max(np.abs(fo)))
        if args.fd:
            po = phi.reshape(tuple(n + 2 for n in ntotal))[1:-1, 1:-1, 1:-1]
            err = max(err, float(np.max(np.abs(gphi - po)) / np.max(np.abs(po))))
        mass = abs(float(sums[0].item()) - s1[0])
        halo_ok = bool(flags.item() > 0.5)
        print("mgpu world=%d grid=%s nvel=%d ndist=%d halo=%s fused=%d overlap=%d direct=%d p2p=%d kt=%g grad=%d walls=%s le=%d fd=%d visc=%d: halo_bit_exact=%s "
              "decomposed==single=%s err_vs_oracle=%.2e mass_diff=%.2e"
              % (world, grid, nvel, ndist, args.halo, args.fused, args.overlap, args.direct, args.p2p, args.kt, args.gradient,
                 walls, args.le, args.fd, args.visc, halo_ok, same, err, mass),
              flush=True)
        if not (halo_ok and same and err < 1e-12 and mass < 1e-9 * n):
            rc = 1
    code = torch.tensor([rc], device="cuda")
    dist.broadcast(code, 0)
    dist.destroy_process_group()
    sys.exit(int(code.item()))


if __name__ == "__main__":
    main()
