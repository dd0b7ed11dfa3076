#!/usr/bin/env python
"""Two processes, one GPU each, peer-memory halo over CUDA IPC -- and nothing else: no
torch, no NCCL (the push needs neither), so compute-sanitizer sees only this library's
kernels. The decomposed run must equal the whole lattice on one GPU bit for bit.

    python tools/p2p_two_process.py [--grid 1 1 2] [--lsize 16 16 16] [--steps 6]
    compute-sanitizer --tool memcheck --target-processes all python tools/p2p_two_process.py
    python tools/p2p_two_process.py --one-gpu        # GPU as its own neighbour, one process
"""

import argparse
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def state(nvel, local, ntotal, offset):
    from ludwig_b200 import models_host, synthetic
    return synthetic.make_state(models_host.cv(nvel), models_host.wv(nvel), local, ntotal=ntotal,
                                offset=offset)


def interior(f, local, nvel):
    nall = tuple(n + 2 for n in local)
    return np.ascontiguousarray(f.reshape(nall + (nvel,))[1:-1, 1:-1, 1:-1])


def worker(rank, grid, local, steps, conn):
    import ludwig_b200 as lb
    ntotal = tuple(local[a]*grid[a] for a in range(3))
    dec = lb.Decomposition(ntotal, rank, 2, grid=grid)
    lat = lb.Lattice(19, local, decomp=dec, device=rank)
    lat.set_option("halo", lb.HALO_REDUCED)
    lat.set_param(eta_shear=0.1, eta_bulk=0.1)
    conn.send(bytes(lat.p2p_export()))
    lat.p2p_import(conn.recv(), 2)
    lat.lb_memcpy_h2d(state(19, local, ntotal, dec.offset))
    lat.step(steps)
    out = interior(lat.lb_memcpy_d2h(), local, 19)
    launches = lat.launch_count()
    lat.sync()
    conn.send((dec.offset, out, launches))
    conn.recv()                       # keep the mappings alive until both are done
    lat.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, nargs=3, default=[1, 1, 2])
    ap.add_argument("--lsize", type=int, nargs=3, default=[16, 16, 16])
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--one-gpu", action="store_true")
    args = ap.parse_args()
    import ludwig_b200 as lb
    grid, local = tuple(args.grid), tuple(args.lsize)
    ntotal = tuple(local[a]*grid[a] for a in range(3))
    if args.one_gpu:
        ntotal = local
    one = lb.Lattice(19, ntotal, device=0)
    one.set_option("halo", lb.HALO_REDUCED)
    one.set_param(eta_shear=0.1, eta_bulk=0.1)
    one.lb_memcpy_h2d(state(19, ntotal, ntotal, (0, 0, 0)))
    one.step(args.steps)
    whole = interior(one.lb_memcpy_d2h(), ntotal, 19)
    one.close()
    if args.one_gpu:
        lat = lb.Lattice(19, local, device=0)
        lat.set_option("halo", lb.HALO_REDUCED)
        lat.set_param(eta_shear=0.1, eta_bulk=0.1)
        lat.set_option("loopback", 1)
        lat.p2p_import(None, 1)
        lat.lb_memcpy_h2d(state(19, local, local, (0, 0, 0)))
        lat.step(args.steps)
        got = interior(lat.lb_memcpy_d2h(), local, 19)
        lat.close()
        ok = bool(np.array_equal(got, whole))
        print("one GPU as its own neighbour, %s, %d steps: identical to the wrapped lattice: %s"
              % (local, args.steps, ok))
        sys.exit(0 if ok else 1)
    ctx = mp.get_context("spawn")
    pipes = [ctx.Pipe() for _ in range(2)]
    procs = [ctx.Process(target=worker, args=(r, grid, local, args.steps, pipes[r][1])) for r in range(2)]
    for p in procs:
        p.start()
    handles = [pipes[r][0].recv() for r in range(2)]
    for r in range(2):
        pipes[r][0].send(handles[0] + handles[1])
    ok = True
    for r in range(2):
        o, out, launches = pipes[r][0].recv()
        blk = whole[o[0]:o[0] + local[0], o[1]:o[1] + local[1], o[2]:o[2] + local[2]]
        same = bool(np.array_equal(blk, out))
        print("rank %d offset %s: identical to the whole lattice: %s (%d launches)" % (r, o, same, launches))
        ok = ok and same
    for r in range(2):
        pipes[r][0].send(b"done")
    for p in procs:
        p.join(60)
    print("grid %s, %s per rank, %d steps, peer-memory halo: %s" % (grid, local, args.steps, "PASS" if ok else "FAIL"))
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
