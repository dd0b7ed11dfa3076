"""Host-side logic of the N > 1 path on CPU (gloo, world_size 2 and 4):
  * ludwig_b200.Decomposition: rank <-> coordinates, neighbours, offsets (cs_init,
    src/coords.c:173-215) and decomposition-independent synthetic input;
  * the halo message protocol libb200lb uses per decomposed direction (x, then y, then z;
    full-extent faces; sends [low face -> lower neighbour, high face -> upper neighbour],
    receives [from upper -> high halo, from lower -> low halo], which is what keeps the
    two-rank periodic case -- both neighbours are the same peer -- matched), modelled
    here with torch.distributed point-to-point on numpy planes and checked against the
    periodic images of the global lattice.
The CUDA/NCCL version of the same protocol is tested on GPUs in tests/test_multi_gpu.py.
"""

import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def _exchange(dec, a, nall, local):
    """Protocol model on array a[nall]; local copies for undivided directions."""
    for d in range(3):
        n = local[d]
        sl = [slice(None)] * 3

        def plane(i):
            s = list(sl)
            s[d] = i
            return tuple(s)
        if dec.cartsz[d] == 1:
            a[plane(0)] = a[plane(n)]
            a[plane(n + 1)] = a[plane(1)]
            continue
        lo, hi = dec.nbr_lo[d], dec.nbr_hi[d]
        to_lo = torch.from_numpy(np.ascontiguousarray(a[plane(1)]))
        to_hi = torch.from_numpy(np.ascontiguousarray(a[plane(n)]))
        from_hi = torch.empty_like(to_lo)
        from_lo = torch.empty_like(to_lo)
        ops = [dist.P2POp(dist.isend, to_lo, lo), dist.P2POp(dist.isend, to_hi, hi),
               dist.P2POp(dist.irecv, from_hi, hi), dist.P2POp(dist.irecv, from_lo, lo)]
        for r in dist.batch_isend_irecv(ops):
            r.wait()
        a[plane(n + 1)] = from_hi.numpy()
        a[plane(0)] = from_lo.numpy()


def _worker(rank, world, grid, ntotal, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import lbinit
    from ludwig_b200 import Decomposition
    dec = Decomposition(ntotal, rank, world, grid=grid)
    local = dec.nlocal
    nall = tuple(n + 2 for n in local)
    ids = lbinit.global_ids(local, ntotal, dec.offset).astype(np.float64)
    a = ids.copy()
    a[0, :, :] = a[-1, :, :] = -1
    a[:, 0, :] = a[:, -1, :] = -1
    a[:, :, 0] = a[:, :, -1] = -1
    _exchange(dec, a, nall, local)
    ok = bool(np.array_equal(a, ids))
    # neighbours are mutual
    t = torch.tensor([dec.nbr_lo[0], dec.nbr_hi[0]], dtype=torch.int64)
    allt = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allt, t)
    mutual = all(int(allt[int(allt[r][1])][0]) == r for r in range(world))
    q.put((rank, ok, mutual))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,grid,ntotal", [(2, (2, 1, 1), (8, 6, 4)),
                                               (2, (1, 1, 2), (4, 6, 8)),
                                               (4, (2, 2, 1), (8, 8, 4))])
def test_halo_protocol_gloo(world, grid, ntotal):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + world + grid[2]
    procs = [ctx.Process(target=_worker, args=(r, world, grid, ntotal, port, q))
             for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok and mutual for _, ok, mutual in res), res


def test_decomposition_matches_mpi_cart_order():
    from ludwig_b200 import Decomposition
    from ludwig_b200.decomp import dims_create
    assert dims_create(8) == (2, 2, 2) and dims_create(4) == (2, 2, 1) and dims_create(2) == (2, 1, 1)
    d = Decomposition((16, 16, 16), 5, 8, grid=(2, 2, 2))
    assert d.coords == (1, 0, 1)               # z fastest
    assert d.offset == (8, 0, 8) and d.nlocal == (8, 8, 8)
    assert d.nbr_lo == (1, 7, 4) and d.nbr_hi == (1, 7, 4)
    e = Decomposition((16, 16, 16), 0, 2, grid=(2, 1, 1), periodic=(0, 1, 1))
    assert e.nbr_lo[0] == -1 and e.nbr_hi[0] == 1 and e.nbr_lo[1] == 0


def test_synthetic_state_is_decomposition_independent():
    import lbinit
    from ludwig_b200 import Decomposition, models_host
    cv, wv = models_host.cv(19), models_host.wv(19)
    ntotal = (8, 6, 4)
    g = lbinit.make_state(cv, wv, ntotal).reshape(10, 8, 6, 1, 19)
    for r in range(4):
        d = Decomposition(ntotal, r, 4, grid=(2, 2, 1))
        f = lbinit.make_state(cv, wv, d.nlocal, ntotal=ntotal, offset=d.offset)
        f = f.reshape(tuple(n + 2 for n in d.nlocal) + (1, 19))
        o = d.offset
        assert np.array_equal(f[1:-1, 1:-1, 1:-1],
                              g[1 + o[0]:1 + o[0] + d.nlocal[0], 1 + o[1]:1 + o[1] + d.nlocal[1],
                                1 + o[2]:1 + o[2] + d.nlocal[2]])
