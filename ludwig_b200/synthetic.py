"""Seeded synthetic lattice states: the input generator of bench.py, smoke() and the tests.

Counter-based generator (splitmix64 finaliser) keyed on the GLOBAL site id and a
stream number, so any decomposition of the same global lattice sees the same data
(SURVEY.md 8(d)). Seed 8361235 is the one used throughout the reference's
regression inputs.
"""

import numpy as np

SEED = 8361235


def _mix(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & np.uint64(0xFFFFFFFFFFFFFFFF)
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def uniform(ids, stream, seed=SEED):
    """U(-1,1) doubles, one per entry of the uint64 array ids."""
    with np.errstate(over="ignore"):
        key = _mix(np.uint64(seed) + np.uint64(stream) * np.uint64(0xD1342543DE82EF95))
        x = _mix(ids.astype(np.uint64) ^ key)
    return (x >> np.uint64(11)).astype(np.float64) * (2.0 / 9007199254740992.0) - 1.0


def global_ids(nlocal, ntotal=None, offset=(0, 0, 0)):
    """Global site id of every site of the halo-inclusive local box (nhalo = 1),
    periodic images included, shape nall."""
    ntotal = tuple(ntotal or nlocal)
    ax = [((np.arange(nlocal[a] + 2) - 1 + offset[a]) % ntotal[a]) for a in range(3)]
    gx, gy, gz = np.meshgrid(*ax, indexing="ij")
    return ((gx * ntotal[1] + gy) * ntotal[2] + gz).astype(np.uint64)


def equilibrium(cv, wv, rho, u):
    """lb_1st_moment_equilib_set (src/model.c:916-941), vectorised.
    rho [n], u [n,3] -> f [n, nvel]."""
    cu = u @ cv.T.astype(np.float64)                     # [n, nvel]
    uu = np.einsum("na,na->n", u, u)
    return rho[:, None] * wv[None, :] * (1.0 + 3.0 * cu + 4.5 * cu * cu
                                           - 1.5 * uu[:, None])


def make_state(cv, wv, nlocal, ndist=1, ntotal=None, offset=(0, 0, 0),
               amp=0.01, noneq=0.01, ndim=3):
    """f in the reference host layout [nsite, ndist, nvel]; every site of the
    halo-inclusive box is filled consistently with periodic images."""
    nvel = len(wv)
    ids = global_ids(nlocal, ntotal, offset).reshape(-1)
    n = ids.size
    rho = 1.0 + amp * uniform(ids, 1)
    u = np.zeros((n, 3))
    for a in range(ndim):
        u[:, a] = amp * uniform(ids, 2 + a)
    f = np.zeros((n, ndist, nvel))
    feq = equilibrium(cv, wv, rho, u)
    for p in range(nvel):
        feq[:, p] *= 1.0 + noneq * uniform(ids, 10 + p)
    f[:, 0, :] = feq
    if ndist == 2:
        phi = 0.5 * uniform(ids, 50)
        g = wv[None, :] * phi[:, None] * 0.0
        g[:, 0] = phi
        for p in range(nvel):
            g[:, p] += 0.01 * wv[p] * uniform(ids, 60 + p)
        f[:, 1, :] = g
    return f


def make_force(nlocal, ntotal=None, offset=(0, 0, 0), amp=1.0e-3, ndim=3):
    ids = global_ids(nlocal, ntotal, offset).reshape(-1)
    force = np.zeros((ids.size, 3))
    for a in range(ndim):
        force[:, a] = amp * uniform(ids, 80 + a)
    return force


def tiled_state(cv, wv, nlocal, ndist=1, ndim=3, block=(37, 41, 43)):
    """A large input in seconds: the interior is a seeded block with pairwise coprime edges,
    repeated and cropped (no two sites a power of two apart see the same data); halo sites
    stay zero (lb_halo fills them). f [nsite, ndist, nvel] in the reference host layout."""
    nvel = len(wv)
    blk = tuple(min(block[a], nlocal[a]) for a in range(3))
    small = make_state(cv, wv, blk, ndist=ndist, ndim=ndim)
    ball = tuple(b + 2 for b in blk)
    small = small.reshape(ball + (ndist, nvel))[1:-1, 1:-1, 1:-1]
    reps = tuple(-(-nlocal[a]//blk[a]) for a in range(3))
    nall = tuple(n + 2 for n in nlocal)
    f0 = np.zeros(nall + (ndist, nvel))
    f0[1:-1, 1:-1, 1:-1] = np.tile(small, reps + (1, 1))[:nlocal[0], :nlocal[1], :nlocal[2]]
    return f0.reshape(-1, ndist, nvel)


def tiled_slabs(cv, wv, nlocal, ndist=1, ndim=3, block=(37, 41, 43), planes=16):
    """tiled_state in pieces of `planes` x-planes of the halo-inclusive box, for lattices
    whose host copy would not fit: yields (first host site, f [count, ndist, nvel]). The
    pieces together are exactly tiled_state(...)."""
    nvel = len(wv)
    blk = tuple(min(block[a], nlocal[a]) for a in range(3))
    small = make_state(cv, wv, blk, ndist=ndist, ndim=ndim)
    ball = tuple(b + 2 for b in blk)
    small = small.reshape(ball + (ndist, nvel))[1:-1, 1:-1, 1:-1]
    reps = tuple(-(-nlocal[a]//blk[a]) for a in range(3))
    # one x-period of the pattern, already repeated and cropped in y and z
    yz = np.tile(small, (1, reps[1], reps[2], 1, 1))[:, :nlocal[1], :nlocal[2]]
    nall = tuple(n + 2 for n in nlocal)
    per_plane = nall[1]*nall[2]
    for x0 in range(0, nall[0], planes):
        x1 = min(x0 + planes, nall[0])
        part = np.zeros((x1 - x0, nall[1], nall[2], ndist, nvel))
        for x in range(x0, x1):
            if 1 <= x <= nlocal[0]:
                part[x - x0, 1:-1, 1:-1] = yz[(x - 1) % blk[0]]
        yield x0*per_plane, part.reshape(-1, ndist, nvel)
