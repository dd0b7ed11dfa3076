"""Seeded synthetic lattice states shared by the oracle, reference and CUDA tests.

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


# ---------------------------------------------------------------------------
# Ludwig's own initial conditions that draw from its lattice generator
# ---------------------------------------------------------------------------

def kiss99(s):
    """One step of the per-site generator (src/noise.c:456-472) for every row of the
    uint32 array s[n, 4], in place; returns the n outputs (uint32)."""
    assert s.dtype == np.uint32
    m16 = np.uint32(65535)
    s[:, 0] = np.uint32(69069) * s[:, 0] + np.uint32(1234567)
    x = s[:, 1].copy()
    x ^= x << np.uint32(17)
    x ^= x >> np.uint32(13)
    x ^= x << np.uint32(5)
    s[:, 1] = x
    s[:, 2] = np.uint32(36969) * (s[:, 2] & m16) + (s[:, 2] >> np.uint32(16))
    s[:, 3] = np.uint32(18000) * (s[:, 3] & m16) + (s[:, 3] >> np.uint32(16))
    return s[:, 1] + (s[:, 0] ^ ((s[:, 2] << np.uint32(16)) + s[:, 3]))


def noise_state_interior(nt, master_seed):
    """noise_init (src/noise.c:136-233) for the interior sites of a one-rank box nt, in
    cs_index order: state[n, 4]."""
    seed0 = np.array([13, 12953, 712357, 22383979], dtype=np.uint32)
    if master_seed > 0:
        seed0[0] = np.uint32(master_seed)
    ig, jg, kg = np.meshgrid(np.arange(1, nt[0] + 1), np.arange(1, nt[1] + 1),
                             np.arange(1, nt[2] + 1), indexing="ij")
    n = ig.size
    local = np.empty((n, 4), dtype=np.uint32)
    local[:, 0] = seed0[0] + ig.reshape(-1).astype(np.uint32)
    local[:, 1] = seed0[1] + jg.reshape(-1).astype(np.uint32)
    local[:, 2] = seed0[2] + kg.reshape(-1).astype(np.uint32)
    local[:, 3] = seed0[3]
    state = np.empty((n, 4), dtype=np.uint32)
    for ia in range(4):
        state[:, ia] = kiss99(local)
    return state


def spinodal_phi(nt, seed, phi0, amp):
    """field_phi_init_spinodal (src/field_phi_init.c:484-519): phi = phi0 + amp (r - 1/2),
    r one full-precision uniform per interior site from a generator seeded with `seed`
    (noise_uniform_double_reap, src/noise.c:436-452). Returns phi[nt]."""
    state = noise_state_interior(nt, seed)
    r = (1.0 / 4294967295.0) * kiss99(state).astype(np.float64)
    return (phi0 + amp * (r - 0.5)).reshape(nt)


def drop_phi(nt, xi0, radius, phistar=1.0, centred=0):
    """field_phi_init_drop (src/field_phi_init.c:38-82): phistar tanh((r - radius)/xi0)
    about the centre of the system (Lmin = 1/2 left out unless centred)."""
    ax = [np.arange(1, nt[a] + 1) - (centred*0.5 + 0.5*nt[a]) for a in range(3)]
    x, y, z = np.meshgrid(*ax, indexing="ij")
    r = np.sqrt(x*x + y*y + z*z)
    return phistar*np.tanh((1.0/xi0)*(r - radius))


def patches_phi(nt, seed, patch, volminus1):
    """field_phi_init_spinodal_patches (src/field_phi_init.c:538-591): cubes of side
    `patch` at +1 or, with probability volminus1, -1; one uniform from the generator of
    the first site of each cube."""
    state = noise_state_interior(nt, seed)
    r = ((1.0/4294967295.0)*kiss99(state).astype(np.float64)).reshape(nt)
    corner = r[::patch, ::patch, ::patch]
    val = np.where(corner < volminus1, -1.0, 1.0)
    phi = np.repeat(np.repeat(np.repeat(val, patch, 0), patch, 1), patch, 2)
    return np.ascontiguousarray(phi[:nt[0], :nt[1], :nt[2]])
