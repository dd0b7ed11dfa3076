"""Test inputs: the seeded synthetic states (ludwig_b200/synthetic.py, shared with bench.py)
and Ludwig's own initial conditions that draw from its lattice generator."""

import numpy as np

from ludwig_b200.synthetic import (SEED, _mix, uniform, global_ids, equilibrium,   # noqa: F401
                                   make_state, make_force)


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
