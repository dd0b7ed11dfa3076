"""Loads tests/golden/ref_*.npz (reference-produced, see tests/golden/make_golden.py)
and regenerates the matching seeded inputs."""

import os

import numpy as np

import lbinit

HERE = os.path.dirname(os.path.abspath(__file__))
NAMES = {19: "d3q19", 27: "d3q27", 9: "d2q9"}
PHYS = dict(eta_shear=0.05, eta_bulk=0.12)
FG = (1e-4, -2e-4, 3e-4)
BIN = dict(a=-0.00625, b=0.00625, kappa=0.004, mobility=3.75, eta=0.00625)


def load(nvel):
    return np.load(os.path.join(HERE, "golden", "ref_%s.npz" % NAMES[nvel]))


def inputs(nvel, cv, wv, ndist=1):
    g = load(nvel)
    nlocal = tuple(int(x) for x in g["nlocal"])
    ndim = 2 if nvel == 9 else 3
    f0 = lbinit.make_state(cv, wv, nlocal, ndist=ndist, ndim=ndim)
    force = lbinit.make_force(nlocal, ndim=ndim)
    fg = FG if ndim == 3 else (FG[0], FG[1], 0.0)
    return g, nlocal, f0, force, fg


def relerr(a, b):
    scale = max(np.max(np.abs(b)), 1e-300)
    return np.max(np.abs(a - b)) / scale


def inputs_grad27(nvel):
    """ref_grad27.npz (tests/golden/make_golden_grad27.py) and its seeded input."""
    from ludwig_b200 import models_host
    g = np.load(os.path.join(HERE, "golden", "ref_grad27.npz"))
    nlocal = tuple(int(x) for x in g["q%d_nlocal" % nvel])
    f0 = lbinit.make_state(models_host.cv(nvel), models_host.wv(nvel), nlocal, ndist=2)
    return g, nlocal, f0


NOISE_CASES = {"q19": (19, 1), "q9": (9, 1), "q19b": (19, 2)}
NOISE_PHYS = {1: dict(eta_shear=0.05, eta_bulk=0.12),
              2: dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75)}


def inputs_noise(tag):
    """ref_noise.npz (tests/golden/make_golden_noise.py) and the seeded inputs of a case."""
    from ludwig_b200 import models_host
    g = np.load(os.path.join(HERE, "golden", "ref_noise.npz"))
    nvel, ndist = NOISE_CASES[tag]
    nlocal = tuple(int(x) for x in g[tag + "_nlocal"])
    ndim = 2 if nvel == 9 else 3
    f0 = lbinit.make_state(models_host.cv(nvel), models_host.wv(nvel), nlocal, ndist=ndist,
                           ndim=ndim)
    force = lbinit.make_force(nlocal, ndim=ndim)
    return g, nvel, ndist, nlocal, f0, force


# "free_energy symmetric" (tests/golden/make_golden_ch.py): nlocal, advection order,
# fe_force_method (2 stress divergence, 3 phi grad mu), gradient stencil, external grad mu
CH_PHYS = dict(a=-0.00625, b=0.00625, kappa=0.004, mobility=1.25, eta=0.00625)
CH_CASES = {
    "o1s27": ((8, 7, 6), 1, 2, 27, (0.0, 0.0, 0.0)),
    "o2s7": ((8, 7, 6), 2, 2, 7, (0.0, 0.0, 0.0)),
    "o3s27": ((8, 7, 6), 3, 2, 27, (0.0, 0.0, 0.0)),
    "o1g27": ((6, 8, 7), 1, 3, 27, (1e-5, 2e-5, 3e-5)),
    "o3g7": ((6, 8, 7), 3, 3, 7, (-2e-5, 0.0, 1e-5)),
}


def ch_fields(nlocal):
    """phi [nsite], u [nsite,3] on the halo-1 box (seeded, decomposition independent)."""
    ids = lbinit.global_ids(nlocal).reshape(-1)
    phi = 0.4 * lbinit.uniform(ids, 50)
    u = np.zeros((ids.size, 3))
    for a in range(3):
        u[:, a] = 0.02 * lbinit.uniform(ids, 2 + a)
    return phi, u


def load_ch():
    return np.load(os.path.join(HERE, "golden", "ref_ch.npz"))
