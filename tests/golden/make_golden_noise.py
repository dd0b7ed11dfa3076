"""Generates tests/golden/ref_noise.npz from the UNMODIFIED reference (oracle/_ref):
isothermal fluctuations -- the per-site generator state after noise_init
(src/noise.c:136) and the fluctuating lb_collide (src/collision.c:476-515, :884-900)
alone and inside ten time steps, for D3Q19, D2Q9 and the D3Q19 binary fluid.

    make -C oracle && python tests/golden/make_golden_noise.py

Inputs are regenerated from tests/lbinit.py (seeded); only reference outputs are stored.
"""

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import lbinit                      # noqa: E402
from oracle import pyoracle, pyref  # noqa: E402

KT = 2.13333e-05
CASES = {"q19": (19, (7, 6, 5), 1), "q9": (9, (9, 7, 1), 1), "q19b": (19, (7, 6, 5), 2)}
PHYS = {1: dict(eta_shear=0.05, eta_bulk=0.12),
        2: dict(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75)}
FE = (-0.00625, 0.00625, 0.004)


def main():
    out = {"kt": np.array(KT)}
    for tag, (nvel, nlocal, ndist) in CASES.items():
        orc = pyoracle.Oracle(nvel, nlocal, ndist=ndist)
        m = orc.interior_mask()
        out[tag + "_nlocal"] = np.array(nlocal)
        f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=ndist, ndim=orc.ndim)
        force = lbinit.make_force(nlocal, ndim=orc.ndim)
        for nrelax, name in ((0, "m10"), (1, "bgk")):
            if ndist == 2 and nrelax:
                continue
            with pyref.Ref(nvel, nlocal, ndist=ndist) as ref:
                ref.physics(**PHYS[ndist])
                ref.relaxation(nrelax)
                if ndist == 2:
                    ref.fe_symm(*FE)
                ref.noise(1, KT)
                if nrelax == 0:
                    out[tag + "_state0"] = ref.noise_state().copy()
                ref.f()[...] = f0
                ref.force()[...] = force
                if ndist == 2:
                    ref.phi_from_lb()
                    ref.phi_halo()
                    ref.grad_compute()
                ref.collide()
                out[tag + "_%s_collide_f" % name] = ref.f()[m].copy()
                out[tag + "_%s_collide_u" % name] = ref.u()[m].copy()
                out[tag + "_%s_collide_state" % name] = ref.noise_state()[m].copy()
            with pyref.Ref(nvel, nlocal, ndist=ndist) as ref:
                ref.physics(**PHYS[ndist])
                ref.relaxation(nrelax)
                if ndist == 2:
                    ref.fe_symm(*FE)
                ref.noise(1, KT)
                ref.f()[...] = f0
                ref.step(10)
                out[tag + "_%s_step10_f" % name] = ref.f()[m].copy()
                out[tag + "_%s_step10_state" % name] = ref.noise_state()[m].copy()
    with pyref.Ref(19, (4, 3, 5)) as ref:
        ref.noise(1, KT, reseed=8361235)
        out["seeded_state0"] = ref.noise_state().copy()
    path = os.path.join(HERE, "ref_noise.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
