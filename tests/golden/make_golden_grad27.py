"""Generates tests/golden/ref_grad27.npz from the UNMODIFIED reference (oracle/_ref):
the 27-point order-parameter gradient grad_3d_27pt_fluid_d2
(src/gradient_3d_27pt_fluid.c:75) selected through the reference's own plug point
field_grad_set(), alone and inside ten binary-fluid time steps.

    make -C oracle && python tests/golden/make_golden_grad27.py

Inputs are regenerated from tests/lbinit.py (seeded); the file stores the reference's
phi (halo included), grad and delsq (interior) and the ten-step outputs.
"""

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import lbinit                      # noqa: E402
from oracle import pyoracle, pyref  # noqa: E402

NLOCAL = {19: (7, 6, 5), 27: (5, 6, 4)}
BIN = dict(a=-0.00625, b=0.00625, kappa=0.004, mobility=3.75, eta=0.00625)


def main():
    out = {}
    for nvel, nlocal in NLOCAL.items():
        tag = "q%d_" % nvel
        orc = pyoracle.Oracle(nvel, nlocal, ndist=2)
        m = orc.interior_mask()
        out[tag + "nlocal"] = np.array(nlocal)
        g0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndist=2)
        with pyref.Ref(nvel, nlocal, ndist=2) as ref:
            ref.physics(eta_shear=BIN["eta"], eta_bulk=BIN["eta"], mobility=BIN["mobility"])
            ref.fe_symm(BIN["a"], BIN["b"], BIN["kappa"])
            ref.grad_select(27)
            ref.f()[...] = g0
            ref.phi_from_lb()
            ref.phi_halo()
            out[tag + "phi"] = ref.phi().copy()
            ref.grad_compute()
            out[tag + "grad"] = ref.grad()[m].copy()
            out[tag + "delsq"] = ref.delsq()[m].copy()
            ref.collide()
            out[tag + "collide_f"] = ref.f()[m].copy()
        with pyref.Ref(nvel, nlocal, ndist=2) as ref:
            ref.physics(eta_shear=BIN["eta"], eta_bulk=BIN["eta"], mobility=BIN["mobility"])
            ref.fe_symm(BIN["a"], BIN["b"], BIN["kappa"])
            ref.grad_select(27)
            ref.f()[...] = g0
            ref.step(10)
            out[tag + "step10_f"] = ref.f()[m].copy()
            out[tag + "step10_phi"] = ref.phi()[m].copy()
    path = os.path.join(HERE, "ref_grad27.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
