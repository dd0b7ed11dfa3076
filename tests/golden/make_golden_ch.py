"""Generates tests/golden/ref_ch.npz from the UNMODIFIED reference (oracle/_ref) built
with a halo of two sites: the finite-difference order-parameter route that
"free_energy symmetric" selects -- phi_force_calculation (src/phi_force.c:74: chemical
stress and its divergence, or phi grad mu) and phi_cahn_hilliard
(src/phi_cahn_hilliard.c:206: advective fluxes of order 1-3, diffusive fluxes, forward
step) -- one stage at a time and inside eight time steps.

    make -C oracle && python tests/golden/make_golden_ch.py

Inputs are regenerated from tests/lbinit.py / tests/goldenlib.py (seeded); the file
stores the reference's interior values only.
"""

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import goldenlib                   # noqa: E402
import lbinit                      # noqa: E402
from oracle import pyoracle, pyref  # noqa: E402


def main():
    out = {}
    nvel = 19
    for tag, (nlocal, order, method, npt, gmu) in goldenlib.CH_CASES.items():
        P = goldenlib.CH_PHYS
        orc = pyoracle.Oracle(nvel, nlocal)
        m = orc.interior_mask()
        phi, u = goldenlib.ch_fields(nlocal)
        f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)

        def setup(ref):
            ref.physics(eta_shear=P["eta"], eta_bulk=P["eta"], mobility=P["mobility"])
            ref.fe_symm(P["a"], P["b"], P["kappa"])
            ref.grad_mu(gmu)
            ref.grad_select(npt)
            ref.inner(ref.phi())[...] = phi.reshape(orc.nall)

        with pyref.Ref(nvel, nlocal, fd=dict(order=order, method=method)) as ref:
            setup(ref)
            ref.inner(ref.u())[...] = u.reshape(orc.nall + (3,))
            ref.phi_halo()
            ref.grad_compute()
            ref.hydro_f_zero()
            ref.phi_force()
            out[tag + "_force"] = ref.inner(ref.force()).reshape(-1, 3)[m].copy()
            ref.cahn_hilliard()
            out[tag + "_phi1"] = ref.inner(ref.phi()).reshape(-1)[m].copy()
        with pyref.Ref(nvel, nlocal, fd=dict(order=order, method=method)) as ref:
            setup(ref)
            ref.inner(ref.f())[...] = f0.reshape(orc.nall + (1, nvel))
            ref.step(8)
            out[tag + "_step8_f"] = ref.inner(ref.f()).reshape(-1, 1, nvel)[m].copy()
            out[tag + "_step8_phi"] = ref.inner(ref.phi()).reshape(-1)[m].copy()
            out[tag + "_step8_u"] = ref.inner(ref.u()).reshape(-1, 3)[m].copy()
    path = os.path.join(HERE, "ref_ch.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
