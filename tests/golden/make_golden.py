"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built in
place from /root/reference by oracle/Makefile). Run in the build container:

    make -C oracle && python tests/golden/make_golden.py

Each file holds the seeded input state and the reference's own output arrays for
one stage of the hot path, so that machines without /root/reference (the GPU box)
can still check the oracle and the CUDA path against reference-produced numbers.
Inputs are regenerated from tests/lbinit.py (seeded), only outputs are stored,
interior sites only, to keep the fixtures small.
"""

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import lbinit                      # noqa: E402
from oracle import pyoracle, pyref  # noqa: E402

CASES = {19: (7, 6, 5), 27: (5, 6, 4), 9: (9, 7, 1)}
PHYS = dict(eta_shear=0.05, eta_bulk=0.12)
FG = (1e-4, -2e-4, 3e-4)
BIN = dict(a=-0.00625, b=0.00625, kappa=0.004, mobility=3.75, eta=0.00625)


def main():
    for nvel, nlocal in CASES.items():
        orc = pyoracle.Oracle(nvel, nlocal)
        ndim = orc.ndim
        m = orc.interior_mask()
        out = {"nlocal": np.array(nlocal)}
        fg = FG if ndim == 3 else (FG[0], FG[1], 0.0)
        f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal, ndim=ndim)
        force = lbinit.make_force(nlocal, ndim=ndim)
        for nrelax, name in ((0, "m10"), (1, "bgk"), (2, "trt")):
            if name == "trt" and nvel != 19:
                continue
            with pyref.Ref(nvel, nlocal) as ref:
                ref.physics(force=fg, **PHYS)
                ref.relaxation(nrelax)
                ref.f()[...] = f0
                ref.force()[...] = force
                ref.collide()
                out["collide_%s_f" % name] = ref.f()[m].copy()
                out["collide_%s_rho" % name] = ref.rho()[m].copy()
                out["collide_%s_u" % name] = ref.u()[m].copy()
                if name == "m10":
                    ref.halo()
                    out["halo_f"] = ref.f().copy()
                    ref.propagate()
                    out["prop_f"] = ref.f()[m].copy()
        with pyref.Ref(nvel, nlocal) as ref:
            ref.physics(eta_shear=0.1, eta_bulk=0.1)
            ref.f()[...] = f0
            ref.step(10)
            out["step10_f"] = ref.f()[m].copy()
            out["step10_rho"] = ref.rho()[m].copy()
            out["step10_u"] = ref.u()[m].copy()
        if nvel == 19:
            orc2 = pyoracle.Oracle(nvel, nlocal, ndist=2)
            g0 = lbinit.make_state(orc2.cv(), orc2.wv(), nlocal, ndist=2)
            with pyref.Ref(nvel, nlocal, ndist=2) as ref:
                ref.physics(eta_shear=BIN["eta"], eta_bulk=BIN["eta"],
                            mobility=BIN["mobility"])
                ref.fe_symm(BIN["a"], BIN["b"], BIN["kappa"])
                ref.f()[...] = g0
                ref.force()[...] = force
                ref.phi_from_lb()
                ref.phi_halo()
                out["bin_phi"] = ref.phi().copy()
                ref.grad_compute()
                out["bin_grad"] = ref.grad()[m].copy()
                out["bin_delsq"] = ref.delsq()[m].copy()
                ref.collide()
                out["bin_collide_f"] = ref.f()[m].copy()
                out["bin_collide_u"] = ref.u()[m].copy()
            with pyref.Ref(nvel, nlocal, ndist=2) as ref:
                ref.physics(eta_shear=BIN["eta"], eta_bulk=BIN["eta"],
                            mobility=BIN["mobility"])
                ref.fe_symm(BIN["a"], BIN["b"], BIN["kappa"])
                ref.f()[...] = g0
                ref.step(10)
                out["bin_step10_f"] = ref.f()[m].copy()
                out["bin_step10_phi"] = ref.phi()[m].copy()
        path = os.path.join(HERE, "ref_%s.npz" % pyref._NAMES[nvel])
        np.savez_compressed(path, **out)
        print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
