"""DEVELOPMENT TOOL (no GPU needed): the finite-difference order parameter across
Lees-Edwards planes. Steps through the source of ch_kernel, the buffer kernels of
le_field_kernels.cuh and the plane passes ch_le_pass_a / ch_le_pass_b on the host
(tools/ch_emulate.cpp, in the order api.cu launches them) and compares force and updated
order parameter with the oracle, which is pinned against the reference bit for bit
(tests/test_oracle_vs_ref.py::test_cahn_hilliard_lees_edwards).

Bars: the updated order parameter bit for bit with the exact build (a few ulp with fused
multiply-add), conserved to round-off; the force within 1e-15 of its largest value (the
reference takes the divergence of stored face fluxes with planes, ch_kernel the
equivalent sum away from them, and the plane average is reduced in another order).

    python tools/check_ch_le_on_host.py
"""

import ctypes
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import goldenlib                      # noqa: E402
from oracle import pyoracle           # noqa: E402


def build(exact):
    so = os.path.join(ROOT, "tools", "_ch_emulate_%s.so" % ("exact" if exact else "fast"))
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-w", "-I/usr/local/cuda/include",
                           "-DCH_EXACT=%d" % exact, "-shared", "-fPIC", "-o", so,
                           os.path.join(ROOT, "tools", "ch_emulate.cpp")])
    return so


CASES = [((16, 8, 6), dict(nplanes=2, uy=0.05), 1, 7),
         ((16, 8, 6), dict(nplanes=2, uy=0.05), 2, 27),
         ((12, 7, 5), dict(nplanes=1, uy=-0.03), 3, 27),
         ((15, 9, 6), dict(nplanes=3, uy=0.31), 3, 7),
         ((12, 8, 6), dict(nplanes=2, uy=0.07), 4, 7),
         ((16, 12, 1), dict(nplanes=2, uy=0.02), 3, 27),
         ((16, 8, 140), dict(nplanes=2, uy=0.04, oscillatory=1, period=40, nt0=0), 2, 7)]


def main(exact):
    E = ctypes.CDLL(build(exact))
    dp = ctypes.POINTER(ctypes.c_double)
    P = goldenlib.CH_PHYS
    bad = 0
    for nlocal, le, order, npt in CASES:
        orc = pyoracle.Oracle(19, nlocal)
        n = orc.nsite
        m = orc.interior_mask()
        prm = pyoracle.make_param(eta_shear=P["eta"], eta_bulk=P["eta"], mobility=P["mobility"],
                                  a=P["a"], b=P["b"], kappa=P["kappa"])
        gmu = (1e-5, -2e-5, 3e-5)
        ch = orc.ch(order, pyoracle.FORCE_STRESS_DIVERGENCE, gmu)
        ole = orc.le(**le)
        loc = (ctypes.c_int*le["nplanes"])(*[orc.le_plane_location(ole, k) for k in range(le["nplanes"])])
        phi0, u0 = goldenlib.ch_fields(nlocal)
        orc.halo(phi0, 1)
        for t in (0.0, 1.0, 6.0, 133.0):
            grad, delsq = np.zeros((n, 3)), np.zeros(n)
            fo, po, uo = np.zeros((n, 3)), phi0.copy(), u0.copy()
            orc.ch_le_step(prm, ch, ole, t, npt, po, uo, fo, grad, delsq, do_force=0, do_update=0)
            orc.ch_le_step(prm, ch, ole, t, npt, po, uo, fo, grad, delsq)
            force, out = np.zeros((n, 3)), np.zeros(n)
            rc = E.ch_le_emulate((ctypes.c_int*3)(*nlocal), le["nplanes"], loc, ctypes.c_double(le["uy"]),
                                 le.get("oscillatory", 0), le.get("period", 0), le.get("nt0", 0),
                                 ctypes.c_double(t), npt, order,
                                 (ctypes.c_double*4)(P["a"], P["b"], P["kappa"], P["mobility"]),
                                 (ctypes.c_double*3)(*gmu),
                                 *[a.ctypes.data_as(dp) for a in (phi0, grad, delsq, u0, force, out)])
            assert rc == 0
            if exact:
                okp = np.array_equal(out[m], po[m])
            else:
                okp = np.max(np.abs(out[m] - po[m])) <= 4e-16*np.max(np.abs(po[m]))
            okp = okp and abs(out[m].sum() - phi0[m].sum()) < 1e-13*m.sum()
            okf = np.max(np.abs(force[m] - fo[m])) <= 1e-15*np.max(np.abs(fo[m]))
            print("%s %s order %d %2d-point t %6.1f  force %s (%.1e)  phi %s" % (
                nlocal, le, order, npt, t, "ok" if okf else "DIFF",
                np.max(np.abs(force[m] - fo[m]))/np.max(np.abs(fo[m])), "ok" if okp else "DIFF"))
            bad += (not okf) + (not okp)
    print("FAILED" if bad else ("exact build: order parameter identical to the oracle, force within 1e-15"
                                if exact else "fast build: within a few ulp of the oracle"))
    return bad


if __name__ == "__main__":
    sys.exit(1 if (main(1) + main(0)) else 0)
