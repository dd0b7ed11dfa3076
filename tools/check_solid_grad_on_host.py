"""DEVELOPMENT TOOL (no GPU needed): steps through the source of grad27_solid_kernel
(ludwig_b200/csrc/solid_grad_kernels.cuh) on the host and compares with the oracle's
orc_grad_27pt_solid (which reproduces the reference's log serial-spin-lb2).

    python tools/check_solid_grad_on_host.py
"""

import ctypes
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle           # noqa: E402

SO = os.path.join(ROOT, "tools", "_solid_grad_emulate.so")


def main():
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-w", "-std=c++17", "-I/usr/local/cuda/include",
                           "-shared", "-fPIC", "-o", SO, os.path.join(ROOT, "tools", "solid_grad_emulate.cpp")])
    E = ctypes.CDLL(SO)
    dp = ctypes.POINTER(ctypes.c_double)
    bad = 0
    for nlocal in ((8, 6, 10), (4, 5, 140), (12, 12, 12)):
        for walls in ((1, 1, 1), (0, 1, 1), (1, 0, 0), (0, 0, 1), (0, 0, 0)):
            orc = pyoracle.Oracle(19, nlocal, ndist=2)
            n = orc.nsite
            m = orc.interior_mask()
            rng = np.random.default_rng(4)
            phi = np.zeros(orc.nall)
            phi[1:-1, 1:-1, 1:-1] = rng.random(nlocal) - 0.5
            phi = phi.reshape(-1)
            orc.halo(phi, 1)
            w = orc.wall(walls) if any(walls) else None
            go, do = np.zeros((n, 3)), np.zeros(n)
            orc.L.orc_grad_27pt_solid.argtypes = [ctypes.c_void_p, ctypes.c_void_p, dp, dp, dp]
            rc = orc.L.orc_grad_27pt_solid(ctypes.byref(orc.box), ctypes.byref(w) if w is not None else None,
                                           phi.ctypes.data_as(dp), go.ctypes.data_as(dp), do.ctypes.data_as(dp))
            assert rc == 0
            ge, de = np.zeros((n, 3)), np.zeros(n)
            E.solid_grad_emulate((ctypes.c_int*3)(*nlocal), (ctypes.c_int*3)(*walls), phi.ctypes.data_as(dp),
                                 ge.ctypes.data_as(dp), de.ctypes.data_as(dp))
            ok = np.array_equal(ge[m], go[m]) and np.array_equal(de[m], do[m]) and np.abs(go[m]).max() > 0
            print("%s walls %s: %s" % (nlocal, walls, "ok" if ok else "DIFF"))
            bad += (not ok)
    print("all identical to the oracle" if bad == 0 else "%d case(s) FAILED" % bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
