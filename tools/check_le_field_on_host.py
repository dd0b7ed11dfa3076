"""DEVELOPMENT TOOL (no GPU needed): steps through the source of the Lees-Edwards
order-parameter kernels (ludwig_b200/csrc/le_field_kernels.cuh) on the host
(tools/le_field_emulate.cpp) and compares them with the oracle, which is pinned against the
reference bit for bit (tests/test_oracle_vs_ref.py::test_lees_edwards_order_parameter).

    python tools/check_le_field_on_host.py
"""

import ctypes
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle   # noqa: E402

SO = os.path.join(ROOT, "tools", "_le_field_emulate.so")


def build():
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-w", "-std=c++17",
                           "-I/usr/local/cuda/include", "-shared", "-fPIC", "-o", SO,
                           os.path.join(ROOT, "tools", "le_field_emulate.cpp")])
    return ctypes.CDLL(SO)


def main():
    lib = build()
    dp = ctypes.POINTER(ctypes.c_double)
    ip = ctypes.POINTER(ctypes.c_int)
    lib.emu_le_grad_fix.argtypes = [ip, ip, ctypes.c_int, ip, ctypes.c_double, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int,
                                    dp, dp, dp, dp]
    bad = 0
    cases = [((16, 6, 4), dict(nplanes=2, uy=0.05)),
             ((12, 7, 5), dict(nplanes=1, uy=-0.03)),
             ((12, 8, 4), dict(nplanes=3, uy=0.02)),
             ((9, 9, 140), dict(nplanes=3, uy=0.31)),
             ((64, 64, 1), dict(nplanes=2, uy=0.05)),
             ((16, 6, 4), dict(nplanes=2, uy=0.04, oscillatory=1, period=40, nt0=0))]
    for nlocal, le in cases:
        orc = pyoracle.Oracle(19, nlocal, ndist=2)
        ole = orc.le(**le)
        n, nall = orc.nsite, orc.nall
        m = orc.interior_mask()
        rng = np.random.default_rng(5)
        phi = np.zeros(nall)
        phi[1:-1, 1:-1, 1:-1] = 0.4*(rng.random(nlocal) - 0.5)
        phi = phi.reshape(-1)
        orc.halo(phi, 1)
        loc = (ctypes.c_int*le["nplanes"])(*[orc.le_plane_location(ole, k) for k in range(le["nplanes"])])
        for npt in (7, 27):
            for wrap in ((1, 1, 1), (1, 1, 0), (0, 1, 1), (0, 0, 0)):
                for t in (0.0, 1.0, 7.0, 133.0, 1000.0):
                    g0, d0 = np.zeros((n, 3)), np.zeros(n)
                    (orc.grad_27pt if npt == 27 else orc.grad_7pt)(phi, g0, d0)
                    buf = np.zeros((2*le["nplanes"], nall[1], nall[2]))
                    rc = lib.emu_le_grad_fix((ctypes.c_int*3)(*nlocal), (ctypes.c_int*3)(*wrap), le["nplanes"], loc,
                                             le["uy"], le.get("oscillatory", 0), le.get("period", 0),
                                             le.get("nt0", 0), t, npt, phi.ctypes.data_as(dp),
                                             g0.ctypes.data_as(dp), d0.ctypes.data_as(dp), buf.ctypes.data_as(dp))
                    assert rc == 0
                    g1, d1 = np.zeros((n, 3)), np.zeros(n)
                    orc.le_field_grad(ole, t, npt, phi, g1, d1)
                    obuf = orc.le_field_buffers(ole, t, phi)
                    zs = slice(1, -1) if wrap[2] else slice(None)
                    ok = (np.array_equal(g0[m], g1[m]) and np.array_equal(d0[m], d1[m])
                          and np.array_equal(buf[:, :, zs], obuf[:, :, zs]))
                    if not ok:
                        bad += 1
                        print("FAIL", nlocal, le, npt, wrap, t)
        print("checked", nlocal, le)
    print("all identical to the oracle" if bad == 0 else "%d case(s) FAILED" % bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
