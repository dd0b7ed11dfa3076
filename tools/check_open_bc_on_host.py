"""DEVELOPMENT TOOL (no GPU needed): steps through the source of open_bc_kernel
(ludwig_b200/csrc/open_bc_kernels.cuh) on the host and compares the lattice it leaves with
the oracle's inflow / outflow "rhou" (oracle/lb_oracle_wall.c: orc_open_apply, which
reproduces the reference's logs serial-open-ru1 and serial-open-ru2).

    python tools/check_open_bc_on_host.py
"""

import ctypes
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import lbinit                         # noqa: E402
from oracle import pyoracle           # noqa: E402

SO = os.path.join(ROOT, "tools", "_open_bc_emulate.so")


def main():
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-w", "-std=c++17", "-I/usr/local/cuda/include",
                           "-shared", "-fPIC", "-o", SO, os.path.join(ROOT, "tools", "open_bc_emulate.cpp")])
    E = ctypes.CDLL(SO)
    dp = ctypes.POINTER(ctypes.c_double)
    bad = 0
    for nvel in (19, 27):
        for nlocal in ((24, 12, 12), (12, 12, 24), (7, 9, 140)):
            for flow_dir in (0, 1, 2):
                orc = pyoracle.Oracle(nvel, nlocal)
                n = orc.nsite
                rng = np.random.default_rng(11)
                f0 = lbinit.make_state(orc.cv(), orc.wv(), nlocal)
                rho = 1.0 + 0.1*rng.random(n)
                u = 0.05*(rng.random((n, 3)) - 0.5)
                u0 = (0.003, -0.002, 0.001)
                o = orc.open_bc(flow_dir, u0, 1.01)
                fo, ro, uo = f0.copy(), rho.copy(), u.copy()
                orc.open_apply(o, ro, uo, fo)
                fe = f0.copy()
                cv = np.ascontiguousarray(orc.cv(), dtype=np.int32)
                wv = np.ascontiguousarray(orc.wv(), dtype=np.float64)
                rc = E.open_bc_emulate((ctypes.c_int*3)(*nlocal), nvel, cv.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                                       wv.ctypes.data_as(dp), flow_dir, (ctypes.c_double*3)(*u0), ctypes.c_double(1.01),
                                       rho.ctypes.data_as(dp), u.ctypes.data_as(dp), fe.ctypes.data_as(dp))
                assert rc == 0
                ok = np.array_equal(fe, fo) and not np.array_equal(fe, f0)
                print("nvel %d %s flow %d: %s (%d populations imposed)" % (nvel, nlocal, flow_dir, "ok" if ok else "DIFF",
                                                                          int((fe != f0).sum())))
                bad += (not ok)
    print("all identical to the oracle" if bad == 0 else "%d case(s) FAILED" % bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
