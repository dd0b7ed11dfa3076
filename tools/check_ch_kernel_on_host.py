"""DEVELOPMENT TOOL (no GPU in the build container): steps through the source of ch_kernel
on the host (tools/ch_emulate.cpp) and compares force and updated order parameter with
the reference-produced golden vectors tests/golden/ref_ch.npz, bit for bit -- with the
periodic wrap inside the kernel and with filled halo planes (what a decomposed direction
or the loopback exchange gives). Proves indexing and order of operations of the kernel
source before GPU time is spent on it; the GPU tests (tests/test_zz_gpu_cahn_hilliard.py)
are the parity evidence proper.

    python tools/check_ch_kernel_on_host.py
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


def same(a, b, exact):
    """exact build: bit for bit; fast build (fused multiply-add, central stress never
    formed): a few ulp of the largest value"""
    if exact:
        return np.array_equal(a, b)
    return np.max(np.abs(a - b)) <= 4e-16 * max(np.max(np.abs(b)), 1e-300) + 1e-19


def main(exact=1):
    E = ctypes.CDLL(build(exact))
    dp = ctypes.POINTER(ctypes.c_double)
    g = goldenlib.load_ch()
    P = goldenlib.CH_PHYS
    bad = 0
    for tag, (nlocal, order, method, npt, gmu) in sorted(goldenlib.CH_CASES.items()):
        orc = pyoracle.Oracle(19, nlocal)
        n = orc.nsite
        m = orc.interior_mask()
        for wrap in ((1, 1, 1), (0, 0, 0), (1, 0, 1)):
            if order == 3 and wrap != (1, 1, 1):
                continue          # refused by b200lb_cahn_hilliard
            phi, u = goldenlib.ch_fields(nlocal)
            orc.halo(phi, 1)
            grad, delsq = np.zeros((n, 3)), np.zeros(n)
            (orc.grad_27pt if npt == 27 else orc.grad_7pt)(phi, grad, delsq)
            if wrap != (1, 1, 1):
                orc.halo(grad, 3)
                orc.halo(delsq, 1)
            else:
                # the kernel must not depend on halo values around wrapped directions
                hal = ~m
                for a in (phi, delsq):
                    a[hal] = np.nan
                grad[hal] = np.nan
                u[hal] = np.nan
            force = np.zeros((n, 3))
            out = np.zeros(n)
            for fused in (1, 0):
                force[...] = 0.0
                args = ((ctypes.c_int * 3)(*nlocal), (ctypes.c_int * 3)(*wrap))
                prm = (ctypes.c_double * 4)(P["a"], P["b"], P["kappa"], P["mobility"])
                gm = (ctypes.c_double * 3)(*gmu)
                fc = (ctypes.c_double * 3)(0.0, 0.0, 0.0)
                ptrs = [a.ctypes.data_as(dp) for a in (phi, grad, delsq, u, force, out)]
                if fused:
                    rc = E.ch_emulate(*args, method, order, prm, gm, fc, 0, *ptrs)
                else:
                    rc = E.ch_emulate(*args, method, 0, prm, gm, fc, 0, *ptrs)
                    f1 = force.copy()
                    rc |= E.ch_emulate(*args, 0, order, prm, gm, fc, 0, *ptrs)
                    assert np.array_equal(f1[m], force[m])
                assert rc == 0
                okf = same(force[m], g[tag + "_force"], exact)
                okp = same(out[m], g[tag + "_phi1"], exact)
                if not exact:
                    # conservation is exact in either build: both sides of a face agree
                    okp = okp and abs(out[m].sum() - phi[m].sum()) < 1e-13 * m.sum()
                print("%-6s wrap %s %-6s force %s  phi %s" % (tag, wrap, "fused" if fused else "split",
                                                             "ok" if okf else "DIFF", "ok" if okp else "DIFF"))
                bad += (not okf) + (not okp)
            # accumulate onto an existing force field
            base = np.random.default_rng(1).standard_normal((n, 3)) * 1e-3
            force[...] = base
            E.ch_emulate((ctypes.c_int * 3)(*nlocal), (ctypes.c_int * 3)(*wrap), method, 0, prm, gm, fc, 1,
                         *[a.ctypes.data_as(dp) for a in (phi, grad, delsq, u, force, out)])
            if method == 2:
                ok = same(force[m], base[m] + g[tag + "_force"], exact)
                print("%-6s wrap %s accumulate %s" % (tag, wrap, "ok" if ok else "DIFF"))
                bad += (not ok)
    # fourth-order advection: no golden vector, the oracle (pinned against the reference
    # for this order, tests/test_oracle_vs_ref.py FD_CASES) stands in
    # ... and a lattice one site thick (wz = 0), third order
    for nlocal, npt, order4 in (((10, 8, 6), 7, 4), ((6, 7, 9), 27, 4), ((12, 9, 1), 27, 3)):
        orc = pyoracle.Oracle(19, nlocal)
        n = orc.nsite
        m = orc.interior_mask()
        prm_o = pyoracle.make_param(eta_shear=0.1, eta_bulk=0.1, mobility=P["mobility"], a=P["a"], b=P["b"],
                                    kappa=P["kappa"])
        ch = orc.ch(order4, 2, (1e-5, -2e-5, 3e-5))
        phi, u = goldenlib.ch_fields(nlocal)
        orc.halo(phi, 1)
        grad, delsq = np.zeros((n, 3)), np.zeros(n)
        (orc.grad_27pt if npt == 27 else orc.grad_7pt)(phi, grad, delsq)
        fo = np.zeros((n, 3))
        po = phi.copy()
        uo = u.copy()
        orc.phi_force(prm_o, ch, po, grad, delsq, fo)
        orc.cahn_hilliard(prm_o, ch, po, delsq, uo)
        force, out = np.zeros((n, 3)), np.zeros(n)
        rc = E.ch_emulate((ctypes.c_int * 3)(*nlocal), (ctypes.c_int * 3)(1, 1, 1), 2, order4,
                          (ctypes.c_double * 4)(P["a"], P["b"], P["kappa"], P["mobility"]),
                          (ctypes.c_double * 3)(1e-5, -2e-5, 3e-5), (ctypes.c_double * 3)(0.0, 0.0, 0.0), 0,
                          *[a.ctypes.data_as(dp) for a in (phi, grad, delsq, u, force, out)])
        assert rc == 0
        okf, okp = same(force[m], fo[m], exact), same(out[m], po[m], exact)
        okp = okp and abs(out[m].sum() - phi[m].sum()) < 1e-13 * m.sum()
        print("order %d %s %d-point force %s  phi %s" % (order4, nlocal, npt, "ok" if okf else "DIFF", "ok" if okp else "DIFF"))
        bad += (not okf) + (not okp)
    print("FAILED" if bad else ("exact build: all identical to the reference" if exact
                                else "fast build: all within a few ulp of the reference"))
    return bad


if __name__ == "__main__":
    sys.exit(1 if (main(1) + main(0)) else 0)
