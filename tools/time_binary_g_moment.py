"""Binary-fluid time step (BASELINE.json configs[2]: D3Q19, ndist 2, symmetric free energy)
on one GPU with option "g_moment" off and on, without torch: CUDA-event times through the
library's own timer, final lattices compared bit for bit.

    python tools/time_binary_g_moment.py [n] [steps]
"""

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import ludwig_b200 as lb              # noqa: E402
from ludwig_b200 import models_host   # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    nt = (n, n, n)
    nall = tuple(x + 2 for x in nt)
    rng = np.random.default_rng(8361235)
    f = np.zeros(nall + (2, 19))
    w = models_host.wv(19)
    f[..., 0, :] = w
    f[..., 1, :] = w * (0.1 * (rng.random(nall) - 0.5))[..., None]
    f = np.ascontiguousarray(f.reshape(-1, 2, 19))
    ref = {}
    for gradient, lazy in ((7, 1), (7, 0), (27, 0)):
        for g in ((0, 1, 0, 1) if (gradient, lazy) == (7, 1) else (0, 1)):
            lat = lb.Lattice(19, nt, ndist=2)
            lat.set_option("halo", lb.HALO_REDUCED)
            lat.set_option("g_moment", g)
            lat.set_option("gradient", gradient)
            lat.set_option("grad_lazy", lazy)
            lat.set_param(eta_shear=0.00625, eta_bulk=0.00625, mobility=3.75, a=-0.00625,
                          b=0.00625, kappa=0.004)
            lat.lb_memcpy_h2d(f)
            lat.step(10)
            lat.sync()
            lat.timer_start()
            lat.step(steps)
            ms = lat.timer_stop()
            s = lat.stats()
            key = (gradient, lazy)
            same = ""
            if g == 0 and key not in ref:
                ref[key] = lat.lb_memcpy_d2h()
            elif g == 1 and key in ref and ref[key] is not None:
                same = ", lattice identical to g_moment 0: %s" % np.array_equal(ref[key], lat.lb_memcpy_d2h())
                ref[key] = None
            print("binary %d^3 gradient %d grad_lazy %d g_moment %d: %.4f ms/step, %.0f MLUPS, mass %.12e%s"
                  % (n, gradient, lazy, g, ms / steps, n**3 * steps / ms / 1e3, s[0], same), flush=True)
            lat.close()


if __name__ == "__main__":
    main()
