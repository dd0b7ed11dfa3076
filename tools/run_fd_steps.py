"""A few time steps of the finite-difference order-parameter workload (free_energy
symmetric) on one GPU, without torch: the command ncu wraps for profiles/*_fd_*.csv.

    python tools/run_fd_steps.py [n] [order] [gradient points] [steps] [ch_planes] [ch_exact]
"""

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import ludwig_b200 as lb              # noqa: E402
from ludwig_b200 import models_host   # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    order = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    npt = int(sys.argv[3]) if len(sys.argv) > 3 else 27
    steps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
    planes = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    exact = int(sys.argv[6]) if len(sys.argv) > 6 else 0
    nt = (n, n, n)
    lat = lb.Lattice(19, nt)
    lat.set_option("halo", lb.HALO_REDUCED)
    lat.set_param(eta_shear=0.00625, eta_bulk=0.00625, mobility=1.25, a=-0.00625, b=0.00625,
                  kappa=0.004)
    lat.phi_ch_create(order, lb.FORCE_STRESS_DIVERGENCE)
    lat.field_grad_set(npt)
    lat.set_option("ch_exact", exact)
    if planes:
        lat.set_option("ch_planes", planes)
    nall = tuple(x + 2 for x in nt)
    f = np.zeros(nall + (1, 19))
    f[1:-1, 1:-1, 1:-1, 0, :] = models_host.wv(19)
    lat.lb_memcpy_h2d(np.ascontiguousarray(f.reshape(-1, 1, 19)))
    del f
    phi = np.zeros(nall)
    phi[1:-1, 1:-1, 1:-1] = 0.1 * (np.random.default_rng(8361235).random(nt) - 0.5)
    lat.field_h2d(lb.FIELD_PHI, np.ascontiguousarray(phi.reshape(-1)))
    lat.step(3)
    lat.sync()
    lat.timer_start()
    lat.step(steps)
    ms = lat.timer_stop()
    s = lat.field_stats(lb.FIELD_PHI)[0]
    # the order-parameter passes on their own
    lat.timer_start()
    for _ in range(steps):
        lat.field_halo()
        lat.field_grad_compute()
        lat.phi_force_calculation()
        lat.phi_cahn_hilliard()
    ms_fd = lat.timer_stop()
    print("fd order %d, %d-point gradient, %d^3, ch_planes %d, ch_exact %d: %.3f ms/step, %.0f MLUPS, "
          "gradient + force/update alone %.3f ms, sum phi %.10e"
          % (order, npt, n, lat.get_option("ch_planes"), lat.get_option("ch_exact"), ms / steps, n**3 * steps / ms / 1e3,
             ms_fd / steps, s[0]))
    lat.close()


if __name__ == "__main__":
    main()
