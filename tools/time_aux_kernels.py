"""Times the order-parameter kernels alone on one GPU (CUDA events on the handle's
stream): phi from g, 7-point and 27-point gradient, at a given binary-fluid lattice.
Prints one JSON line per kernel with the algorithmic bytes per site and GB/s.

    python tools/time_aux_kernels.py [--local 256 256 256] [--reps 20]
"""

import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import ludwig_b200 as lb  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--local", type=int, nargs=3, default=[256, 256, 256])
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    local = tuple(args.local)
    lat = lb.Lattice(19, local, ndist=2)
    lat.set_option("grad_lazy", 0)            # the 7-point kernel itself, not the in-sweep stencil
    rng = np.random.default_rng(0)
    phi = rng.standard_normal(lat.nsite)
    lat.field_h2d(lb.FIELD_PHI, phi)
    lat.field_halo()
    sites = float(np.prod(local))
    peak = None
    try:
        with open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) as fh:
            peak = json.load(fh)
    except Exception:
        pass
    for name, npt, nbytes, direct in (("grad7_kernel", 7, 40, 1), ("grad27_direct_kernel", 27, 40, 1),
                                      ("grad27_kernel (planes staged in shared memory)", 27, 40, 0)):
        lat.set_option("grad_direct", direct)
        lat.field_grad_set(npt)
        for _ in range(3):
            lat.field_grad_compute()
        lat.sync()
        lat.timer_start()
        for _ in range(args.reps):
            lat.field_grad_compute()
        ms = lat.timer_stop()/args.reps
        print(json.dumps({"kernel": name, "local": list(local), "ms": ms,
                          "bytes_per_site": nbytes,
                          "GBps": nbytes*sites/(ms*1e-3)/1e9,
                          "what": "8 B phi read + 32 B grad,delsq written per site"}), flush=True)
    lat.close()


if __name__ == "__main__":
    main()
