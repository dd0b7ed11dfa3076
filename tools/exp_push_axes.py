"""Development tool: cost of the peer-memory halo push per decomposed axis, measured on ONE
GPU with option loopback (the GPU is its own neighbour, so NVLink is not involved: what is
left is the instruction, launch and synchronisation overhead of the protocol).

    python tools/exp_push_axes.py [--size 256] [--steps 100]
"""

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import ludwig_b200 as lb                       # noqa: E402
from ludwig_b200 import synthetic             # noqa: E402
from ludwig_b200 import models_host           # noqa: E402


def run(nvel, local, mask, p2p, steps, hydro_write, f0, zpack=1):
    lat = lb.Lattice(nvel, local)
    lat.set_option("halo", lb.HALO_REDUCED)
    lat.set_option("hydro_write", hydro_write)
    lat.set_option("zpack", zpack)
    lat.set_param(eta_shear=0.1, eta_bulk=0.1)
    if mask:
        lat.set_option("loopback", 8 | mask)
        if p2p:
            lat.p2p_import(None, 1)
        else:
            lat.set_option("p2p", 0)
    lat.lb_memcpy_h2d(f0)
    lat.step(5)
    lat.sync()
    l0 = lat.launch_count()
    lat.timer_start()
    lat.step(steps)
    ms = lat.timer_stop() / steps
    nl = (lat.launch_count() - l0) / steps
    sites = float(local[0]) * local[1] * local[2]
    st = lat.stats()
    lat.close()
    return {"mask": mask, "p2p": p2p, "zpack": zpack if (p2p and mask & 4) else None,
            "ms_per_step": ms, "mlups": sites / ms / 1e3,
            "launches_per_step": nl, "mass": st[0]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--nvel", type=int, default=19)
    ap.add_argument("--hydro-write", type=int, default=0)
    args = ap.parse_args()
    local = (args.size,) * 3
    cv, wv = models_host.cv(args.nvel), models_host.wv(args.nvel)
    f0 = synthetic.make_state(cv, wv, local)
    # z faces and all three axes: row-end halo words (zpack 0) against the packed planes, one
    # word (1) or one sector (2) per column
    cases = [(0, 0, 0), (1, 1, 0), (2, 1, 0)]
    cases += [(4, 1, z) for z in (0, 1, 2)]
    cases += [(7, 1, z) for z in (0, 1, 2)]
    for mask, p2p, zpack in cases:
        print(json.dumps(run(args.nvel, local, mask, p2p, args.steps, args.hydro_write, f0, zpack)),
              flush=True)

if __name__ == "__main__":
    main()
