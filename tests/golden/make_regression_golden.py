"""Extracts the statistics block printed after the last time step from the reference's own
regression golden logs (tests/regression/d3q19-short/serial-dist-*.log, d2q9/serial-2khz-bm1.log
the fluctuating-hydrodynamics case d2q9/serial-flct-t01.log and the binary-fluid spinodal
cases d3q19-short/serial-spin-lb{1,3,4}.log (symmetric_lb, 27-point gradient, lb3/lb4 with
fluctuations, lb1/lb4 with ghost modes off), and the wall cases d3q19-short/serial-rect-ct1
(body-force driven duct flow between y and z walls) and serial-wall-st1 (closed box with
fluctuations), the finite-difference order-parameter cases (free_energy symmetric)
serial-spin-fd1, serial-symm-dr1, serial-symm-dr2 (viscosity model), serial-le2d-lb1 (symmetric_lb sheared by two Lees-Edwards planes), serial-symm-pat, serial-muex-st1 and the 400-step
d3q19/pmpi08-spin-fd1, serial-rect-ct2..5 (the duct with partial-slip walls), and the 400-step, eight-rank spinodal case d3q19/pmpi08-spin-lb1 (its log does
not depend on the decomposition), Ludwig v0.20.1)
into tests/golden/regression_logs.json, together with the input keys that define each
case. Run in the build container (needs /root/reference)."""

import json
import os
import re

REF = "/root/reference/tests/regression"
HERE = os.path.dirname(os.path.abspath(__file__))
CASES = ["d3q19-short/serial-dist-3du", "d3q19-short/serial-dist-1dp",
         "d3q19-short/serial-dist-2kh", "d3q19-short/serial-dist-2sw", "d2q9/serial-2khz-bm1",
         "d2q9/serial-flct-t01", "d3q19-short/serial-spin-lb1", "d3q19-short/serial-spin-lb3",
         "d3q19-short/serial-spin-lb4", "d3q19-short/serial-rect-ct1",
         "d3q19-short/serial-wall-st1", "d3q19/pmpi08-spin-lb1",
         "d3q19-short/serial-rect-ct2", "d3q19-short/serial-rect-ct3",
         "d3q19-short/serial-rect-ct4", "d3q19-short/serial-rect-ct5",
         # free_energy symmetric: the order parameter as a finite-difference field
         "d3q19-short/serial-spin-fd1", "d3q19-short/serial-symm-dr1",
         "d3q19-short/serial-symm-pat", "d3q19-short/serial-muex-st1", "d3q19/pmpi08-spin-fd1",
         # ... with viscosity_model arrhenius (the collision reads a per-site viscosity)
         "d3q19-short/serial-symm-dr2",
         # binary fluid across two Lees-Edwards planes (a sheared 64 x 64 x 1 system)
         "d3q19-short/serial-le2d-lb1",
         # the finite-difference order parameter sheared by two planes: 7- and 27-point
         # gradients, advection orders 1-4, and the two-dimensional case
         "d3q19-short/serial-le3d-st1", "d3q19-short/serial-le3d-st2",
         "d3q19-short/serial-le3d-st3", "d3q19-short/serial-le3d-st4",
         "d3q19-short/serial-le3d-st5", "d3q19-short/serial-le3d-st6",
         "d3q19-short/serial-le3d-st7", "d3q19-short/serial-le3d-st8",
         "d3q19-short/serial-le2d-fd1",
         # the eight-rank long runs of the same (logs do not depend on the decomposition)
         "d3q19/pmpi08-le2d-fd1", "d3q19/pmpi08-le2d-lb1", "d3q19/pmpi08-le3d-fd1",
         "d3q19/pmpi08-le3d-lb1",
         # a drop relaxing for 20 steps (the run the reference's i/o tests restart from)
         "d3q19-io/iodrop-mpi1-io1",
         # open boundaries (inflow / outflow "rhou") along the one direction without walls
         "d3q19-short/serial-open-ru1", "d3q19-short/serial-open-ru2",
         # the binary fluid in a closed box (walls on all sides, 27-point stencil next to solid)
         "d3q19-short/serial-spin-lb2",
         # the finite-difference order parameter between two walls (solid stencil, no-flux
         # faces, one-sided stress at the walls)
         "d3q19-short/serial-spin-fd2"]
NVEL = {"d3q19-short": 19, "d3q19": 19, "d2q9": 9, "d3q19-io": 19}


def floats(line):
    return [float(x) for x in re.findall(r"[-+]?\d+\.\d+(?:[eE][-+]?\d+)?", line)]


def parse_block(lines):
    """The statistics block printed after the last time step of a Ludwig log."""
    start = max(i for i, l in enumerate(lines) if l.startswith("Scalars - total"))
    blk = lines[start:start + 20]
    rec = {"_start": start}
    for l in blk:
        if l.startswith("[rho]"):
            rec["rho"] = floats(l)                 # total mean variance min max
        if l.startswith("[phi]"):
            rec["phi"] = floats(l)                 # total mean variance min max
        if l.startswith("[fe]"):
            rec["fe"] = floats(l)                  # with walls: f, v, f/v, f_s1, f_s2
        if l.startswith("[fed]"):
            rec["fed"] = floats(l)                 # total/volume, fluid mean
        if l.startswith("[total   ]"):
            rec["momentum"] = floats(l)
        if l.startswith("[fluid   ]"):
            rec["momentum_fluid"] = floats(l)
        if l.startswith("[walls   ]"):
            rec["momentum_walls"] = floats(l)
        if l.startswith("[vol flux]"):
            rec["vol_flux"] = floats(l)
        if l.startswith("[minimum ]"):
            rec["umin"] = floats(l)
        if l.startswith("[maximum ]"):
            rec["umax"] = floats(l)
        if l.startswith("[eqipart.]"):
            rec["eqipart"] = floats(l)             # isothermal fluctuations
        if l.startswith("[measd/kT]"):
            rec["measd_kt"] = floats(l)
    return rec


def main():
    out = {}
    for case in CASES:
        inp = {}
        for line in open(os.path.join(REF, case + ".inp")):
            line = line.strip()
            if not line or line.startswith("#"):
                continue
            parts = line.split()
            inp[parts[0]] = parts[1] if len(parts) > 1 else ""
        lines = open(os.path.join(REF, case + ".log")).read().splitlines()
        rec = parse_block(lines)
        start = rec.pop("_start")
        rec["input"] = inp
        rec["nvel"] = NVEL[case.split("/")[0]]
        rec["source"] = "tests/regression/%s.log:%d-%d" % (case, start + 1, start + 10)
        out[case.split('/')[1]] = rec
    with open(os.path.join(HERE, "regression_logs.json"), "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out, indent=1)[:600])


if __name__ == "__main__":
    main()
