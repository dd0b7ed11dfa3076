#!/usr/bin/env python
"""Turns the ncu captures of tools/ncu_traffic.sh (gpurun_out/ncu/*.csv) into
profiles/r2_ncu_traffic.json -- DRAM bytes per launch of the fused sweep for every named
shape, the figure bench.py reports as roofline.traffic -- and a readable summary,
profiles/r2_ncu_traffic.md, with the launch list of the default bench command."""

import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench      # noqa: E402  (traffic_key, B_ALG)

SRC = os.path.join(ROOT, "gpurun_out", "ncu")
CASES = {   # capture -> (nvel, ndist, local, hydro stored)
    "q19-256-ondemand": (19, 1, (256, 256, 256), False),
    "q19-256-store": (19, 1, (256, 256, 256), True),
    "q19-64-ondemand": (19, 1, (64, 64, 64), False),
    "q19-binary-128": (19, 2, (128, 128, 128), False),
    "q27-256-ondemand": (27, 1, (256, 256, 256), False),
    "q9-4096-ondemand": (9, 1, (4096, 4096, 1), False),
}


def rows(path):
    with open(path) as fh:
        lines = [l for l in fh if l.startswith('"')]
    return list(csv.DictReader(lines))


def launches(path):
    """[(kernel name, {metric: value})] in launch order."""
    out = collections.OrderedDict()
    for r in rows(path):
        key = int(r["ID"])
        out.setdefault(key, (r["Kernel Name"], {}))[1][r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    return list(out.values())


def main():
    table, md = {}, ["# ncu: DRAM traffic per launch of the dominant kernel (round 2, one B200)", "",
                     "`tools/ncu_traffic.sh` (`ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,"
                     "gpu__time_duration.sum --clock-control none`), converted by `tools/ncu_to_traffic.py`. "
                     "Durations under ncu are serialised and cold; traffic is what is taken from here.", "",
                     "| shape | kernel | launches | DRAM read | DRAM written | total per launch | algorithmic | ratio | ns under ncu |",
                     "|---|---|---|---|---|---|---|---|---|"]
    for name, (nvel, ndist, local, store) in CASES.items():
        ls = launches(os.path.join(SRC, name + ".csv"))
        # the fused pull+collide sweep: template arguments <M, VEC, PULL=1, COLLIDE=1, ...>
        fused = [(k, m) for k, m in ls if "lb_sweep_kernel" in k and k.split("<")[1].split(",")[2].strip() == "1"
                 and k.split("<")[1].split(",")[3].strip() == "1"]
        phi = [(k, m) for k, m in ls if "phi_from_g_kernel" in k]
        sites = float(local[0])*local[1]*local[2]
        for label, group, alg in (("fused sweep", fused, None), ("phi_from_g", phi, None)):
            if not group:
                continue
            rd = sum(m["dram__bytes_read.sum"] for _, m in group)/len(group)
            wr = sum(m["dram__bytes_write.sum"] for _, m in group)/len(group)
            ns = sum(m["gpu__time_duration.sum"] for _, m in group)/len(group)
            if label == "fused sweep":
                alg = (bench.B_ALG[nvel] if ndist == 1 else 3*bench.B_ALG[nvel]//2)*sites
                table[bench.traffic_key(nvel, ndist, local, store)] = {
                    "dram_bytes_per_launch": rd + wr, "dram_read": rd, "dram_write": wr,
                    "algorithmic_bytes_per_launch": alg, "kernel": group[0][0], "launches_averaged": len(group),
                    "capture": "gpurun_out/ncu/%s.csv (tools/ncu_traffic.sh)" % name}
            else:
                alg = (19*8 + 8 + 24)*sites       # g in, phi and its flux out
            md.append("| %s%s | %s | %d | %.3f GB | %.3f GB | %.3f GB | %.3f GB | %.3f | %.0f |"
                      % (name, "", label, len(group), rd/1e9, wr/1e9, (rd + wr)/1e9, alg/1e9, (rd + wr)/alg, ns))
    with open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json"), "w") as fh:
        json.dump(table, fh, indent=1, sort_keys=True)
    # launch list of the default bench command: share of the step per kernel
    ll = launches(os.path.join(SRC, "r2_launches_default.csv"))
    tot = collections.Counter()
    cnt = collections.Counter()
    for k, m in ll:
        short = k.split("(")[0].replace("void ", "")
        tot[short] += m["gpu__time_duration.sum"]
        cnt[short] += 1
    md += ["", "## Launch list of `python bench.py --steps 2 --warmup 1 --no-cpu` (every kernel of the run)", "",
           "| kernel | launches | total us | share |", "|---|---|---|---|"]
    allns = sum(tot.values())
    for k, v in tot.most_common(14):
        md.append("| `%s` | %d | %.1f | %.1f %% |" % (k[:110], cnt[k], v/1e3, 100.0*v/allns))
    with open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.md"), "w") as fh:
        fh.write("\n".join(md) + "\n")
    print("\n".join(md))


if __name__ == "__main__":
    main()
