"""Golden logs for the BASELINE configurations the reference ships no log for, made by
running the UNMODIFIED reference executable (oracle/_ref/Ludwig_<model>.exe, built from
/root/reference by oracle/Makefile) on an input derived from one of its own regression
inputs. Run in the build container; the records go to tests/golden/named_size_logs.json
in the format of regression_logs.json (tests/test_regression_logs.py runs both).

SURVEY.md 8(c): "symmetric_lb + 3d_7pt_fluid (BASELINE config 3): no golden log -- generate
with the oracle build from tests/regression/d3q19-short/serial-spin-lb1.inp (:21 size, :48
fd_gradient_calculation) with the gradient key changed and size 128_128_128".
"""

import json
import os
import subprocess
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
import make_regression_golden as mrg      # noqa: E402  (the log parser)

REF = "/root/reference/tests/regression"
# name -> (source input, model, keys changed)
CASES = {
    "named-spin-lb1-7pt-128": ("d3q19-short/serial-spin-lb1", "d3q19",
                               {"size": "128_128_128", "fd_gradient_calculation": "3d_7pt_fluid",
                                "N_cycles": "100", "freq_statistics": "100", "grid": "1_1_1"}),
}


def read_input(path):
    inp = {}
    for line in open(path):
        line = line.strip()
        if not line or line.startswith("#"):
            continue
        parts = line.split()
        inp[parts[0]] = parts[1] if len(parts) > 1 else ""
    return inp


def main():
    out = {}
    for name, (src, model, changes) in CASES.items():
        inp = read_input(os.path.join(REF, src + ".inp"))
        inp.update(changes)
        exe = os.path.join(ROOT, "oracle", "_ref", "Ludwig_%s.exe" % model)
        with tempfile.TemporaryDirectory() as tmp:
            with open(os.path.join(tmp, "input"), "w") as fh:
                for k, v in inp.items():
                    fh.write("%s %s\n" % (k, v))
            t0 = time.time()
            env = dict(os.environ, OMP_NUM_THREADS=str(os.cpu_count() or 1))
            log = subprocess.run([exe], cwd=tmp, env=env, check=True, capture_output=True,
                                 text=True).stdout
            print("%s: %.1f s" % (name, time.time() - t0), file=sys.stderr)
        assert "Ludwig finished normally" in log, log[-2000:]
        lines = log.splitlines()
        rec = mrg.parse_block(lines)
        rec.pop("_start")
        rec["input"] = inp
        rec["nvel"] = {"d3q19": 19, "d3q27": 27, "d2q9": 9}[model]
        rec["source"] = ("oracle/_ref/Ludwig_%s.exe (unmodified reference) on %s.inp with %s"
                         % (model, src, ", ".join("%s %s" % kv for kv in changes.items())))
        out[name] = rec
    with open(os.path.join(HERE, "named_size_logs.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)
    print("wrote", len(out), "records")


if __name__ == "__main__":
    main()
