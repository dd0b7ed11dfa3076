"""CPU: host-side logic added in round 2 -- the slab generator behind piecewise uploads,
bench.py's bookkeeping and its reference arm (the one leg that may execute oracle/), and
the anchors tools/make_ludwig_b200.py patches the reference at."""

import json
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.mark.parametrize("nvel,nlocal,ndist", [(19, (50, 45, 47), 2), (9, (90, 85, 1), 1), (27, (8, 40, 44), 1)])
def test_tiled_slabs_are_the_tiled_state(nvel, nlocal, ndist):
    from ludwig_b200 import models_host, synthetic
    cv, wv = models_host.cv(nvel), models_host.wv(nvel)
    ndim = 2 if nvel == 9 else 3
    whole = synthetic.tiled_state(cv, wv, nlocal, ndist=ndist, ndim=ndim)
    first_expected, parts = 0, []
    for first, part in synthetic.tiled_slabs(cv, wv, nlocal, ndist=ndist, ndim=ndim, planes=7):
        assert first == first_expected
        first_expected += part.shape[0]
        parts.append(part)
    assert first_expected == whole.shape[0]
    assert np.array_equal(np.concatenate(parts), whole)
    # halo sites are left to lb_halo; the interior repeats a block with coprime edges
    nall = tuple(n + 2 for n in nlocal)
    w = whole.reshape(nall + (ndist, nvel))
    assert not w[0].any() and not w[:, 0].any()
    if nlocal[0] > 37:
        assert np.array_equal(w[1], w[38])


def test_bench_bookkeeping():
    import bench
    assert bench.grid_for(8) == (2, 2, 2) and bench.grid_for(4) == (2, 2, 1) and bench.grid_for(2) == (2, 1, 1)
    assert bench.grid_for(3) is None

    class A:
        global_size, local, nvel = (1024, 1024, 1024), None, 19
    assert bench.local_size(A, (2, 1, 1)) == (512, 1024, 1024)
    A.global_size = None
    assert bench.local_size(A, (2, 2, 2)) == (256, 256, 256)
    A.nvel = 9
    assert bench.local_size(A, (1, 1, 1)) == (4096, 4096, 1)
    with pytest.raises(SystemExit):
        A.global_size = (100, 100, 100)
        bench.local_size(A, (3, 1, 1))
    # every named shape has a committed ncu figure, between 0.4 and 1.2 of the algorithmic bytes
    for nvel, ndist, local, store in [(19, 1, (256, 256, 256), False), (19, 1, (256, 256, 256), True),
                                      (19, 2, (128, 128, 128), False), (27, 1, (256, 256, 256), False),
                                      (9, 1, (4096, 4096, 1), False), (19, 1, (64, 64, 64), False)]:
        t = bench.ncu_traffic(nvel, ndist, local, store)
        alg = (bench.B_ALG[nvel] if ndist == 1 else 3*bench.B_ALG[nvel]//2)*float(np.prod(local))
        assert t is not None and 0.4 < t/alg < 1.2, (nvel, ndist, local, t/alg)
    r = bench.roofline(19, 1, (256, 256, 256), 256.0**3, 0.8, 0.79, 1.0, False)
    assert r["frac"] <= 1.0 and abs(r["achieved"] - 304*256.0**3/0.79e-3/1e9) < 1e-6
    assert r["bytes_per_site"] == 304 and r["traffic"] is not None


@pytest.mark.parametrize("extra,expect", [([], "D3Q19 16x16x16"), (["--ndist", "2"], "binary symmetric_lb 7-point"),
                                          (["--walls", "0", "1", "1"], None)])
def test_reference_arm(extra, expect):
    """--impl reference under a launcher's OMP_NUM_THREADS=1: all host cores, the named
    workload or an honest refusal, one JSON line."""
    env = dict(os.environ, OMP_NUM_THREADS="1")
    env.pop("B200LB_BENCH_REEXEC", None)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2",
                        "--warmup", "1", "--ref-size", "16", "16", "16"] + extra,
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference"
    if expect is None:
        assert "unavailable" in line and "walls" in line["unavailable"]
        return
    assert line["value"] > 0 and line["unit"] == "MLUPS" and line["e2e"]["value"] == line["value"]
    assert expect in line["cpu_baseline"]["sample"]
    ncores = len(os.sched_getaffinity(0))
    assert line["cpu_baseline"]["cores"] == ncores
    assert "%d OpenMP threads" % ncores in line["cpu_baseline"]["sample"]


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="the reference sources are not on this machine")
def test_dropin_patch_anchors(tmp_path):
    """Every anchor of tools/make_ludwig_b200.py matches the reference exactly once, and the
    patched files call the glue."""
    import make_ludwig_b200 as mk
    src = tmp_path / "src"
    src.mkdir()
    names = {n for n, _, _ in mk.AT_FUNCTION_HEAD} | {n for n, _, _ in mk.REPLACE}
    for n in names:
        shutil.copy(os.path.join("/root/reference/src", n), str(src / n))
        os.chmod(str(src / n), 0o644)
    touched = mk.patch(str(src))
    assert set(touched) == {"collision.c", "field.c", "field_grad.c", "hydro.c", "model.c",
                            "phi_lb_coupler.c", "propagation.c", "wall.c"}
    for n in touched:
        text = open(str(src / n)).read()
        assert '#include "lb_b200_glue.h"' in text and "lb_b200_" in text
    assert "LB_HALO_B200" in open(str(src / "lb_data_options.h")).read()
    assert "lb_halo_b200" in open(str(src / "distribution_rt.c")).read()
    assert os.path.exists(str(src / "lb_b200_glue.c"))
