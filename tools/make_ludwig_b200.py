#!/usr/bin/env python
"""Builds the REAL drop-in: a scratch copy of the reference (zazu29/ludwig v0.20.1,
/root/reference, never modified) whose hot-path functions forward to libb200lb.so when
the input file says `lb_halo_scheme lb_halo_b200`, linked into Ludwig.exe by the
reference's own Makefiles (serial build, gcc). Everything else -- ludwig.c's time loop, the
input parser, statistics, i/o -- is the reference's, unchanged.

    python tools/make_ludwig_b200.py [--models d3q19 d2q9] [--keep]

Output: ludwig_b200/host/_ludwig/Ludwig_<model>_b200.exe (git-ignored; travels to the GPU
box). The forwarding lines are inserted at anchors (the function heads below), so this file
holds no reference code; INTEGRATION.md shows the same lines in context. The glue they
call is ludwig_b200/host/dropin/lb_b200_glue.{h,c}.
"""

import argparse
import os
import re
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "ludwig_b200", "host", "_ludwig")
GLUE = os.path.join(ROOT, "ludwig_b200", "host", "dropin")
MODELS = {"d3q19": "-D_D3Q19_", "d3q27": "-D_D3Q27_", "d2q9": "-D_D2Q9_"}

INCLUDE = '#include "lb_b200_glue.h"\n'

# (file, regex matching the head of the function up to and including its opening brace,
#  lines inserted right behind it)
AT_FUNCTION_HEAD = [
    ("model.c", r"__host__ int lb_free\(lb_t \* lb\) \{\n",
     "  if (lb_b200_active()) lb_b200_free();\n"),
    ("model.c", r"__host__ int lb_memcpy\(lb_t \* lb, tdpMemcpyKind flag\) \{\n",
     "  if (lb_b200_active()) return lb_b200_memcpy(lb, flag);\n"),
    ("model.c", r"__host__ int lb_halo_swap\(lb_t \* lb, lb_halo_enum_t flag\) \{\n",
     "  if (lb_b200_active()) return lb_b200_halo(lb);\n"),
    ("collision.c", r"int lb_collide\(lb_t \* lb, hydro_t \* hydro, map_t \* map, noise_t \* noise,\n\s*fe_t \* fe, visc_t \* visc\) \{\n",
     "  if (lb_b200_active()) return lb_b200_collide(lb, hydro, noise, fe, visc);\n"),
    ("propagation.c", r"__host__ int lb_propagation\(lb_t \* lb\) \{\n",
     "  if (lb_b200_active()) return lb_b200_propagation(lb);\n"),
    ("hydro.c", r"__host__ int hydro_f_zero\(hydro_t \* obj, const double fzero\[NHDIM\]\) \{\n",
     "  if (lb_b200_active()) return lb_b200_hydro_f_zero(fzero);\n"),
    ("hydro.c", r"__host__ int hydro_u_zero\(hydro_t \* obj, const double uzero\[NHDIM\]\) \{\n",
     "  if (lb_b200_active()) return lb_b200_hydro_u_zero(uzero);\n"),
    ("hydro.c", r"__host__ int hydro_memcpy\(hydro_t \* obj, tdpMemcpyKind flag\) \{\n",
     "  if (lb_b200_active()) return lb_b200_hydro_memcpy(obj, flag);\n"),
    ("wall.c", r"__host__ int wall_bbl\(wall_t \* wall\) \{\n",
     "  if (lb_b200_active()) return lb_b200_wall_bbl(wall);\n"),
    ("wall.c", r"__host__ int wall_momentum\(wall_t \* wall, double f\[3\]\) \{\n",
     "  if (lb_b200_active()) return lb_b200_wall_momentum(wall, f);\n"),
    ("wall.c", r"__host__ int wall_set_wall_distributions\(wall_t \* wall\) \{\n",
     "  if (lb_b200_active()) return 0;    /* wall_bbl overwrites every population this sets */\n"),
    ("phi_lb_coupler.c", r"__host__ int phi_lb_to_field\(field_t \* phi, lb_t  \* lb\) \{\n",
     "  if (lb_b200_active()) return lb_b200_phi_lb_to_field(phi, lb);\n"),
    ("field.c", r"__host__ int field_halo\(field_t \* obj\) \{\n",
     "  if (lb_b200_is_phi(obj)) return lb_b200_field_halo(obj);\n"),
    ("field.c", r"__host__ int field_memcpy\(field_t \* obj, tdpMemcpyKind flag\) \{\n",
     "  if (lb_b200_is_phi(obj)) return lb_b200_field_memcpy(obj, flag);\n"),
    ("field_grad.c", r"int field_grad_compute\(field_grad_t \* obj\) \{\n",
     "  if (lb_b200_is_phi(obj->field)) return lb_b200_field_grad_compute(obj);\n"),
    ("field_grad.c", r"__host__ int field_grad_memcpy\(field_grad_t \* obj, tdpMemcpyKind flag\) \{\n",
     "  if (lb_b200_is_phi(obj->field)) return lb_b200_field_grad_memcpy(obj, flag);\n"),
]

# (file, exact text, replacement): the run-time switch and the object's creation
REPLACE = [
    # src/lb_data_options.h:26-28: one more halo scheme
    ("lb_data_options.h", "LB_HALO_OPENMP_REDUCED} lb_halo_enum_t;",
     "LB_HALO_OPENMP_REDUCED,\n                           LB_HALO_B200} lb_halo_enum_t;"),
    # src/lb_data_options.c:63-75: valid with two distributions as well
    ("lb_data_options.c", "opts->halo != LB_HALO_TARGET) valid = 0;",
     "opts->halo != LB_HALO_TARGET\n      && opts->halo != LB_HALO_B200) valid = 0;"),
    # src/distribution_rt.c:167-188: the key's new value
    ("distribution_rt.c", "    else if (havetype) {\n      pe_fatal(pe, \"lb_halo_scheme not recognised",
     "    else if (strcmp(htype, \"lb_halo_b200\") == 0) {\n      options.halo = LB_HALO_B200;\n    }\n"
     "    else if (havetype) {\n      pe_fatal(pe, \"lb_halo_scheme not recognised"),
    # src/wall.c:168 wall_commit: the walls the host has just laid out
    ("wall.c", "  wall_memcpy(wall, tdpMemcpyHostToDevice);\n\n  return 0;\n",
     "  wall_memcpy(wall, tdpMemcpyHostToDevice);\n  if (lb_b200_active()) lb_b200_wall_commit(wall);\n\n  return 0;\n"),
    # src/model.c:132 lb_data_create: the device object next to the host one
    ("model.c", "  lb_halo_create(obj, &obj->h, obj->haloscheme);\n",
     "  if (obj->haloscheme == LB_HALO_B200) {\n"
     "    lb_b200_create(pe, cs, obj->ndim, obj->nvel, obj->ndist);\n  }\n"
     "  lb_halo_create(obj, &obj->h, obj->haloscheme);\n"),
]


def patch(src):
    touched = set()
    for name, pattern, text in AT_FUNCTION_HEAD:
        path = os.path.join(src, name)
        s = open(path).read()
        m = list(re.finditer(pattern, s))
        if len(m) != 1:
            raise SystemExit("anchor for %s matched %d times in %s" % (text.strip(), len(m), name))
        s = s[:m[0].end()] + "\n" + text + s[m[0].end():]
        open(path, "w").write(s)
        touched.add(name)
    for name, old, new in REPLACE:
        path = os.path.join(src, name)
        s = open(path).read()
        if s.count(old) != 1:
            raise SystemExit("text to replace found %d times in %s: %r" % (s.count(old), name, old))
        open(path, "w").write(s.replace(old, new))
        if name.endswith(".c") and "lb_b200_" in new:
            touched.add(name)
    for name in sorted(touched):
        path = os.path.join(src, name)
        s = open(path).read()
        i = s.index('#include "')
        open(path, "w").write(s[:i] + INCLUDE + s[i:])
    for f in ("lb_b200_glue.h", "lb_b200_glue.c"):
        shutil.copy(os.path.join(GLUE, f), os.path.join(src, f))
    return sorted(touched)


def build(model, keep):
    work = tempfile.mkdtemp(prefix="ludwig_b200_%s_" % model)
    top = os.path.join(work, "ludwig")
    shutil.copytree(REF, top, symlinks=True,
                    ignore=shutil.ignore_patterns(".git", "docs", "tests", "util"))
    subprocess.check_call(["chmod", "-R", "u+w", top])
    touched = patch(os.path.join(top, "src"))
    libdir = os.path.join(ROOT, "ludwig_b200")
    with open(os.path.join(top, "config.mk"), "w") as fh:
        # the reference's serial gcc build (config/unix-gcc-default.mk) with
        # OpenMP; libb200lb.so rides in on the library line, found at run time relative to
        # the executable (ludwig_b200/host/_ludwig -> ludwig_b200)
        fh.write("BUILD=serial\nMODEL=%s\nCC=gcc -fopenmp\n" % MODELS[model])
        fh.write("CFLAGS=-O2 -g -Wall -I%s\n" % os.path.join(ROOT, "include"))
        fh.write("AR=ar\nARFLAGS=-cr\nLDFLAGS=-fopenmp\n")
        fh.write("MPI_INC_PATH=./mpi_s\nMPI_LIB_PATH=./mpi_s\n")
        fh.write("MPI_LIB=-lmpi\n")
        fh.write("LAUNCH_SERIAL_CMD=\nLAUNCH_MPIRUN_CMD=\n")
    ncpu = str(os.cpu_count() or 4)
    log = open(os.path.join(work, "build.log"), "w")
    # Makefile.mk fixes the MPI library of a serial build itself: the whole list goes on the
    # command line, the reference's own (src/Makefile:24) plus libb200lb.so
    libs = ("LIBS=-L../mpi_s -lmpi ../target/libtarget.a -lm -L%s -lb200lb "
            "-Wl,-rpath,'$$ORIGIN/../..'" % libdir)
    for cmd in (["make", "serial"], ["make", "-C", "target"], ["make", "-j", ncpu, "-C", "src", libs]):
        r = subprocess.run(cmd, cwd=top, stdout=log, stderr=subprocess.STDOUT)
        if r.returncode != 0:
            log.flush()
            sys.stderr.write(open(os.path.join(work, "build.log")).read()[-4000:])
            raise SystemExit("build failed: %s (tree kept in %s)" % (" ".join(cmd), work))
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, "Ludwig_%s_b200.exe" % model)
    shutil.copy(os.path.join(top, "src", "Ludwig.exe"), exe)
    print("%s: patched %s -> %s" % (model, ", ".join(touched), os.path.relpath(exe, ROOT)))
    if keep:
        print("scratch tree kept in", work)
    else:
        shutil.rmtree(work)
    return exe


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--models", nargs="+", default=["d3q19", "d2q9"], choices=sorted(MODELS))
    ap.add_argument("--keep", action="store_true", help="keep the scratch tree")
    args = ap.parse_args()
    if not os.path.isdir(os.path.join(REF, "src")):
        raise SystemExit("the reference sources are not here (%s): nothing to build; "
                         "the GPU box uses the prebuilt executables" % REF)
    if not os.path.exists(os.path.join(ROOT, "ludwig_b200", "libb200lb.so")):
        raise SystemExit("build libb200lb.so first (make -C ludwig_b200/csrc)")
    for model in args.models:
        build(model, args.keep)


if __name__ == "__main__":
    main()
