"""The host-side C mirror of the reference call surface (ludwig_b200/host): builds
everywhere (CPU check), and its C test program -- test_prop.c patterns and the time loop
against the oracle, all through lb_collide / lb_halo / lb_propagation in C -- runs on a GPU."""

import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "host", "test_host_shim.exe")


def build_exe():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "ludwig_b200", "host")])
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    subprocess.check_call([
        "gcc", "-std=c99", "-O1", "-Wall", "-o", EXE,
        os.path.join(ROOT, "tests", "host", "test_host_shim.c"),
        "-L" + os.path.join(ROOT, "ludwig_b200"), "-lludwig_b200_host", "-lb200lb",
        "-L" + os.path.join(ROOT, "oracle"), "-llb_oracle", "-lm",
        "-Wl,-rpath," + os.path.join(ROOT, "ludwig_b200"),
        "-Wl,-rpath," + os.path.join(ROOT, "oracle")])


def test_host_mirror_builds_and_fails_loudly_without_gpu():
    import torch
    build_exe()
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: see test_host_mirror_on_gpu")
    r = subprocess.run([EXE], capture_output=True, text=True)
    assert r.returncode == 2 and "no CUDA device" in r.stdout   # no CPU fallback


@pytest.mark.gpu
def test_host_mirror_on_gpu():
    build_exe()
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=300)
    print(r.stdout)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all passed" in r.stdout


@pytest.mark.gpu
def test_host_mirror_cahn_hilliard_on_gpu():
    """"free_energy symmetric" through phi_force_calculation / phi_cahn_hilliard in C."""
    build_exe()
    r = subprocess.run([EXE, "fd"], capture_output=True, text=True, timeout=300)
    print(r.stdout)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all passed" in r.stdout
