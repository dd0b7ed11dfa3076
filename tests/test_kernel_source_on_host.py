"""The SOURCE of the CUDA kernels, stepped through on the host (no GPU): the headers
ludwig_b200/csrc/*.cuh are compiled with g++ behind thin shims for blockIdx / threadIdx /
__ldg / __dmul_rn (tools/*_emulate.cpp) and driven over the launch grids api.cu uses.

This is not the parity evidence proper -- that is tests/test_*gpu*.py on a B200 -- and
nothing in the product links it; it pins indexing, wrap addressing and order of operations
of the kernel source against the oracle / the reference's golden vectors in the CPU suite,
and it is what checked the kernels written after a round's GPU allocation was spent:

  ch_kernel                         vs tests/golden/ref_ch.npz (bit for bit; fma build: ulps)
  lb_sweep_kernel GM bit 3          vs the sweep that pulls g itself (bit for bit)
  le_phi_buffer / le_grad_fix       vs the oracle's field_leesedwards + stencils (bit for bit)
  ch_kernel + buffer kernels + ch_le_pass_a/_b (finite-difference order parameter across
  Lees-Edwards planes)              vs the oracle (phi bit for bit, force 1e-15)
  open_bc_kernel (inflow/outflow)   vs the oracle (bit for bit)
  lb_sweep_kernel WALL + BINARY     the propagation of both distributions with bounce-back
  grad27_solid_kernel               vs the oracle's 27-point stencil next to walls (bit for bit)
"""

import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOLS = os.path.join(ROOT, "tools")
CUDA_INC = "/usr/local/cuda/include"

pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(CUDA_INC, "cuda_runtime.h")),
                                reason="needs the CUDA headers (vector types) for g++")


def run_tool(name):
    r = subprocess.run([sys.executable, os.path.join(TOOLS, name)], capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def test_ch_kernel_source_matches_reference_golden():
    out = run_tool("check_ch_kernel_on_host.py")
    assert "exact build: all identical to the reference" in out
    assert "fast build: all within a few ulp of the reference" in out
    assert "DIFF" not in out


def test_binary_sweep_with_handed_over_flux_is_identical():
    exe = os.path.join(TOOLS, "_sweep_emulate")
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-w", "-std=c++17", "-I" + CUDA_INC,
                           "-o", exe, os.path.join(TOOLS, "sweep_emulate.cpp")])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
    assert "all cases identical" in r.stdout and "FAIL" not in r.stdout
    assert r.stdout.count("ok  ") == 70          # 54 flux hand-over cases + 16 wall propagations


def test_peer_memory_sweeps_identical_to_wrap_sweeps():
    """tools/push_emulate.cpp: the PUSH variants of the sweep with the lattice as its own
    neighbour, halo words / whole sectors / packed z-face planes, and zhalo_unpack_kernel."""
    exe = os.path.join(TOOLS, "_push_emulate")
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-w", "-std=c++17", "-I" + CUDA_INC,
                           "-o", exe, os.path.join(TOOLS, "push_emulate.cpp")])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
    assert "all cases identical" in r.stdout and "FAIL" not in r.stdout
    assert r.stdout.count("ok  ") == 330         # 264 three-step cases + 66 unpack cases
    assert r.stdout.count("packed planes") == 132  # both arrangements (option zpack = 1, 2)


def test_lees_edwards_order_parameter_kernels_match_oracle():
    out = run_tool("check_le_field_on_host.py")
    assert "all identical to the oracle" in out and "FAIL" not in out


def test_finite_difference_order_parameter_across_planes_matches_oracle():
    out = run_tool("check_ch_le_on_host.py")
    assert "exact build: order parameter identical to the oracle, force within 1e-15" in out
    assert "fast build: within a few ulp of the oracle" in out
    assert "DIFF" not in out


def test_open_boundary_kernel_matches_oracle():
    out = run_tool("check_open_bc_on_host.py")
    assert "all identical to the oracle" in out and "DIFF" not in out


def test_solid_gradient_kernel_matches_oracle():
    out = run_tool("check_solid_grad_on_host.py")
    assert "all identical to the oracle" in out and "DIFF" not in out
