"""Multi-GPU parity (needs >= 2 visible GPUs; skipped otherwise): launches
tests/mgpu_worker.py under torch.distributed.run, one rank per GPU, NCCL halo exchange."""

import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def ngpu():
    import torch
    return torch.cuda.device_count()


def launch(n, extra, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node",
           str(n), "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(HERE, "mgpu_worker.py")] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    sys.stdout.write(r.stdout[-2000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("extra", [
    ["--nvel", "19", "--halo", "full"],
    ["--nvel", "19", "--halo", "reduced"],
    ["--nvel", "19", "--halo", "reduced", "--fused", "0"],
    ["--nvel", "19", "--halo", "reduced", "--overlap", "0", "--p2p", "0"],
    ["--nvel", "19", "--halo", "reduced", "--p2p", "0"],
    ["--nvel", "19", "--halo", "reduced", "--direct", "0"],
    ["--nvel", "19", "--halo", "full", "--overlap", "0", "--lsize", "8", "8", "8"],
    ["--nvel", "27", "--halo", "reduced", "--lsize", "8", "6", "10"],
    ["--nvel", "19", "--ndist", "2", "--halo", "full", "--lsize", "8", "8", "6"],
    ["--nvel", "9", "--halo", "reduced", "--lsize", "16", "12", "1"],
    # widened rows: fluctuations (decomposition-independent generators), the 27-point
    # gradient (corners of the phi halo across ranks), walls held by the edge ranks only
    ["--nvel", "19", "--halo", "reduced", "--kt", "1e-4"],
    ["--nvel", "19", "--halo", "reduced", "--kt", "1e-4", "--p2p", "0"],
    ["--nvel", "19", "--ndist", "2", "--halo", "full", "--lsize", "8", "8", "6", "--gradient", "27"],
    ["--nvel", "19", "--ndist", "2", "--halo", "full", "--lsize", "8", "8", "6", "--gradient", "27", "--kt", "1e-5"],
    ["--nvel", "19", "--halo", "reduced", "--walls", "1", "1", "0"],
    ["--nvel", "19", "--halo", "reduced", "--walls", "1", "0", "1", "--p2p", "0", "--kt", "1e-4"],
    ["--nvel", "9", "--halo", "reduced", "--lsize", "16", "12", "1", "--walls", "1", "1", "0"],
    # Lees-Edwards planes inside every rank's block (x and z decomposed)
    ["--nvel", "19", "--halo", "reduced", "--le", "1", "--steps", "9"],
    ["--nvel", "19", "--halo", "full", "--le", "1", "--p2p", "0", "--overlap", "0", "--steps", "9"],
    ["--nvel", "9", "--halo", "reduced", "--lsize", "16", "12", "1", "--le", "2", "--steps", "9"],
    # free_energy symmetric: the order parameter as a finite-difference field (phi, its
    # gradient and u cross the rank boundaries one site wide), with and without the
    # viscosity model
    ["--nvel", "19", "--halo", "reduced", "--fd", "1", "--gradient", "27", "--steps", "8"],
    ["--nvel", "19", "--halo", "reduced", "--fd", "2", "--gradient", "7", "--p2p", "0", "--steps", "8"],
    ["--nvel", "27", "--halo", "full", "--lsize", "8", "6", "10", "--fd", "2", "--gradient", "27", "--visc", "1", "--steps", "8"],
    # single-lattice streaming across ranks: the even sweep pushes into the neighbours' halo,
    # the odd sweep stores into their interior; looked at after an odd and an even sweep
    ["--nvel", "19", "--halo", "reduced", "--lattices", "1", "--steps", "6"],
    ["--nvel", "19", "--halo", "reduced", "--lattices", "1", "--steps", "5"],
    ["--nvel", "27", "--halo", "reduced", "--lattices", "1", "--steps", "7", "--lsize", "8", "6", "10"],
])
def test_decomposed_equals_single(n, extra):
    if ngpu() < n:
        pytest.skip("needs %d GPUs" % n)
    launch(n, extra, 29600 + n)


@pytest.mark.parametrize("extra", [
    # two ranks along the CONTIGUOUS direction: row-end halo sites written into the
    # neighbour's memory as whole 32-byte sectors; and along y
    ["--nvel", "19", "--halo", "reduced", "--grid", "1", "1", "2", "--lsize", "8", "6", "32"],
    ["--nvel", "19", "--halo", "reduced", "--grid", "1", "1", "2", "--lsize", "6", "5", "16", "--fused", "0"],
    ["--nvel", "19", "--ndist", "2", "--halo", "full", "--grid", "1", "1", "2", "--lsize", "8", "6", "32"],
    ["--nvel", "27", "--halo", "reduced", "--grid", "1", "1", "2", "--lsize", "6", "6", "48"],
    ["--nvel", "9", "--halo", "reduced", "--grid", "1", "2", "1", "--lsize", "16", "32", "1"],
    ["--nvel", "19", "--halo", "reduced", "--grid", "1", "2", "1", "--lsize", "8", "6", "32", "--kt", "1e-4"],
    ["--nvel", "19", "--halo", "reduced", "--grid", "1", "1", "2", "--lsize", "8", "6", "32", "--lattices", "1", "--steps", "6"],
    ["--nvel", "19", "--halo", "reduced", "--grid", "1", "1", "2", "--lsize", "7", "5", "16", "--lattices", "1", "--steps", "5"],
    ["--nvel", "19", "--halo", "reduced", "--grid", "1", "2", "1", "--lsize", "8", "6", "32", "--lattices", "1", "--steps", "7"],
    ["--nvel", "9", "--halo", "reduced", "--grid", "1", "2", "1", "--lsize", "16", "32", "1", "--lattices", "1", "--steps", "6"],
])
def test_two_ranks_along_other_axes(extra):
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    launch(2, extra, 29650)


@pytest.mark.parametrize("grid", [("1", "1", "2"), ("2", "1", "1"), ("1", "2", "1")])
def test_peer_memory_halo_two_processes_without_nccl(grid):
    """tools/p2p_two_process.py: two plain processes (no torch, no NCCL), CUDA IPC mappings
    exchanged over pipes, halo sites written into the neighbour's memory from the sweep;
    both blocks identical to the whole lattice on one GPU, bit for bit."""
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    tool = os.path.join(os.path.dirname(HERE), "tools", "p2p_two_process.py")
    r = subprocess.run([sys.executable, tool, "--grid"] + list(grid), capture_output=True, text=True,
                       timeout=300)
    sys.stdout.write(r.stdout[-1500:])
    assert r.returncode == 0 and "PASS" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
