"""CPU: a selection of the `gpu` tests run against the shipped kernels compiled for the host (tests/emu/README.md).

The build container has no GPU; this keeps the index logic, the barrier / mbarrier protocol and the arithmetic order of the CUDA
sources under test in the CPU suite: tendencies, halo fills and substeps bit for bit against the oracle for every advection /
topology / closure configuration, the fused sweeps (tile and TMA plane sweep) against each other and against the complete path, the
reference's halo test and the checkpoint → pickup tests.  The emulated library is test infrastructure: it is built into
tests/emu/_build/, loaded only when NHSTEP_EMU=1 is set (here, for the child process), and never by the product.
"""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SELECTION = ("tendencies_and_halos_bit_exact or substeps_and_normal or fused_fast_path_is_bit or "
             "(plane_marching and size0 and RungeKutta3) or halo_regions or checkpointer or weno_grid_tendencies")


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++ to compile the kernels for the host")
def test_gpu_tests_pass_against_the_kernels_compiled_for_the_host():
    env = dict(os.environ, NHSTEP_EMU="1")
    for k in ("NHSTEP_EMU_GUARD", "NHSTEP_EMU_ORDER", "NHSTEP_EMU_POISON", "NHSTEP_TILE_SWEEP", "NHSTEP_PLANE_SWEEP"):
        env.pop(k, None)
    cmd = [sys.executable, "-m", "pytest", "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider", "-k", SELECTION,
           "tests/test_gpu_parity.py", "tests/test_reference_halo_tests.py", "tests/test_zy_checkpoint_gpu.py",
           "tests/test_reference_weno_grid.py"]
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=1500)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert " passed" in r.stdout and "failed" not in r.stdout.splitlines()[-1], tail


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++ to compile the kernels for the host")
def test_multi_rank_slab_tests_pass_against_the_kernels_compiled_for_the_host():
    """Two and four rank PROCESSES on host cores (the deployment shape: one process per GPU): halo rows and FFT transpose blocks
    stored into the peers' buffers (files under /dev/shm standing in for CUDA IPC over NVLink), the flag barrier between them."""
    env = dict(os.environ, NHSTEP_EMU="1")
    for k in ("NHSTEP_EMU_GUARD", "NHSTEP_EMU_ORDER", "NHSTEP_EMU_POISON", "NHSTEP_TILE_SWEEP", "NHSTEP_PLANE_SWEEP", "NHSTEP_OVERLAP"):
        env.pop(k, None)
    selection = ("test_multi_gpu_matches_single_gpu and float64-RungeKutta3-size0-2-p2p or "
                 "test_transpose_round_trip_is_bit_identical and (2-p2p or 4-p2p or 2-nccl) or "
                 "test_x_slabs_match_single_gpu and float64-RungeKutta3-2-p2p or "
                 "test_complete_path_x_slabs_match_single_gpu and ppb-les-2")
    cmd = [sys.executable, "-m", "pytest", "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider", "-k", selection,
           "tests/test_gpu_distributed.py", "tests/test_zx_slab_partition.py"]
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=1500)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert " passed" in r.stdout and "failed" not in r.stdout.splitlines()[-1], tail
