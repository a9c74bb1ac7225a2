"""Host logic of the y-slab multi-GPU path on CPU: world_size-2/3 ``gloo`` process groups.

What is covered here without a GPU: the neighbour ordering of the halo exchange, and the blocked spectral layout +
all-to-all data flow of the distributed Poisson solve (a NumPy stand-in plays the role of the FFT kernels, using the
same index helpers as csrc/dist.cuh) against the oracle's FFTBasedPoissonSolver on the global problem.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oceananigans_jl_b200 as oc                      # noqa: E402
from oceananigans_jl_b200 import distributed as D      # noqa: E402
import oracle as o                                     # noqa: E402


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(fn, world, *args):
    port = _free_port()
    mp.spawn(_entry, args=(world, port, fn, args), nprocs=world, join=True)


def _entry(rank, world, port, fn, args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fn(rank, world, *args)
    finally:
        dist.destroy_process_group()


def test_partition_arithmetic():
    assert D.local_slab(64, 4, 2) == (32, 16)
    with pytest.raises(ValueError):
        D.local_slab(30, 4, 0)
    assert D.neighbours(0, 4) == (3, 1) and D.neighbours(3, 4) == (2, 0) and D.neighbours(0, 2) == (1, 1)
    for Nx in (16, 64, 256, 1024):
        for R in (1, 2, 4, 8):
            Cx = D.spectral_block_width(Nx, R)
            assert Cx % 8 == 0 and Cx * R >= Nx // 2 + 1 and (Cx - 8) * R < Nx // 2 + 1 + 8 * R
    # Partition(R) / Partition(x=R): x-slabs (run as the mirror image); Partition(y=R), Partition(1, R), Partition(): y-slabs
    assert (D.Partition(2).direction, D.Partition(x=2).ranks, D.Partition(x=1).direction) == ("x", 2, "x")
    assert (D.Partition(y=4).direction, D.Partition(1, 4).direction, D.Partition().direction, D.Partition(y=1).direction) == ("y",) * 4
    with pytest.raises(NotImplementedError):
        D.Partition(x=2, y=2)                      # pencils
    with pytest.raises(NotImplementedError):
        D.Partition(z=2)


def test_blocked_layout_is_a_bijection():
    Nx, NyL, Nz, R = 32, 4, 8, 2
    Cx = D.spectral_block_width(Nx, R)
    seen = set()
    for kx in range(Nx // 2 + 1):
        for jl in range(NyL):
            for kz in range(Nz):
                off = D.blocked_index(kx, jl, kz, Cx, NyL, Nz)
                assert 0 <= off < R * Cx * NyL * Nz
                seen.add(off)
                # after the all-to-all, block d of the sender (rank s) is block s of rank d
                d, inner = divmod(off, Cx * NyL * Nz)
                assert d == kx // Cx
                for s in range(R):
                    assert D.transposed_index(kx % Cx, s * NyL + jl, kz, Cx, NyL, Nz) == inner + s * Cx * NyL * Nz
    assert len(seen) == (Nx // 2 + 1) * NyL * Nz


def _halo_worker(rank, world):
    n = 5
    lo = torch.full((n,), 10.0 * rank + 1)       # my first rows
    hi = torch.full((n,), 10.0 * rank + 2)       # my last rows
    rlo, rhi = torch.zeros(n), torch.zeros(n)
    D.exchange_y_halos(lo, hi, rlo, rhi, rank, world)
    prev, nxt = D.neighbours(rank, world)
    assert torch.all(rlo == 10.0 * prev + 2), "lower halo must hold the previous rank's last rows"
    assert torch.all(rhi == 10.0 * nxt + 1), "upper halo must hold the next rank's first rows"


@pytest.mark.parametrize("world", [2, 3])
def test_halo_exchange_neighbour_order(world):
    _run(_halo_worker, world)


def test_single_rank_messages_are_copies():
    lo, hi = torch.arange(4.0), torch.arange(4.0) + 10
    rlo, rhi = torch.zeros(4), torch.zeros(4)
    D.exchange_y_halos(lo, hi, rlo, rhi, 0, 1)
    assert torch.equal(rhi, lo) and torch.equal(rlo, hi)
    a, b = torch.arange(6.0), torch.zeros(6)
    D.transpose_blocks(b, a, 1)
    assert torch.equal(a, b)


def _poisson_worker(rank, world, Nx, Ny, Nz, seed):
    """Distributed solve with NumPy FFTs in the library's buffer layouts; compare with the oracle on the global grid."""
    rng = np.random.default_rng(seed)
    rhs = rng.standard_normal((Nx, Ny, Nz))
    rhs -= rhs.mean()
    grid = o.RectilinearGrid(np.float64, size=(Nx, Ny, Nz), x=(0, 1.0), y=(0, 2.0), z=(0, 0.5),
                             topology=(o.Periodic, o.Periodic, o.Periodic))
    solver = o.FFTBasedPoissonSolver(grid)
    expect = solver.solve(b=rhs).copy()
    lx, ly, lz = solver.lam

    j0, NyL = D.local_slab(Ny, world, rank)
    Cx = D.spectral_block_width(Nx, world)
    M = Nx // 2 + 1
    # phase 0: x r2c and z c2c on the local rows, stored in the blocked layout (what pf_x_forward + pd_z_kernel write)
    loc = np.fft.fft(np.fft.rfft(rhs[:, j0:j0 + NyL, :], axis=0), axis=2)
    A = np.zeros(world * Cx * NyL * Nz, dtype=np.complex128)
    kx, jl, kz = np.meshgrid(np.arange(M), np.arange(NyL), np.arange(Nz), indexing="ij")
    A[D.blocked_index(kx, jl, kz, Cx, NyL, Nz)] = loc
    A, B = torch.from_numpy(A), torch.zeros(world * Cx * NyL * Nz, dtype=torch.complex128)
    D.transpose_blocks(B, A, world)
    # phase 1: y c2c, division, inverse y c2c on this rank's columns (what pd_y_solve_kernel does)
    Bn = B.numpy()
    kxl, j, kz = np.meshgrid(np.arange(Cx), np.arange(Ny), np.arange(Nz), indexing="ij")
    tix = D.transposed_index(kxl, j, kz, Cx, NyL, Nz)
    cols = Bn[tix]
    kxg = rank * Cx + np.arange(Cx)
    active = kxg <= Nx // 2
    spec = np.fft.fft(cols, axis=1)
    lam = np.where(active, lx[np.minimum(kxg, Nx - 1)], 1.0)[:, None, None] + ly[None, :, None] + lz[None, None, :]
    with np.errstate(divide="ignore", invalid="ignore"):
        spec = -spec / lam
    if rank == 0:
        spec[0, 0, 0] = 0
    spec[~active] = 0
    Bn[tix] = np.fft.ifft(spec, axis=1)
    D.transpose_blocks(A, B, world)
    # phase 2: inverse z c2c, inverse x c2r
    An = A.numpy()
    kx, jl, kz = np.meshgrid(np.arange(M), np.arange(NyL), np.arange(Nz), indexing="ij")
    loc = An[D.blocked_index(kx, jl, kz, Cx, NyL, Nz)]
    phi = np.fft.irfft(np.fft.ifft(loc, axis=2), n=Nx, axis=0)
    err = np.abs(phi - expect[:, j0:j0 + NyL, :]).max() / np.abs(expect).max()
    assert err < 1e-12, err


@pytest.mark.parametrize("world,size", [(2, (16, 16, 8)), (2, (32, 8, 16)), (3, (16, 12, 8))])
def test_distributed_poisson_data_flow_matches_oracle(world, size):
    _run(_poisson_worker, world, *size, 7)
