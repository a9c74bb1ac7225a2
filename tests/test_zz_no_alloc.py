"""GPU: nh_tridiag_solve_static (the batched tridiagonal solve with its scratch in a static __device__ array, no device
allocation inside the call) against nh_tridiag_solve, through the C ABI, on two calls in a row.

No reference counterpart (include/nhstep.h).  The file
name keeps this test last in collection order."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oc():
    import oceananigans_jl_b200 as oc
    assert oc.device_count() >= 1
    return oc


def _static_solve(oc, torch, FT, size, direction, a, b, c, f, phi0=None):
    """Calls the entry point with the caller's (torch) device arrays; returns phi as a host (Nx,Ny,Nz) array."""
    L = oc._lib
    tdt = torch.float64 if FT == np.float64 else torch.float32
    dev = lambda x: torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=FT).ravel(order="F"))).to("cuda:0")
    a_d, b_d, c_d, f_d = dev(a), dev(b), dev(c), dev(f)
    phi_d = torch.zeros(int(np.prod(size)), dtype=tdt, device="cuda:0") if phi0 is None else dev(phi0)
    torch.cuda.synchronize()
    N = (C.c_int32 * 3)(*size)
    L.check(L.lib().nh_tridiag_solve_static(0, L.F64 if FT == np.float64 else L.F32, N, direction, a_d.data_ptr(), b_d.data_ptr(),
                                            int(np.asarray(b).ndim == 3), c_d.data_ptr(), f_d.data_ptr(), phi_d.data_ptr(), None))
    torch.cuda.synchronize()
    return phi_d.cpu().numpy().reshape(size, order="F")


@pytest.mark.parametrize("FT,rtol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_static_scratch_tridiagonal_solve_matches_nh_tridiag_solve_twice_in_a_row(oc, FT, rtol):
    import torch
    rng = np.random.default_rng(2024)
    # two calls in a row sharing the static scratch: different sizes, directions and diagonals (1-D then 3-D b), the second
    # one smaller so that stale scratch contents of the first call lie beyond and inside its range
    cases = [((24, 20, 32), 2, False), ((16, 28, 12), 1, True), ((12, 9, 7), 0, False)]
    for size, direction, b3d in cases:
        n = size[direction]
        g = oc.RectilinearGrid(FT, size=size, extent=(1, 1, 1), topology=("Periodic", "Periodic", "Bounded"))
        a, c = rng.random(n - 1), rng.random(n - 1)
        b = 3 + (rng.random(size) if b3d else rng.random(n))
        f = rng.standard_normal(size)
        expect = oc.BatchedTridiagonalSolver(g, a, b, c, direction=direction).solve(f)
        got = _static_solve(oc, torch, FT, size, direction, a, b, c, f)
        assert got.dtype == expect.dtype
        scale = np.abs(expect).max()
        assert np.abs(got - expect).max() <= rtol * scale, (size, direction, np.abs(got - expect).max() / scale)
        # and against the dense solve, like test_batched_tridiagonal_vs_dense, for the 1-D diagonals
        if not b3d and FT == np.float64:
            M = np.diag(b) + np.diag(a, -1) + np.diag(c, 1)
            dense = np.moveaxis(np.linalg.solve(M, np.moveaxis(f, direction, 0).reshape(n, -1)).reshape(
                [n] + [s for d, s in enumerate(size) if d != direction]), 0, direction)
            assert np.abs(got - dense).max() <= 1e-12 * scale


def test_static_scratch_tridiagonal_solve_rejects_sizes_beyond_the_static_array(oc):
    import torch
    L = oc._lib
    x = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    N = (C.c_int32 * 3)(128, 128, 65)                      # 1 064 960 > 2^20 elements; rejected before anything is read
    assert 128 * 128 * 65 > L.TRIDIAG_STATIC_MAX_ELEMENTS
    status = L.lib().nh_tridiag_solve_static(0, L.F64, N, 2, x.data_ptr(), x.data_ptr(), 0, x.data_ptr(), x.data_ptr(),
                                             x.data_ptr(), None)
    assert status != 0
    assert "exceeds the static scratch array" in L.lib().nh_last_error().decode()
    torch.cuda.synchronize()
