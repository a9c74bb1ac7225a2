"""GPU: the reference's own solver tests run against the CUDA path, plus value parity with the oracle.
  test/test_poisson_solvers.jl:58-106, test/dependencies_for_poisson_solvers.jl:13-141
  test/test_poisson_solvers_stretched_grids.jl:12-50, test/test_batched_tridiagonal_solver.jl:7-42
"""
import itertools

import numpy as np
import pytest

import oracle as o
from oracle.operators import divergence_ccc, laplacian_ccc, interior_region
from oracle.fields import Field, fill_halo_regions

pytestmark = pytest.mark.gpu
P, B, F = "Periodic", "Bounded", "Flat"


@pytest.fixture(scope="module")
def oc():
    import oceananigans_jl_b200 as oc
    assert oc.device_count() >= 1
    return oc


def isapprox(a, b, rtol=None):
    a, b = np.asarray(a), np.asarray(b)
    rtol = np.sqrt(np.finfo(a.dtype).eps) if rtol is None else rtol
    return np.linalg.norm((a - b).ravel()) <= rtol * max(np.linalg.norm(a.ravel()), np.linalg.norm(b.ravel()))


def grids(oc, FT, topo, size3, z=None):
    size = tuple(n for n, t in zip(size3, topo) if t != F)
    kw = {n: (0, 1) for n, t in zip("xyz", topo) if t != F}
    if z is not None:
        kw["z"] = z
    return oc.RectilinearGrid(FT, size=size, topology=topo, **kw), o.RectilinearGrid(FT, size=size, topology=topo, **kw)


def random_rhs(og, rng):
    U = [Field(og, loc, name=n) for n, loc in zip("uvw", (("f", "c", "c"), ("c", "f", "c"), ("c", "c", "f")))]
    for f in U:
        f.set(rng.random(f.size))
        fill_halo_regions(f)
    return np.array(divergence_ccc(og, *U, interior_region(og))).astype(og.FT)


def residual_ok(og, phi, rhs):
    f = Field(og, ("c", "c", "c"), prognostic=False)
    f.interior[...] = phi
    fill_halo_regions(f)
    return isapprox(np.array(laplacian_ccc(og, f, interior_region(og))), rhs)


TOPOS3 = list(itertools.product((P, B), repeat=3))
TOPOS2 = [(F, P, P), (F, P, B), (F, B, B), (P, F, P), (P, F, B), (B, F, B), (P, P, F), (P, B, F), (B, B, F)]


@pytest.mark.parametrize("topo", TOPOS3 + TOPOS2)
def test_fft_poisson_property_and_oracle(oc, topo):
    rng = np.random.default_rng(1)
    for size3 in [(7, 7, 7), (16, 16, 16), (11, 16, 11), (16, 11, 16), (1, 16, 16), (16, 16, 1)]:
        g, og = grids(oc, np.float64, topo, size3)
        rhs = random_rhs(og, rng)
        phi = oc.FFTBasedPoissonSolver(g).solve(rhs)
        assert residual_ok(og, phi, rhs), (topo, size3)
        ref = o.FFTBasedPoissonSolver(og).solve(None, rhs)
        scale = np.abs(ref).max()
        assert np.abs(phi - ref).max() <= 1e-12 * scale, (topo, size3)


def test_fft_poisson_float32(oc):
    rng = np.random.default_rng(2)
    for topo, size3 in [((P, B, B), (16, 16, 16)), ((B, P, P), (7, 11, 16))]:
        g, og = grids(oc, np.float32, topo, size3)
        rhs = random_rhs(og, rng)
        phi = oc.FFTBasedPoissonSolver(g).solve(rhs)
        assert residual_ok(og, phi, rhs)
        ref = o.FFTBasedPoissonSolver(og).solve(None, rhs)
        assert np.abs(phi - ref).max() <= 2e-5 * np.abs(ref).max()


@pytest.mark.parametrize("topo", TOPOS3)
def test_fft_poisson_convergence(oc, topo):
    def err(N, mode):
        two_pi = 2 * np.pi
        g = oc.RectilinearGrid(np.float64, size=(N, N, N), x=(0, two_pi), y=(0, two_pi), z=(0, two_pi), topology=topo)
        psi = {B: lambda x: np.cos(mode * x / 2), P: lambda x: np.cos(mode * x)}
        k2 = {B: (mode / 2) ** 2, P: mode ** 2}
        c = [g.nodes(d, "c") for d in range(3)]
        Psi = psi[topo[0]](c[0])[:, None, None] * psi[topo[1]](c[1])[None, :, None] * psi[topo[2]](c[2])[None, None, :]
        f = -(k2[topo[0]] + k2[topo[1]] + k2[topo[2]]) * Psi
        return np.mean(np.abs(oc.FFTBasedPoissonSolver(g).solve(f) - Psi))
    for N1, N2 in ((64, 128), (67, 131)):
        rate = np.log(err(N1, 2) / err(N2, 2)) / np.log(N2 / N1)
        assert abs(rate - 2) <= 1e-2, (topo, N1, N2, rate)


FACES = np.array([1, 2, 4, 7, 11, 16, 22, 29, 37], float)


@pytest.mark.parametrize("topo", [(P, P, B), (P, B, B), (B, P, B), (B, B, B), (F, P, B), (P, F, B)])
@pytest.mark.parametrize("FT", [np.float64, np.float32])
def test_fourier_tridiagonal_property_and_oracle(oc, topo, FT):
    rng = np.random.default_rng(4)
    for N1, N2 in ((8, 8), (7, 11)):
        g, og = grids(oc, FT, topo, (N1, N2, len(FACES) - 1), z=FACES)
        rhs = random_rhs(og, rng)
        rhs = (rhs - rhs.mean()).astype(FT)        # make it solvable (the reference builds a divergence-free U instead)
        phi = oc.FourierTridiagonalPoissonSolver(g).solve(rhs)
        ref = o.FourierTridiagonalPoissonSolver(og).solve(None, rhs)
        tol = 1e-11 if FT == np.float64 else 5e-4
        assert np.abs(phi - ref).max() <= tol * np.abs(ref).max(), (topo, N1, N2)


@pytest.mark.parametrize("direction", [0, 1, 2])
def test_batched_tridiagonal_vs_dense(oc, direction):
    rng = np.random.default_rng(11)
    N = 8
    size = [3, 4, 5]
    size[direction] = N
    g = oc.RectilinearGrid(np.float64, size=tuple(size), extent=(1, 1, 1), topology=(P, P, B))
    a, c = rng.random(N - 1), rng.random(N - 1)
    b = 3 + rng.random(N)
    f = rng.random(size)
    phi = oc.BatchedTridiagonalSolver(g, a, b, c, direction=direction).solve(f)
    M = np.diag(b) + np.diag(a, -1) + np.diag(c, 1)
    expect = np.moveaxis(np.linalg.solve(M, np.moveaxis(f, direction, 0).reshape(N, -1)).reshape(
        [N] + [s for d, s in enumerate(size) if d != direction]), 0, direction)
    assert np.allclose(phi, expect, rtol=1e-12, atol=0)


def test_fft_poisson_large_roundtrip_property(oc):
    """Size-independent property at a bench-sized grid: ∇²(solve(rhs)) == rhs - mean(rhs) on (P,P,P) 256³."""
    N = 256
    g = oc.RectilinearGrid(np.float64, size=(N, N, N), extent=(1, 1, 1), topology=(P, P, P))
    rng = np.random.default_rng(9)
    rhs = rng.standard_normal((N, N, N))
    rhs -= rhs.mean()
    phi = oc.FFTBasedPoissonSolver(g).solve(rhs)
    d = 1.0 / N
    lap = sum((np.roll(phi, -1, ax) - 2 * phi + np.roll(phi, 1, ax)) for ax in range(3)) / d ** 2
    assert np.linalg.norm((lap - rhs).ravel()) <= 1e-9 * np.linalg.norm(rhs.ravel())


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 3e-5)])
@pytest.mark.parametrize("topo", [(P, P, P), (P, P, B)])
@pytest.mark.parametrize("size3", [(32, 16, 64), (64, 32, 16), (16, 64, 32), (128, 128, 128)])
def test_fused_standalone_solver_vs_oracle_and_complete_path(oc, topo, size3, FT, tol, monkeypatch):
    """The fused FFT / DCT passes (power-of-two (P,P,P|B) grids) against the oracle and against the complete path, with and
    without a Helmholtz shift m (fft_based_poisson_solver.jl:95-125)."""
    rng = np.random.default_rng(5)
    g, og = grids(oc, FT, topo, size3)
    rhs = rng.standard_normal(size3).astype(FT)
    rhs -= rhs.mean(dtype=np.float64).astype(FT)
    fused = oc.FFTBasedPoissonSolver(g)
    monkeypatch.setenv("NHSTEP_NO_FAST_POISSON", "1")
    complete = oc.FFTBasedPoissonSolver(g)
    monkeypatch.delenv("NHSTEP_NO_FAST_POISSON")
    assert fused.launch_count() == 0
    for m in (0.0, -3.5):
        phi = fused.solve(rhs, m)
        ref = o.FFTBasedPoissonSolver(og).solve(None, rhs, m)
        scale = np.abs(ref).max()
        assert np.abs(phi - ref).max() <= tol * scale, (topo, size3, m)
        assert np.abs(phi - complete.solve(rhs, m)).max() <= tol * scale
    assert fused.launch_count() == 10 and complete.launch_count() > 10      # 5 fused launches per solve


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-11), (np.float32, 1e-4)])
@pytest.mark.parametrize("size3", [(16, 16, 8), (32, 16, 21), (64, 32, 40), (128, 128, 64)])
def test_fused_fourier_tridiagonal_vs_oracle_and_complete_path(oc, size3, FT, tol, monkeypatch):
    """FFT x/y passes around the Thomas sweep on the half spectrum (stretched Bounded z, any Nz) against the oracle and the
    complete path; two solves in a row exercise the storage that persists between solves
    (fourier_tridiagonal_poisson_solver.jl:119-147, batched_tridiagonal_solver.jl:209-235)."""
    rng = np.random.default_rng(11)
    Nz = size3[2]
    faces = np.cumsum(np.concatenate([[0.0], 0.5 + rng.random(Nz)]))
    g, og = grids(oc, FT, (P, P, B), size3, z=faces)
    fused = oc.FourierTridiagonalPoissonSolver(g)
    monkeypatch.setenv("NHSTEP_NO_FAST_POISSON", "1")
    complete = oc.FourierTridiagonalPoissonSolver(g)
    monkeypatch.delenv("NHSTEP_NO_FAST_POISSON")
    ref_solver = o.FourierTridiagonalPoissonSolver(og)
    base = fused.launch_count()
    for trial in range(2):
        rhs = random_rhs(og, rng)
        phi = fused.solve(rhs)
        ref_solver.set_source_term(rhs)
        ref = np.array(ref_solver.solve()).real.astype(FT)
        scale = np.abs(ref).max()
        assert np.abs(phi - ref).max() <= tol * scale, (size3, trial, np.abs(phi - ref).max() / scale)
        assert np.abs(phi - complete.solve(rhs)).max() <= tol * scale
        if FT == np.float64:                   # in Float32 the 1/Δ² of the residual check swamps √eps at these sizes
            assert residual_ok(og, phi, rhs)
    assert fused.launch_count() - base == 12   # x, y, Thomas, y, x + the one-block stale-value gather, per solve
