"""GPU parity at the sizes BASELINE.json's `configs` name: full-size runs against the compiled C oracle
(oracle/cpu_step.c, itself pinned to the NumPy oracle in tests/test_cpu_step.py) where it covers the configuration, and
size-independent properties (∇²ϕ = rhs, ∇·u = 0, fused path = complete path) elsewhere."""
import numpy as np
import pytest

import oracle as o
import parity_log
from oracle.cpu_step import CpuStepModel

pytestmark = pytest.mark.gpu
P, B, F = "Periodic", "Bounded", "Flat"


@pytest.fixture(scope="module")
def oc():
    import oceananigans_jl_b200 as oc
    assert oc.device_count() >= 1, "no CUDA device: the product has no CPU fallback"
    return oc


def rel_l2(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm((a - b).ravel()) / np.linalg.norm(b.ravel())


def test_config0_2d_turbulence_128_readme_example(oc):
    """configs[0]: 128² (Periodic, Periodic, Flat), x, y ∈ (0, 2π), WENO5, RK3, u, v ~ U(-1, 1) (README.md:113-120), Δt = 0.01,
    400 steps (t = 4); parity with the NumPy oracle at step 100 and 400, Float64, rel. L2 ≤ 1e-12 (chaotic growth of the FFT
    rounding difference is measured by the oracle's own 1-ulp sensitivity and the bound is 10× that where larger)."""
    N = 128
    rng = np.random.default_rng(1234)
    ics = {n: rng.uniform(-1, 1, (N, N, 1)) for n in "uv"}      # fields of a Flat-z grid are N×N×1
    kw = dict(size=(N, N), x=(0, 2 * np.pi), y=(0, 2 * np.pi), topology=(P, P, F))
    m1 = oc.NonhydrostaticModel(oc.RectilinearGrid(np.float64, **kw), advection=oc.WENO(order=5))
    m2 = o.NonhydrostaticModel(o.RectilinearGrid(np.float64, **kw), advection=o.WENO(np.float64, 5))
    m3 = o.NonhydrostaticModel(o.RectilinearGrid(np.float64, **kw), advection=o.WENO(np.float64, 5))
    oc.set_model(m1, **ics)
    o.set_model(m2, **ics)
    eps = np.finfo(np.float64).eps
    o.set_model(m3, **{n: v * (1 + eps * rng.choice([-1, 1], v.shape)) for n, v in ics.items()})
    done = 0
    for upto in (100, 400):
        oc.time_step(m1, 0.01, n_steps=upto - done)
        for _ in range(upto - done):
            o.time_step(m2, 0.01)
            o.time_step(m3, 0.01)
        done = upto
        for a, n in enumerate("uv"):
            ref = m2.velocities[a].interior
            err = rel_l2(m1.velocities[n].interior(), ref)
            floor = rel_l2(m3.velocities[a].interior, ref)
            assert err <= max(1e-12, 10 * floor), (upto, n, err, floor)
    oc.fill_halo_regions(list(m1.velocities.values()))
    assert m1.max_abs_divergence() < 1e-10


def test_config1_256_cubed_full_size_100_steps_against_the_compiled_oracle(oc):
    """configs[1] at its FULL size: 256³ (P,P,P) Float64 WENO5 RK3, the bench's initial condition, 100 complete steps on the GPU
    (fused fast path) in BOTH arithmetic modes — exact (reference expression order, no FMA) and contracted (the bench's default)
    — against oracle/cpu_step.c on the host cores, compared after 2, 10 and 100 steps: u, v, w ≤ 1e-12 and p ≤ 1e-12 strictly
    (the north_star bound, no floor relaxation), the discrete divergence at rounding level.  The achieved errors go to the
    parity log (profiles/r2_parity_errors.jsonl)."""
    from bench import smooth_ics
    N = 256
    ics = smooth_ics(N, 1234)
    dt = 0.2 * (2 * np.pi / N) / 2.0
    g = oc.RectilinearGrid(np.float64, size=(N, N, N), extent=(2 * np.pi,) * 3, topology=(P, P, P))
    models = {name: oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), exact_arithmetic=name == "exact")
              for name in ("exact", "contracted")}
    for m in models.values():
        oc.set_model(m, **ics)
    CpuStepModel.use_all_cores()
    c = CpuStepModel((N, N, N), (2 * np.pi,) * 3)
    c.set(ics["u"], ics["v"], ics["w"])
    done, failures = 0, []
    for upto in (2, 10, 100):
        c.time_step(dt, upto - done)
        ref = {n: c.get(n) for n in ("u", "v", "w", "p")}
        for name, m in models.items():
            oc.time_step(m, dt, n_steps=upto - done)
            err = {n: rel_l2(m.velocities[n].interior(), ref[n]) for n in "uvw"}
            err["p"] = rel_l2(m.pressures.pNHS.interior(), ref["p"])
            err["max_abs_divergence"] = m.max_abs_divergence()
            parity_log.record(err, 1e-12, note=f"256^3 configs[1], {name} arithmetic, after {upto} steps, vs oracle/cpu_step.c")
            for n in ("u", "v", "w", "p"):
                if not err[n] <= 1e-12:
                    failures.append((name, upto, n, err[n]))
            if not err["max_abs_divergence"] < 1e-10:
                failures.append((name, upto, "div", err["max_abs_divergence"]))
        done = upto
    assert not failures, failures


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-9), (np.float32, 2e-2)])
def test_config2_fft_solver_512_cubed_ppb(oc, FT, tol):
    """configs[2] at full size: FFTBasedPoissonSolver on 512³ (P,P,B), DCT in z: the reference's own test property
    ∇²ϕ ≈ rhs (test_poisson_solvers.jl:61-97) with the second-order Laplacian and Neumann walls, evaluated in Float64."""
    N = 512
    g = oc.RectilinearGrid(FT, size=(N, N, N), extent=(1, 1, 1), topology=(P, P, B))
    rng = np.random.default_rng(1234)
    rhs = rng.standard_normal((N, N, N)).astype(FT)
    rhs -= rhs.mean(dtype=np.float64).astype(FT)
    phi = oc.FFTBasedPoissonSolver(g).solve(rhs).astype(np.float64)
    d = 1.0 / N
    lap = (np.roll(phi, -1, 0) - 2 * phi + np.roll(phi, 1, 0)) + (np.roll(phi, -1, 1) - 2 * phi + np.roll(phi, 1, 1))
    pz = np.concatenate([phi[:, :, :1], phi, phi[:, :, -1:]], axis=2)          # no-flux walls: mirror
    lap += pz[:, :, 2:] - 2 * phi + pz[:, :, :-2]
    lap /= d * d
    r = rhs.astype(np.float64)
    assert np.linalg.norm((lap - r).ravel()) <= tol * np.linalg.norm(r.ravel())


def test_config3_fourier_tridiagonal_512x512x256_stretched(oc):
    """configs[3] at full size: FourierTridiagonalPoissonSolver on 512×512×256 with the reference's tanh-stretched faces
    (test_time_stepping.jl:439-445): ∇²ϕ ≈ rhs with the stretched-z Laplacian 1/Δzᶜ·δz(δzϕ/Δzᶠ)."""
    Nx = Ny = 512
    Nz = 256
    k = np.arange(1, Nz + 2)
    zf = np.tanh(1.3 * (2 * (k - 1) / Nz - 1)) / np.tanh(1.3)
    g = oc.RectilinearGrid(np.float64, size=(Nx, Ny, Nz), x=(0, 1), y=(0, 1), z=zf, topology=(P, P, B))
    rng = np.random.default_rng(1234)
    dzc = np.diff(zf)
    rhs = rng.standard_normal((Nx, Ny, Nz))
    rhs -= (rhs * dzc).sum(axis=2, keepdims=True).mean(axis=(0, 1), keepdims=True) / dzc.sum()     # volume mean zero
    phi = oc.FourierTridiagonalPoissonSolver(g).solve(rhs)
    d = 1.0 / Nx
    lap = ((np.roll(phi, -1, 0) - 2 * phi + np.roll(phi, 1, 0)) + (np.roll(phi, -1, 1) - 2 * phi + np.roll(phi, 1, 1))) / d ** 2
    zc = 0.5 * (zf[1:] + zf[:-1])
    dzf = np.diff(zc)                                     # Δzᶠ at interior faces 2…Nz
    flux = np.zeros((Nx, Ny, Nz + 1))
    flux[:, :, 1:-1] = (phi[:, :, 1:] - phi[:, :, :-1]) / dzf
    lap += (flux[:, :, 1:] - flux[:, :, :-1]) / dzc
    assert np.linalg.norm((lap - rhs).ravel()) <= 1e-8 * np.linalg.norm(rhs.ravel())


def test_config3_stretched_model_step_stays_incompressible(oc):
    """configs[3], the full model on the stretched grid (256×256×128 sample): ∇·u after 3 RK3 steps stays below the
    reference's own bound (test_time_stepping.jl:125-158: 5e-8), on the fused Thomas passes."""
    Nx, Nz = 256, 128
    k = np.arange(1, Nz + 2)
    zf = np.tanh(1.3 * (2 * (k - 1) / Nz - 1)) / np.tanh(1.3)
    g = oc.RectilinearGrid(np.float64, size=(Nx, Nx, Nz), x=(0, 1), y=(0, 1), z=zf, topology=(P, P, B))
    m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), buoyancy=oc.SeawaterBuoyancy(), tracers=("T", "S"))
    rng = np.random.default_rng(1)
    T = np.full((Nx, Nx, Nz), 20.0)
    T[Nx // 4:3 * Nx // 4, Nx // 4:3 * Nx // 4, Nz // 4:3 * Nz // 4] += 0.01
    oc.set_model(m, u=0.1 * rng.uniform(-1, 1, (Nx, Nx, Nz)), T=T, S=np.full((Nx, Nx, Nz), 35.0))
    oc.time_step(m, 1e-3, n_steps=3)
    oc.fill_halo_regions(list(m.velocities.values()))
    assert m.max_abs_divergence() < 5e-8
