"""CPU oracle for the NonhydrostaticModel time step (TEST INFRASTRUCTURE ONLY).

This package is a NumPy/SciPy restatement of the *CPU code path* of the reference
(samlewin/Oceananigans.jl v0.96.19) for the one hot path this repository accelerates:
tendencies -> RK3/AB2 substep -> halo fill -> FFT / Fourier-tridiagonal pressure solve ->
pressure correction.  Every function cites the reference file:line it follows.

It is the *checker*, never the product:
  * only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
    ``--impl reference`` legs may import it;
  * the product path (``oceananigans.jl_b200/``) never imports it and has no CPU fallback.

Pinning status: the reference is pure Julia and cannot run in the build container (no
``julia``), and its golden regression files are remote DataDeps
(``test/data_dependencies.jl:17-38``).  The oracle is therefore pinned by the reference's
own *property / analytic* tests re-run against it (``tests/test_oracle_*.py``): Poisson
``∇²ϕ ≈ div U`` over the topology matrix, second-order convergence, Thomas vs dense solve,
Taylor–Green decay, incompressibility after 1/10/100 steps, first-AB2-step exactness, tracer
conservation.  There are no golden *value* vectors from a live reference run:
**parity unpinned (values)** — see DESIGN.md.

FFT backend: ``scipy.fft`` (pocketfft) stands in for FFTW (FFTW.jl "1", not vendored in
/root/reference); both compute the same DFT / REDFT10 / REDFT01 to rounding.
"""

from .grid import RectilinearGrid, Periodic, Bounded, Flat  # noqa: F401
from .fields import Field, fill_halo_regions  # noqa: F401
from .advection import WENO, Centered, UpwindBiased  # noqa: F401
from .solvers import (FFTBasedPoissonSolver, FourierTridiagonalPoissonSolver,  # noqa: F401
                      BatchedTridiagonalSolver)
from .model import (NonhydrostaticModel, time_step, set_model, FPlane,  # noqa: F401
                    BuoyancyTracer, SeawaterBuoyancy, ScalarDiffusivity, SmagorinskyLilly, Smagorinsky,
                    AnisotropicMinimumDissipation)
