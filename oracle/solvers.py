"""Pressure Poisson solvers (oracle; test infrastructure only).

Follows the reference's CPU branches of
  src/Solvers/poisson_eigenvalues.jl:8-31
  src/Solvers/fft_based_poisson_solver.jl:52-74,95-137
  src/Solvers/plan_transforms.jl:16-34,78-82,119-146,149-179   (r2r REDFT10/REDFT01 over Bounded dims, fft/ifft over Periodic)
  src/Solvers/discrete_transforms.jl:26-34,108-122,178-184     (CPU backward Bounded normalisation 1/2N)
  src/Solvers/batched_tridiagonal_solver.jl:100-123,133-235    (Thomas sweep with the |β| > 10·eps guard)
  src/Solvers/fourier_tridiagonal_poisson_solver.jl:41-147,155-177

FFT backend: scipy.fft (pocketfft) in place of FFTW.jl; ``dct(type=2)`` ≡ REDFT10 and ``dct(type=3)`` ≡ REDFT01
(both unnormalised), applied to real and imaginary parts as FFTW's r2r on a complex array does.
"""
import numpy as np
import scipy.fft as sfft

from .grid import Periodic, Bounded, Flat


def poisson_eigenvalues(N, L, topo):
    """poisson_eigenvalues.jl:8-31 — always Float64 arithmetic on (inds - 1)·π."""
    if topo == Flat:
        return np.zeros(N)
    inds = np.arange(1, N + 1)
    h = L / N          # L is FT; for Float32 grids this division happens in Float32 like the reference's L / N
    if topo == Periodic:
        s = np.sin((inds - 1) * np.pi / N)
    else:
        s = np.sin((inds - 1) * np.pi / (2 * N))
    x = 2 * s / np.float64(h)
    return x * x


def _complex_type(FT):
    return np.complex128 if np.dtype(FT) == np.float64 else np.complex64


def _forward(b, topo, dims=(0, 1, 2)):
    bounded = [d for d in dims if topo[d] == Bounded]
    periodic = [d for d in dims if topo[d] == Periodic]
    for d in bounded:                                   # one multi-dim r2r plan in the reference; separable
        b = sfft.dct(b.real, type=2, axis=d) + 1j * sfft.dct(b.imag, type=2, axis=d)
    if periodic:
        b = sfft.fftn(b, axes=periodic)
    return b


def _backward(b, topo, dims=(0, 1, 2)):
    bounded = [d for d in dims if topo[d] == Bounded]
    periodic = [d for d in dims if topo[d] == Periodic]
    if periodic:
        b = sfft.ifftn(b, axes=periodic)
    norm = 1.0
    for d in bounded:
        b = sfft.dct(b.real, type=3, axis=d) + 1j * sfft.dct(b.imag, type=3, axis=d)
        norm = norm * (1 / (2 * b.shape[d]))
    if bounded:
        b = b * norm
    return b


class FFTBasedPoissonSolver:
    def __init__(self, grid):
        for d in range(3):
            if not grid.regular_dim(d):
                raise ValueError("FFTBasedPoissonSolver requires a grid that is regular in x, y and z")
        self.grid = grid
        self.ctype = _complex_type(grid.FT)
        L = (grid.Lx, grid.Ly, grid.Lz)
        self.lam = [poisson_eigenvalues(grid.N[d], L[d], grid.topology[d]) for d in range(3)]
        self.storage = np.zeros(grid.N, dtype=self.ctype)

    def solve(self, phi_field=None, b=None, m=0):
        """solve!(ϕ, solver, b, m): ϕ.interior ← Re(backward(-forward(b) / (λx+λy+λz-m)))."""
        topo = self.grid.topology
        if b is not None:
            self.storage[...] = b
        bh = _forward(self.storage, topo).astype(self.ctype)
        lx, ly, lz = self.lam
        denom = lx[:, None, None] + ly[None, :, None] + lz[None, None, :] - m
        with np.errstate(divide="ignore", invalid="ignore"):
            ph = (-bh / denom).astype(self.ctype)
        if m == 0:
            ph[0, 0, 0] = 0
        self.storage[...] = _backward(ph, topo).astype(self.ctype)
        if phi_field is not None:
            phi_field.interior[...] = self.storage.real
        return self.storage.real


class BatchedTridiagonalSolver:
    """Thomas algorithm batched over the two other dims; a, c: 1-D (length N-1, a[k] couples k+1 to k) or 3-D;
    b: 1-D or 3-D; direction ∈ {0,1,2} (XDirection, YDirection, ZDirection)."""

    def __init__(self, grid, lower_diagonal, diagonal, upper_diagonal, direction=2, scratch_dtype=None):
        self.grid, self.a, self.b, self.c, self.dir = grid, lower_diagonal, diagonal, upper_diagonal, direction
        self.t = np.zeros(grid.N, dtype=scratch_dtype or grid.FT)

    def _coef(self, arr, k):
        arr = np.asarray(arr)
        if arr.ndim == 1:
            return arr[k]
        sl = [slice(None)] * 3
        sl[self.dir] = k
        return arr[tuple(sl)]

    def solve(self, phi, f):
        """In-place on phi (which keeps its previous content where the guard elides the update)."""
        d, N = self.dir, self.grid.N[self.dir]
        real_t = np.float32 if phi.dtype in (np.float32, np.complex64) else np.float64
        guard = 10 * np.finfo(real_t).eps

        def S(arr, k):
            sl = [slice(None)] * 3
            sl[d] = k
            return arr[tuple(sl)]
        beta = self._coef(self.b, 0)
        phi[_ix(d, 0)] = S(f, 0) / beta
        for k in range(1, N):
            ck, bk, ak = self._coef(self.c, k - 1), self._coef(self.b, k), self._coef(self.a, k - 1)
            tk = ck / beta
            self.t[_ix(d, k)] = tk
            tk = self.t[_ix(d, k)]
            beta = bk - ak * tk
            with np.errstate(divide="ignore", invalid="ignore"):
                star = (S(f, k) - ak * S(phi, k - 1)) / beta
            ok = np.abs(beta) > guard
            phi[_ix(d, k)] = np.where(ok, star, S(phi, k))
        for k in range(N - 2, -1, -1):
            phi[_ix(d, k)] = S(phi, k) - self.t[_ix(d, k + 1)] * S(phi, k + 1)
        return phi


def _ix(d, k):
    sl = [slice(None)] * 3
    sl[d] = k
    return tuple(sl)


class FourierTridiagonalPoissonSolver:
    def __init__(self, grid, tridiagonal_direction=None):
        irregular = [d for d in range(3) if not grid.regular_dim(d)]
        if tridiagonal_direction is None:
            if len(irregular) != 1:
                raise ValueError("FourierTridiagonalPoissonSolver needs exactly one stretched direction "
                                 "(or an explicit tridiagonal_direction)")
            tridiagonal_direction = irregular[0]
        td = tridiagonal_direction
        if grid.topology[td] != Bounded:
            raise ValueError("`FourierTridiagonalPoissonSolver` can only be used when the stretched "
                             "direction's topology is `Bounded`.")
        self.grid, self.td = grid, td
        self.ctype = _complex_type(grid.FT)
        self.other = [d for d in range(3) if d != td]
        L = (grid.Lx, grid.Ly, grid.Lz)
        lam = [poisson_eigenvalues(grid.N[d], L[d], grid.topology[d]) for d in self.other]
        N = grid.N[td]
        c = grid.coords[td]

        FT = grid.FT
        # lower = upper = 1/Δξᶠ(q), q = 2..N, computed in FT like the reference's comprehension
        lower = np.array([FT(1) / (c.df if c.regular else c.df[q]) for q in range(2, N + 1)], dtype=FT)
        self.lower = lower
        # main diagonal (Float64 array: zeros(size(grid)...)), compute_main_diagonal! :41-51
        l1 = lam[0].reshape([-1 if e == self.other[0] else 1 for e in range(3)])
        l2 = lam[1].reshape([-1 if e == self.other[1] else 1 for e in range(3)])
        lsum = l1 + l2
        D = np.zeros(grid.N, dtype=np.float64)

        def put(q, val):          # q is the 1-based index along td
            D[_ix(td, q - 1)] = np.broadcast_to(np.squeeze(val, axis=td), D[_ix(td, q - 1)].shape)
        dfv = (lambda q: (c.df if c.regular else c.df[q]))
        dcv = (lambda q: (c.dc if c.regular else c.dc[q]))
        put(1, -1 / dfv(2) - dcv(1) * lsum)
        for q in range(2, N):
            put(q, -(1 / dfv(q + 1) + 1 / dfv(q)) - dcv(q) * lsum)
        put(N, -1 / dfv(N) - dcv(N) * lsum)          # written last, as in the reference (wins when N == 1)
        self.D = D
        self.tri = BatchedTridiagonalSolver(grid, lower, D, lower, direction=td)
        self.source_term = np.zeros(grid.N, dtype=self.ctype)
        self.storage = np.zeros(grid.N, dtype=self.ctype)

    def set_source_term(self, b):
        """set_source_term!: source_term = b · Δξᶜ along the tridiagonal direction (:155-177)."""
        c = self.grid.coords[self.td]
        N = self.grid.N[self.td]
        dc = c.dc if c.regular else c.dc.r(1, N).reshape([-1 if e == self.td else 1 for e in range(3)])
        self.source_term[...] = np.asarray(b).astype(self.ctype) * dc

    def solve(self, phi_field=None, b=None):
        if b is not None:
            self.set_source_term(b)
        topo = self.grid.topology
        self.source_term[...] = _forward(self.source_term, topo, self.other).astype(self.ctype)
        self.tri.solve(self.storage, self.source_term)
        self.storage[...] = _backward(self.storage, topo, self.other).astype(self.ctype)
        self.storage[...] = self.storage - np.mean(self.storage)
        if phi_field is not None:
            phi_field.interior[...] = self.storage.real
        return self.storage.real
