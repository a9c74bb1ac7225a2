// kernels_basic.cuh — halo fills, RK3/AB2 substeps, source term, pressure correction, hydrostatic pressure,
// layout conversion.  Complete (any topology) versions in the reference's expression order; the fused fast kernels
// live in step_fast.cu.
#pragma once
#include "tendency_generic.cuh"
#include "closure_amd.cuh"

#define NH_MAXF (3 + NH_MAXTR)

// ---- halo fill ------------------------------------------------------------------------------------------------
// src/BoundaryConditions/fill_halo_regions_periodic.jl:40-111, fill_halo_regions_flux.jl:14-33,
// fill_halo_regions_open.jl:65-70, fill_halo_regions_value_gradient.jl:7-99.  One launch fills one direction of up to
// NH_MAXF fields (the tupled fill of src/Fields/field_tuples.jl:88-101).
template <typename FT>
struct HaloJob {
  GridDev<FT> g;
  FT* p[NH_MAXF];
  int mode[NH_MAXF];      // 0 skip, 1 periodic, 2 wall (per-side kind below), 3 open (wall-normal velocity := 0)
  int side_kind[NH_MAXF][2];   // mode 2: NH_BC_DEFAULT / NH_BC_FLUX (mirror), NH_BC_VALUE, NH_BC_GRADIENT at the left / right wall
  FT side_val[NH_MAXF][2];
  const FT* side_arr[NH_MAXF][2];   // array condition (N_e1 × N_e2 over the transverse grid range, e1 fastest) or nullptr
  // slab runs (csrc/dist.cuh): the periodic fill along y reads the previous / next rank's field over NVLink instead of this
  // rank's own far end (same index arithmetic, different base pointer); nullptr = this rank's own field
  const FT* pull_prev[NH_MAXF];
  const FT* pull_next[NH_MAXF];
  int nf;
  int d;                  // direction being filled
  int lo[3], hi[3];       // transverse logical index range (inclusive); entry d unused
};

template <typename FT>
__global__ void halo_fill_kernel(const __grid_constant__ HaloJob<FT> J) {
  const GridDev<FT>& g = J.g;
  const int d = J.d;
  const int e1 = d == 0 ? 1 : 0, e2 = d == 2 ? 1 : 2;
  int n1 = J.hi[e1] - J.lo[e1] + 1, n2 = J.hi[e2] - J.lo[e2] + 1;
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= (long long)n1 * n2) return;
  int a1 = J.lo[e1] + (int)(t % n1), a2 = J.lo[e2] + (int)(t / n1);
  int f = blockIdx.y;
  int mode = J.mode[f];
  if (mode == 0) return;
  FT* p = J.p[f];
  const int N = g.N[d], H = g.H[d];
  int c[3];
  c[e1] = a1; c[e2] = a2;
  auto at = [&](int q) -> FT& { c[d] = q; return p[g.idx(c[0], c[1], c[2])]; };
  if (mode == 1) {
    const FT* pp = J.pull_prev[f] ? J.pull_prev[f] : p;
    const FT* pn = J.pull_next[f] ? J.pull_next[f] : p;
    for (int h = 1; h <= H; ++h) {
      c[d] = N - H + h; const FT lo = pp[g.idx(c[0], c[1], c[2])];
      c[d] = h;         const FT hi = pn[g.idx(c[0], c[1], c[2])];
      at(h - H) = lo;                  // west halo  ← east end of the interior (of the previous rank on a slab)
      at(N + h) = hi;                  // east halo  ← west end of the interior (of the next rank)
    }
  } else if (mode == 2) {
    // Δ between the first interior and the first halo point: the Face spacing at index 1 / N+1
    for (int side = 0; side < 2; ++side) {
      const int kind = J.side_kind[f][side];
      const int iI = side ? N : 1, iH = side ? N + 1 : 0, iB = side ? N + 1 : 1;
      const FT cI = at(iI);
      if (kind != NH_BC_VALUE && kind != NH_BC_GRADIENT) { at(iH) = cI; continue; }
      const FT D = d == 0 ? g.dx : d == 1 ? g.dy : g.Dzf(iB);
      const FT* arr = J.side_arr[f][side];
      const FT v = arr ? arr[(a1 - 1) + (long long)(a2 - 1) * g.N[e1]] : J.side_val[f][side];
      FT grad;
      if (kind == NH_BC_GRADIENT) grad = v;
      else grad = side ? (v - cI) / (D / FT(2)) : (cI - v) / (D / FT(2));
      at(iH) = cI + grad * (side ? D : -D);
    }
  } else {
    at(1) = FT(0);
    at(N + 1) = FT(0);
  }
}

// ---- flux boundary conditions ------------------------------------------------------------------------------------
// apply_{x,y,z}_bcs! — src/BoundaryConditions/apply_flux_bcs.jl:50-160: Gc[1] += q·A/V, Gc[N] -= q·A/V over the
// grid-interior transverse range (`:yz`, `:xz`, `:xy`), A at the boundary face, V at the field's own location.
template <typename FT>
struct FluxBcJob {
  GridDev<FT> g;
  FT* G;
  int d;                  // direction of the wall pair
  int zloc_face;          // the field sits at Faces in z (w): Δz of its volume / transverse areas is Δzᶠ
  int has[2];
  FT q[2];
  const FT* qarr[2];      // array condition or nullptr
};

template <typename FT>
__global__ void flux_bc_kernel(const __grid_constant__ FluxBcJob<FT> J) {
  const GridDev<FT>& g = J.g;
  const int d = J.d;
  const int e1 = d == 0 ? 1 : 0, e2 = d == 2 ? 1 : 2;
  const int n1 = g.N[e1], n2 = g.N[e2];
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= (long long)n1 * n2) return;
  int c[3];
  c[e1] = 1 + (int)(t % n1); c[e2] = 1 + (int)(t / n1);
  const int N = g.N[d];
  for (int side = 0; side < 2; ++side) {
    if (!J.has[side]) continue;
    c[d] = side ? N : 1;
    const int k = c[2];
    // Δz at the field's own vertical location (index k), and at the flipped location on the wall for d == 2
    const FT dzl = J.zloc_face ? g.Dzf(k) : g.Dzc(k);
    FT A;
    if (d == 0) A = g.dy * dzl;             // Ax = Δy·Δz
    else if (d == 1) A = g.dx * dzl;        // Ay = Δx·Δz
    else A = g.dx * g.dy;                   // Az = Δx·Δy
    const FT V = (g.dx * g.dy) * dzl;       // V = Az·Δz
    FT& G = J.G[g.idx(c[0], c[1], c[2])];
    const FT q = J.qarr[side] ? J.qarr[side][t] : J.q[side];
    const FT inc = q * A / V;
    if (side == 0) G += inc; else G -= inc;
  }
}

// ---- vertically implicit diffusion -----------------------------------------------------------------------------------
// implicit_step!(field, …, Δt) — vertically_implicit_diffusion_solver.jl:57-121,231-273: (1 - Δt ∂z κ ∂z) ϕⁿ⁺¹ = ϕ★ per
// column with the Thomas sweep of batched_tridiagonal_solver.jl:209-235 (coefficient functions, |β| < 10·eps guard,
// scratch t in the field's type).  CT is the type the reference forms the coefficients in: Float64 under RK3 (stage Δt
// is the caller's Float64 Δt times γ + ζ), the field's type under AB2 (Δt converted).
// With a constant κ on a horizontally regular grid the coefficient recurrence (a, β, t per level) is the same in every
// column, so it is evaluated once per (Δt, κ, location) by implicit_coeff_kernel — the same expressions in the same order
// as the reference evaluates per column — into a table of 4 × (Nz + 2) doubles (a Float32 value is held exactly), and
// implicit_step_kernel streams the columns: one thread per column, 8 levels of loads in flight, ϕ read and written twice.
template <typename FT, typename CT>
struct ImplicitCoeffJob {
  GridDev<FT> g;
  double* tab;            // [0] a_k = lower(k-1), [1] β_k, [2] t_k, [3] 1 if |β_k| ≥ 10 eps — each Nz + 2 long, indexed by k
  CT dt, kappa;
  int w_field;            // the field sits at Faces in z (w)
};

template <typename FT, typename CT>
__global__ void implicit_coeff_kernel(const __grid_constant__ ImplicitCoeffJob<FT, CT> J) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const GridDev<FT>& g = J.g;
  const int Nz = g.N[2], L = Nz + 2;
  const CT dt = J.dt, kap = J.kappa;
  auto dzc = [&](int k) { return (CT)g.Dzc(k); };
  auto dzf = [&](int k) { return (CT)g.Dzf(k); };
  auto upper = [&](int k) -> CT {
    if (!J.w_field) return k > Nz - 1 ? CT(0) : -dt * kap / (dzc(k) * dzf(k + 1));
    return k < 1 ? CT(0) : -dt * kap / (dzc(k) * dzf(k));
  };
  auto lower = [&](int k) -> CT {      // Tridiagonal convention: row k holds the coupling of level k+1 to level k
    if (!J.w_field) return k < 1 ? CT(0) : -dt * kap / (dzc(k + 1) * dzf(k + 1));
    return k < 1 ? CT(0) : -dt * kap / (dzc(k + 2) * dzf(k + 1));
  };
  auto diag = [&](int k) -> CT { return CT(1) - dt * CT(0) - upper(k) - lower(k - 1); };
  const CT eps10 = sizeof(CT) == 8 ? CT(10 * 2.220446049250313e-16) : CT(10 * 1.1920929e-7f);
  CT beta = diag(1);
  J.tab[1] = 0.0; J.tab[L + 1] = (double)beta; J.tab[2 * L + 1] = 0.0; J.tab[3 * L + 1] = 1.0;
  for (int k = 2; k <= Nz; ++k) {
    const CT c = upper(k - 1), b = diag(k), a = lower(k - 1);
    const FT tk = (FT)(c / beta);
    beta = b - a * (CT)tk;
    J.tab[k] = (double)a; J.tab[L + k] = (double)beta; J.tab[2 * L + k] = (double)tk;
    J.tab[3 * L + k] = fabs((double)beta) < (double)eps10 ? 0.0 : 1.0;
  }
}

template <typename FT, typename CT>
struct ImplicitJob {
  GridDev<FT> g;
  FT* phi;
  const double* tab;
};

template <typename FT, typename CT>
__global__ void __launch_bounds__(64) implicit_step_kernel(const __grid_constant__ ImplicitJob<FT, CT> J) {
  const GridDev<FT>& g = J.g;
  const int i = blockIdx.x * blockDim.x + threadIdx.x + 1, j = blockIdx.y + 1;
  if (i > g.N[0]) return;
  constexpr int U = 8;
  const int Nz = g.N[2], L = Nz + 2;
  const double* __restrict__ ta = J.tab;
  const double* __restrict__ tb = J.tab + L;
  const double* __restrict__ tt = J.tab + 2 * L;
  const double* __restrict__ tok = J.tab + 3 * L;
  const long long sz = g.sz;
  FT* col = J.phi + g.idx(i, j, 1) - sz;                 // col[k·sz] = ϕ[i, j, k]
  FT prev = (FT)((CT)col[sz] / (CT)__ldg(tb + 1));
  col[sz] = prev;
  for (int k0 = 2; k0 <= Nz; k0 += U) {
    FT v[U];
#pragma unroll
    for (int q = 0; q < U; ++q) if (k0 + q <= Nz) v[q] = col[(long long)(k0 + q) * sz];
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int k = k0 + q;
      if (k <= Nz) {
        if (__ldg(tok + k) != 0.0) v[q] = (FT)(((CT)v[q] - (CT)__ldg(ta + k) * (CT)prev) / (CT)__ldg(tb + k));
        prev = v[q];
        col[(long long)k * sz] = prev;
      }
    }
  }
  // back-substitution ϕ_k -= t_{k+1} ϕ_{k+1}, k = Nz-1 … 1 (prev = ϕ_Nz)
  for (int k0 = Nz - 1; k0 >= 1; k0 -= U) {
    FT v[U];
#pragma unroll
    for (int q = 0; q < U; ++q) if (k0 - q >= 1) v[q] = col[(long long)(k0 - q) * sz];
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const int k = k0 - q;
      if (k >= 1) {
        prev = v[q] - (FT)__ldg(tt + k + 1) * prev;
        col[(long long)k * sz] = prev;
      }
    }
  }
}

// ---- time-stepper substeps ------------------------------------------------------------------------------------
template <typename FT>
struct StepJob {
  GridDev<FT> g;
  FT* U[NH_MAXF];
  const FT* Gn[NH_MAXF];
  FT* Gm[NH_MAXF];
  int face_dim[NH_MAXF];   // 0,1,2 for u,v,w; -1 for tracers
  int nf;
  FT dt, c1, c2;           // RK3: c1 = γ, c2 = ζ; AB2: c1 = χ
  int mode;                // 0: RK3 first stage, 1: RK3 stages 2,3, 2: AB2, 3: AB2 Euler (χ = -1/2), 4: G⁻ ← Gⁿ
};

template <typename FT>
__global__ void __launch_bounds__(256) substep_kernel(const __grid_constant__ StepJob<FT> J) {
  const GridDev<FT>& g = J.g;
  int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  long long n = g.idx(i, j, k);
  for (int f = 0; f < J.nf; ++f) {
    if (J.mode == 4) { J.Gm[f][n] = J.Gn[f][n]; continue; }   // cache_previous_tendencies! over the interior
    int fd = J.face_dim[f];
    if (fd >= 0 && g.topo[fd] == NH_BOUNDED && g.N[fd] > 1 && (fd == 0 ? i : (fd == 1 ? j : k)) == 1) continue;
    FT u = J.U[f][n];
    if (J.mode == 0) {
      u = u + J.dt * J.c1 * J.Gn[f][n];
    } else if (J.mode == 1) {
      u = u + J.dt * (J.c1 * J.Gn[f][n] + J.c2 * J.Gm[f][n]);
    } else if (J.mode == 2) {
      FT G = (FT(1.5) + J.c1) * J.Gn[f][n] - (FT(0.5) + J.c1) * J.Gm[f][n];
      u = u + J.dt * G;
    } else {
      FT G = (FT(1.5) + J.c1) * J.Gn[f][n] - FT(0);
      u = u + J.dt * G;
    }
    J.U[f][n] = u;
  }
}

// ---- divergence, source term, pressure correction -------------------------------------------------------------
// divᶜᶜᶜ — src/Operators/divergence_operators.jl:16-19
template <typename FT>
__device__ __forceinline__ FT div_ccc(const GridDev<FT>& g, const FT* u, const FT* v, const FT* w, int i, int j, int k) {
  FT tx = FT(0), ty = FT(0), tz = FT(0);
  if (g.topo[0] != NH_FLAT) { FT A = area_at(g, 0, 1, k); tx = A * u[g.idx(i + 1, j, k)] - A * u[g.idx(i, j, k)]; }
  if (g.topo[1] != NH_FLAT) { FT A = area_at(g, 1, 1, k); ty = A * v[g.idx(i, j + 1, k)] - A * v[g.idx(i, j, k)]; }
  if (g.topo[2] != NH_FLAT) { FT A = area_at(g, 2, 0, k); tz = A * w[g.idx(i, j, k + 1)] - A * w[g.idx(i, j, k)]; }
  FT total = tx + ty;
  total = total + tz;
  return FT(1) / volume_at(g, 1, k) * total;
}

// _compute_source_term! / _fourier_tridiagonal_source_term! — solve_for_pressure.jl:12-38.  Δt is a Float64 at the
// reference's call sites, so for Float32 grids the division is done in Float64 and rounded on store.
template <typename FT, typename CT>
__global__ void __launch_bounds__(256) source_term_kernel(GridDev<FT> g, const FT* u, const FT* v, const FT* w, double dt,
                                                          int times_dz, CT* rhs) {
  int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  FT d = div_ccc(g, u, v, w, i, j, k);
  if (times_dz) d = g.Dzc(k) * d;
  FT r;
  if (sizeof(FT) == 4) r = (FT)((double)d / dt); else r = (FT)(d / (FT)dt);
  CT out; out.x = r; out.y = FT(0);
  rhs[(long long)(i - 1) + (long long)g.N[0] * ((j - 1) + (long long)g.N[1] * (k - 1))] = out;
}

// the same source term as a dense real (Nx, Ny, Nz) array: input of the fused solver passes (poisson_fast.cuh, RHS mode)
template <typename FT>
__global__ void __launch_bounds__(256) source_term_real_kernel(GridDev<FT> g, const FT* u, const FT* v, const FT* w, double dt,
                                                               int times_dz, FT* rhs) {
  int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  FT d = div_ccc(g, u, v, w, i, j, k);
  if (times_dz) d = g.Dzc(k) * d;
  FT r;
  if (sizeof(FT) == 4) r = (FT)((double)d / dt); else r = (FT)(d / (FT)dt);
  rhs[(long long)(i - 1) + (long long)g.N[0] * ((j - 1) + (long long)g.N[1] * (k - 1))] = r;
}

// _pressure_correct_velocities! — pressure_correction.jl:31-37 (∂p·Δt evaluated in Float64 for Float32 grids)
template <typename FT>
__global__ void __launch_bounds__(256) pressure_correct_kernel(GridDev<FT> g, FT* u, FT* v, FT* w, const FT* p, double dt) {
  int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  long long n = g.idx(i, j, k);
  FT p0 = p[n];
  FT gx = g.topo[0] == NH_FLAT ? FT(0) : (p0 - p[g.idx(i - 1, j, k)]) / g.dx;
  FT gy = g.topo[1] == NH_FLAT ? FT(0) : (p0 - p[g.idx(i, j - 1, k)]) / g.dy;
  FT gz = g.topo[2] == NH_FLAT ? FT(0) : (p0 - p[g.idx(i, j, k - 1)]) / g.Dzf(k);
  if (sizeof(FT) == 4) {
    u[n] = (FT)((double)u[n] - (double)gx * dt);
    v[n] = (FT)((double)v[n] - (double)gy * dt);
    w[n] = (FT)((double)w[n] - (double)gz * dt);
  } else {
    u[n] = u[n] - gx * (FT)dt;
    v[n] = v[n] - gy * (FT)dt;
    w[n] = w[n] - gz * (FT)dt;
  }
}

// _update_hydrostatic_pressure! — update_hydrostatic_pressure.jl:12-20,41-49: one thread per column, top-down
template <typename FT>
__global__ void hydrostatic_pressure_kernel(const __grid_constant__ TendParams<FT> P, FT* pHY) {
  const GridDev<FT>& g = P.g;
  int i0 = g.topo[0] == NH_FLAT ? 1 : 0, j0 = g.topo[1] == NH_FLAT ? 1 : 0;
  int ni = g.topo[0] == NH_FLAT ? g.N[0] : g.N[0] + 2, nj = g.topo[1] == NH_FLAT ? g.N[1] : g.N[1] + 2;
  int ti = blockIdx.x * blockDim.x + threadIdx.x, tj = blockIdx.y * blockDim.y + threadIdx.y;
  if (ti >= ni || tj >= nj) return;
  int i = i0 + ti, j = j0 + tj;
  const int Nz = g.N[2];
  FT b_up = buoyancy_at(P, i, j, Nz + 1);
  FT b_k = buoyancy_at(P, i, j, Nz);
  FT ph = -(FT(0.5) * (b_k + b_up)) * g.Dzf(Nz + 1);
  pHY[g.idx(i, j, Nz)] = ph;
  for (int k = Nz - 1; k >= 1; --k) {
    b_up = b_k;
    b_k = buoyancy_at(P, i, j, k);
    ph = ph - (FT(0.5) * (b_k + b_up)) * g.Dzf(k + 1);
    pHY[g.idx(i, j, k)] = ph;
  }
}

// copy_real_component! — fft_based_poisson_solver.jl:129-137 (dense complex N³ → halo'd real field)
template <typename FT, typename CT>
__global__ void __launch_bounds__(256) copy_real_kernel(GridDev<FT> g, const CT* phi, FT* p) {
  int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  p[g.idx(i, j, k)] = phi[(long long)(i - 1) + (long long)g.N[0] * ((j - 1) + (long long)g.N[1] * (k - 1))].x;
}

// dense (reference layout) ↔ internal layout; region given by logical lower corner and extents
template <typename FT>
__global__ void layout_copy_kernel(GridDev<FT> g, FT* field, FT* dense, int i0, int j0, int k0, int nx, int ny, int nz, int to_field) {
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  long long tot = (long long)nx * ny * nz;
  if (t >= tot) return;
  int a = (int)(t % nx), b = (int)((t / nx) % ny), c = (int)(t / ((long long)nx * ny));
  long long n = g.idx(i0 + a, j0 + b, k0 + c);
  if (to_field) field[n] = dense[t]; else dense[t] = field[n];
}

// max |div U| over the interior (test/utils_for_runtests.jl:139-151): block maxima, finished on the host
template <typename FT>
__global__ void __launch_bounds__(256) max_div_kernel(GridDev<FT> g, const FT* u, const FT* v, const FT* w, double* out) {
  long long tot = (long long)g.N[0] * g.N[1] * g.N[2];
  double m = 0;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < tot; t += (long long)gridDim.x * blockDim.x) {
    int i = (int)(t % g.N[0]) + 1, j = (int)((t / g.N[0]) % g.N[1]) + 1, k = (int)(t / ((long long)g.N[0] * g.N[1])) + 1;
    double d = fabs((double)div_ccc(g, u, v, w, i, j, k));
    m = d > m || d != d ? d : m;
  }
  __shared__ double sm[256];
  sm[threadIdx.x] = m;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) { double o = sm[threadIdx.x + s]; if (o > sm[threadIdx.x] || o != o) sm[threadIdx.x] = o; }
    __syncthreads();
  }
  if (threadIdx.x == 0) out[blockIdx.x] = sm[0];
}
