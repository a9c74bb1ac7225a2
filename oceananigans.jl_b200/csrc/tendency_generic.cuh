// tendency_generic.cuh — pointwise tendency functions for every supported configuration (any topology, stretched z,
// Centered/UpwindBiased/WENO, FPlane, buoyancy, ScalarDiffusivity), in the reference's expression order:
//   src/Models/NonhydrostaticModels/nonhydrostatic_tendency_kernel_functions.jl:47-78,107-138,170-200,228-259
//   src/Advection/momentum_advection_operators.jl:46-83, tracer_advection_operators.jl:30-34
//   src/Advection/upwind_biased_advective_fluxes.jl:23-121, centered_advective_fluxes.jl:15-33
//   src/Coriolis/f_plane.jl:44-46, src/Operators/interpolation_operators.jl:121-131
//   src/TurbulenceClosures/closure_kernel_operators.jl:22-48, velocity_tracer_gradients.jl:25-42
// The fused fast kernels (tendency_fast.cu) cover the periodic WENO5 case; this file is the complete one.
#pragma once
#include "recon.cuh"

template <typename FT>
struct FieldGet {
  const FT* p; const GridDev<FT>* g;
  __device__ __forceinline__ FT operator()(int i, int j, int k) const { return p[g->idx(i, j, k)]; }
};

// Δz at a z-location (1 = centre, 0 = face)
template <typename FT>
__device__ __forceinline__ FT dz_at(const GridDev<FT>& g, int zc, int k) { return zc ? g.Dzc(k) : g.Dzf(k); }

// A_d at a location whose z-stagger is zc: Ax = Δy·Δz, Ay = Δx·Δz, Az = Δx·Δy
template <typename FT>
__device__ __forceinline__ FT area_at(const GridDev<FT>& g, int d, int zc, int k) {
  if (d == 0) return g.dy * dz_at(g, zc, k);
  if (d == 1) return g.dx * dz_at(g, zc, k);
  return g.dx * g.dy;
}
template <typename FT>
__device__ __forceinline__ FT volume_at(const GridDev<FT>& g, int zc, int k) { return (g.dx * g.dy) * dz_at(g, zc, k); }

template <typename FT>
__device__ __forceinline__ FT delta_at(const GridDev<FT>& g, int d, int zc, int k) {
  return d == 0 ? g.dx : (d == 1 ? g.dy : dz_at(g, zc, k));
}

// Ax_qᶠᶜᶜ-style getter: area of velocity component b at its own location times U_b
template <typename FT, int b>
struct AreaVelGet {
  const FT* p; const GridDev<FT>* g;
  __device__ __forceinline__ FT operator()(int i, int j, int k) const {
    return area_at(*g, b, b == 2 ? 0 : 1, k) * p[g->idx(i, j, k)];
  }
};

// advective_momentum_flux_{U_b}{u_a} at the flux location index (i,j,k)
// ZT: reconstruct along z with the stretched WENO tables of TendParams::wz (WENO(grid)); only the per-cell kernel instantiates it
template <typename FT, int a, int b, bool ZT = false>
__device__ __forceinline__ FT momentum_flux(const TendParams<FT>& P, int i, int j, int k) {
  const GridDev<FT>& g = P.g;
  SchemeDev s = P.adv[b];
  if (g.topo[b] == NH_FLAT || s.kind == NH_ADV_NONE) return FT(0);
  constexpr int sym_dim = (a == b) ? b : a;
  constexpr int stag_c = (a == b) ? 1 : 0;
  FieldGet<FT> ua{P.U[a], &g};
  if (s.kind != NH_ADV_CENTERED) {
    AreaVelGet<FT, b> Aq{P.U[b], &g};
    FT ut = symmetric_interp<FT>(P.tab, s, g.topo[sym_dim], g.N[sym_dim], sym_dim, stag_c, i, j, k, Aq);
    FT qR = biased_interp<FT>(P.tab, s, g.topo[b], g.N[b], b, stag_c, ut > FT(0), i, j, k, ua, (ZT && b == 2) ? P.wz[stag_c] : nullptr);
    return ut * qR;
  }
  constexpr int zc = (a != b && (a == 2 || b == 2)) ? 0 : 1;
  FieldGet<FT> ub{P.U[b], &g};
  FT A = area_at(g, b, zc, k);
  return (A * symmetric_interp<FT>(P.tab, s, g.topo[sym_dim], g.N[sym_dim], sym_dim, stag_c, i, j, k, ub)) *
         symmetric_interp<FT>(P.tab, s, g.topo[b], g.N[b], b, stag_c, i, j, k, ua);
}

template <typename FT, int a, int b, bool ZT = false>
__device__ __forceinline__ FT momentum_flux_diff(const TendParams<FT>& P, int i, int j, int k) {
  if (P.g.topo[b] == NH_FLAT) return FT(0);
  constexpr int di = b == 0, dj = b == 1, dk = b == 2;
  if (a == b)  // target is a Face: F[i] - F[i-1]
    return momentum_flux<FT, a, b, ZT>(P, i, j, k) - momentum_flux<FT, a, b, ZT>(P, i - di, j - dj, k - dk);
  return momentum_flux<FT, a, b, ZT>(P, i + di, j + dj, k + dk) - momentum_flux<FT, a, b, ZT>(P, i, j, k);
}

template <typename FT, int a, bool ZT = false>
__device__ __forceinline__ FT div_Uu(const TendParams<FT>& P, int i, int j, int k) {
  FT total = momentum_flux_diff<FT, a, 0, ZT>(P, i, j, k) + momentum_flux_diff<FT, a, 1, ZT>(P, i, j, k);
  total = total + momentum_flux_diff<FT, a, 2, ZT>(P, i, j, k);
  return FT(1) / volume_at(P.g, a == 2 ? 0 : 1, k) * total;
}

template <typename FT, int b, bool ZT = false>
__device__ __forceinline__ FT tracer_flux(const TendParams<FT>& P, const FT* c, int i, int j, int k) {
  const GridDev<FT>& g = P.g;
  SchemeDev s = P.adv[b];
  if (g.topo[b] == NH_FLAT || s.kind == NH_ADV_NONE) return FT(0);
  FT A = area_at(g, b, b == 2 ? 0 : 1, k);
  FT ub = P.U[b][g.idx(i, j, k)];
  FieldGet<FT> cg{c, &g};
  if (s.kind != NH_ADV_CENTERED) {
    FT cR = biased_interp<FT>(P.tab, s, g.topo[b], g.N[b], b, 0, ub > FT(0), i, j, k, cg, (ZT && b == 2) ? P.wz[0] : nullptr);
    return A * ub * cR;
  }
  return (A * ub) * symmetric_interp<FT>(P.tab, s, g.topo[b], g.N[b], b, 0, i, j, k, cg);
}

template <typename FT, bool ZT = false>
__device__ __forceinline__ FT div_Uc(const TendParams<FT>& P, const FT* c, int i, int j, int k) {
  const GridDev<FT>& g = P.g;
  FT tx = g.topo[0] == NH_FLAT ? FT(0) : tracer_flux<FT, 0, ZT>(P, c, i + 1, j, k) - tracer_flux<FT, 0, ZT>(P, c, i, j, k);
  FT ty = g.topo[1] == NH_FLAT ? FT(0) : tracer_flux<FT, 1, ZT>(P, c, i, j + 1, k) - tracer_flux<FT, 1, ZT>(P, c, i, j, k);
  FT tz = g.topo[2] == NH_FLAT ? FT(0) : tracer_flux<FT, 2, ZT>(P, c, i, j, k + 1) - tracer_flux<FT, 2, ZT>(P, c, i, j, k);
  FT total = tx + ty;
  total = total + tz;
  return FT(1) / volume_at(g, 1, k) * total;
}

// ---- Coriolis (FPlane): active-weighted ℑxy ------------------------------------------------------------------
template <typename FT>
__device__ __forceinline__ int inactive_cell(const GridDev<FT>& g, int i, int j, int k) {
  int bad = 0;
  if (g.topo[0] == NH_BOUNDED) bad |= (i < 1) | (i > g.N[0]);
  if (g.topo[1] == NH_BOUNDED) bad |= (j < 1) | (j > g.N[1]);
  if (g.topo[2] == NH_BOUNDED) bad |= (k < 1) | (k > g.N[2]);
  return bad;
}

// not_peripheral_node at the location of velocity component q (face in dim q)
template <typename FT, int q>
struct ActiveGet {
  const GridDev<FT>* g;
  __device__ __forceinline__ FT operator()(int i, int j, int k) const {
    int bad = inactive_cell(*g, i, j, k) | inactive_cell(*g, i - (q == 0), j - (q == 1), k - (q == 2));
    return bad ? FT(0) : FT(1);
  }
};

// ℑxy onto (tx_c, ty_c): ℑy(ℑx ·); each ℑ = FT(0.5)·(a + b), identity when Flat
template <typename FT, typename G>
__device__ __forceinline__ FT interp_x(const GridDev<FT>& g, int tc, int i, int j, int k, const G& get) {
  if (g.topo[0] == NH_FLAT) return get(i, j, k);
  return tc ? FT(0.5) * (get(i, j, k) + get(i + 1, j, k)) : FT(0.5) * (get(i - 1, j, k) + get(i, j, k));
}
template <typename FT, typename G>
__device__ __forceinline__ FT interp_xy(const GridDev<FT>& g, int txc, int tyc, int i, int j, int k, const G& get) {
  if (g.topo[1] == NH_FLAT) return interp_x<FT>(g, txc, i, j, k, get);
  if (tyc) return FT(0.5) * (interp_x<FT>(g, txc, i, j, k, get) + interp_x<FT>(g, txc, i, j + 1, k, get));
  return FT(0.5) * (interp_x<FT>(g, txc, i, j - 1, k, get) + interp_x<FT>(g, txc, i, j, k, get));
}

template <typename FT, int a>
__device__ __forceinline__ FT f_cross_U(const TendParams<FT>& P, int i, int j, int k) {
  if (!P.has_coriolis || a == 2) return FT(0);
  const GridDev<FT>& g = P.g;
  if (a == 0) {   // -f · ℑxyᶠᶜ(v) / active
    FieldGet<FT> v{P.U[1], &g};
    ActiveGet<FT, 1> act{&g};
    FT active = interp_xy<FT>(g, 0, 1, i, j, k, act);
    FT val = interp_xy<FT>(g, 0, 1, i, j, k, v);
    FT aw = active == FT(0) ? FT(0) : val / active;
    return (-P.fcor) * aw;
  }
  FieldGet<FT> u{P.U[0], &g};
  ActiveGet<FT, 0> act{&g};
  FT active = interp_xy<FT>(g, 1, 0, i, j, k, act);
  FT val = interp_xy<FT>(g, 1, 0, i, j, k, u);
  FT aw = active == FT(0) ? FT(0) : val / active;
  return P.fcor * aw;
}

// ---- buoyancy ---------------------------------------------------------------------------------------------
template <typename FT>
__device__ __forceinline__ FT buoyancy_at(const TendParams<FT>& P, int i, int j, int k) {
  long long n = P.g.idx(i, j, k);
  if (P.buoy_kind == NH_BUOYANCY_TRACER) return P.c[P.iT][n];
  return P.grav * (P.alpha * P.c[P.iT][n] - P.beta * P.c[P.iS][n]);
}

// ---- ScalarDiffusivity ------------------------------------------------------------------------------------
// ∂_d q onto a Face (fc = 0: q[idx] - q[idx-1]) or a Centre (fc = 1: q[idx+1] - q[idx]), divided by Δ at the target
template <typename FT, int d>
__device__ __forceinline__ FT deriv(const GridDev<FT>& g, const FT* q, int target_c, int zc, int i, int j, int k) {
  if (g.topo[d] == NH_FLAT) return FT(0);
  constexpr int di = d == 0, dj = d == 1, dk = d == 2;
  FT num = target_c ? q[g.idx(i + di, j + dj, k + dk)] - q[g.idx(i, j, k)]
                    : q[g.idx(i, j, k)] - q[g.idx(i - di, j - dj, k - dk)];
  return num / delta_at(g, d, zc, k);
}

// Σ_ab at its natural location (ccc when a == b, Face in both a and b otherwise)
template <typename FT, int a, int b>
__device__ __forceinline__ FT strain(const TendParams<FT>& P, int i, int j, int k) {
  const GridDev<FT>& g = P.g;
  if (a == b) return deriv<FT, a>(g, P.U[a], 1, 1, i, j, k);
  constexpr int lo = a < b ? a : b, hi = a < b ? b : a;
  constexpr int zc = (hi == 2) ? 0 : 1;     // location is a Face in z iff z is one of the two dims
  return FT(0.5) * (deriv<FT, hi>(g, P.U[lo], 0, zc, i, j, k) + deriv<FT, lo>(g, P.U[hi], 0, zc, i, j, k));
}

// ν at the location of Σ_ab: the constant, or the eddy viscosity νₑ[i,j,k] at ccc / ℑ_hi(ℑ_lo νₑ) on the edge
// (abstract_scalar_diffusivity_closure.jl:60-75 with a Center-located viscosity field; interpolation_operators.jl:45-53)
template <typename FT, int d>
__device__ __forceinline__ FT interp_to_face(const GridDev<FT>& g, const FT* q, int i, int j, int k) {
  if (g.topo[d] == NH_FLAT) return q[g.idx(i, j, k)];
  constexpr int di = d == 0, dj = d == 1, dk = d == 2;
  return FT(0.5) * (q[g.idx(i - di, j - dj, k - dk)] + q[g.idx(i, j, k)]);
}

template <typename FT, int a, int b>
__device__ __forceinline__ FT nu_at(const TendParams<FT>& P, int i, int j, int k) {
  if (!P.nue) return P.nu;
  const GridDev<FT>& g = P.g;
  if (a == b) return P.nue[g.idx(i, j, k)];
  constexpr int lo = a < b ? a : b, hi = a < b ? b : a;
  if (g.topo[hi] == NH_FLAT) return interp_to_face<FT, lo>(g, P.nue, i, j, k);
  constexpr int di = hi == 0, dj = hi == 1, dk = hi == 2;
  return FT(0.5) * (interp_to_face<FT, lo>(g, P.nue, i - di, j - dj, k - dk) + interp_to_face<FT, lo>(g, P.nue, i, j, k));
}

template <typename FT, int a, int b>
__device__ __forceinline__ FT viscous_Aflux(const TendParams<FT>& P, int i, int j, int k) {
  constexpr int zc = (a != b && (a == 2 || b == 2)) ? 0 : 1;
  FT flux = FT(-2) * (nu_at<FT, a, b>(P, i, j, k) * strain<FT, a, b>(P, i, j, k));
  if (b == 2 && P.vitd && !(k == 1 || k == P.g.N[2] + 1)) {
    // VerticallyImplicitTimeDiscretization (abstract_scalar_diffusivity_closure.jl:219-246): off the walls only the part that
    // the tridiagonal solve does not carry stays explicit: -(ν·∂x w) for u, -(ν·∂y w) for v, nothing for w
    if (a == 2) flux = FT(0);
    else flux = -(nu_at<FT, a, b>(P, i, j, k) * deriv<FT, a == 2 ? 0 : a>(P.g, P.U[2], 0, 0, i, j, k));
  }
  return area_at(P.g, b, zc, k) * flux;
}

template <typename FT, int a, int b>
__device__ __forceinline__ FT viscous_diff(const TendParams<FT>& P, int i, int j, int k) {
  if (P.g.topo[b] == NH_FLAT) return FT(0);
  constexpr int di = b == 0, dj = b == 1, dk = b == 2;
  if (a == b) return viscous_Aflux<FT, a, b>(P, i, j, k) - viscous_Aflux<FT, a, b>(P, i - di, j - dj, k - dk);
  return viscous_Aflux<FT, a, b>(P, i + di, j + dj, k + dk) - viscous_Aflux<FT, a, b>(P, i, j, k);
}

template <typename FT, int a>
__device__ __forceinline__ FT div_viscous_flux(const TendParams<FT>& P, int i, int j, int k) {
  FT total = viscous_diff<FT, a, 0>(P, i, j, k) + viscous_diff<FT, a, 1>(P, i, j, k);
  total = total + viscous_diff<FT, a, 2>(P, i, j, k);
  return FT(1) / volume_at(P.g, a == 2 ? 0 : 1, k) * total;
}

// diffusive_flux = -κ·∂c with κ the constant, ℑ_b(νₑ) / Pr (Smagorinskys/smagorinsky.jl:135-137) or ℑ_b(κₑ[n]) (AMD)
template <typename FT, int b>
__device__ __forceinline__ FT diffusive_Aflux(const TendParams<FT>& P, const FT* c, int n, int i, int j, int k) {
  constexpr int zc = b == 2 ? 0 : 1;
  const FT kappa = P.nue ? interp_to_face<FT, b>(P.g, P.kape[n], i, j, k) / P.Pr[n] : P.kappa[n];
  FT flux = -kappa * deriv<FT, b>(P.g, c, 0, zc, i, j, k);
  if (b == 2 && P.vitd && !(k == 1 || k == P.g.N[2] + 1)) flux = FT(0);     // diffusive_flux_z is implicit off the walls (:248-252)
  return area_at(P.g, b, zc, k) * flux;
}

template <typename FT>
__device__ __forceinline__ FT div_diffusive_flux(const TendParams<FT>& P, const FT* c, int kappa, int i, int j, int k) {
  const GridDev<FT>& g = P.g;
  FT tx = g.topo[0] == NH_FLAT ? FT(0) : diffusive_Aflux<FT, 0>(P, c, kappa, i + 1, j, k) - diffusive_Aflux<FT, 0>(P, c, kappa, i, j, k);
  FT ty = g.topo[1] == NH_FLAT ? FT(0) : diffusive_Aflux<FT, 1>(P, c, kappa, i, j + 1, k) - diffusive_Aflux<FT, 1>(P, c, kappa, i, j, k);
  FT tz = g.topo[2] == NH_FLAT ? FT(0) : diffusive_Aflux<FT, 2>(P, c, kappa, i, j, k + 1) - diffusive_Aflux<FT, 2>(P, c, kappa, i, j, k);
  FT total = tx + ty;
  total = total + tz;
  return FT(1) / volume_at(g, 1, k) * total;
}

// ---- Smagorinsky / Smagorinsky-Lilly eddy viscosity ---------------------------------------------------------------
// _compute_smagorinsky_viscosity! over :xyz — Smagorinskys/smagorinsky.jl:88-105, lilly_coefficient.jl:125-139,
// scale_invariant_operators.jl:10-13 (ΣᵢⱼΣᵢⱼᶜᶜᶜ), BuoyancyFormulations ∂z_b.  Δᶠ² = cbrt(ΔxΔyΔz)² comes from a host
// table (one entry per k) with a correctly rounded cube root (grid_host.cpp), the same bits as the oracle's.
template <typename FT, int lo, int hi>
__device__ __forceinline__ FT strain_sq_to_center(const TendParams<FT>& P, int i, int j, int k) {
  // ℑ_hi(ℑ_lo Σ_lohi²) onto Centers: 0.5·(X(q) + X(q+1)) along hi of X = 0.5·(Σ²(p) + Σ²(p+1)) along lo; identity if Flat
  const GridDev<FT>& g = P.g;
  constexpr int li = lo == 0, lj = lo == 1, lk = lo == 2, hi_i = hi == 0, hi_j = hi == 1, hi_k = hi == 2;
  auto sq = [&](int i2, int j2, int k2) { FT x = strain<FT, lo, hi>(P, i2, j2, k2); return x * x; };
  auto X = [&](int i2, int j2, int k2) {
    if (g.topo[lo] == NH_FLAT) return sq(i2, j2, k2);
    return FT(0.5) * (sq(i2, j2, k2) + sq(i2 + li, j2 + lj, k2 + lk));
  };
  if (g.topo[hi] == NH_FLAT) return X(i, j, k);
  return FT(0.5) * (X(i, j, k) + X(i + hi_i, j + hi_j, k + hi_k));
}

template <typename FT>
__device__ __forceinline__ FT dz_b_at(const TendParams<FT>& P, int i, int j, int k) {
  if (P.buoy_kind == NH_BUOYANCY_NONE) return FT(0);
  if (P.buoy_kind == NH_BUOYANCY_TRACER) return deriv<FT, 2>(P.g, P.c[P.iT], 0, 0, i, j, k);
  return P.grav * (P.alpha * deriv<FT, 2>(P.g, P.c[P.iT], 0, 0, i, j, k) - P.beta * deriv<FT, 2>(P.g, P.c[P.iS], 0, 0, i, j, k));
}

template <typename FT>
struct SmagJob {
  TendParams<FT> P;
  FT* nue;
  const FT* df2;          // Δᶠ·Δᶠ at k (index k-1)
  int lilly;
  FT C, Cb;               // SmagorinskyLilly: converted to FT
  double C64;             // Smagorinsky(coefficient = number): NOT converted (the product is formed in Float64)
};

template <typename FT>
__global__ void __launch_bounds__(128) smagorinsky_kernel(const __grid_constant__ SmagJob<FT> J) {
  const TendParams<FT>& P = J.P;
  const GridDev<FT>& g = P.g;
  int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  const FT s11 = strain<FT, 0, 0>(P, i, j, k), s22 = strain<FT, 1, 1>(P, i, j, k), s33 = strain<FT, 2, 2>(P, i, j, k);
  FT S2 = s11 * s11 + s22 * s22 + s33 * s33;
  S2 = S2 + FT(2) * strain_sq_to_center<FT, 0, 1>(P, i, j, k);
  S2 = S2 + FT(2) * strain_sq_to_center<FT, 0, 2>(P, i, j, k);
  S2 = S2 + FT(2) * strain_sq_to_center<FT, 1, 2>(P, i, j, k);
  const FT df2 = J.df2[k - 1];
  const FT root = sqrt(FT(2) * S2);
  FT nu;
  if (J.lilly) {
    const FT N2 = g.topo[2] == NH_FLAT ? dz_b_at(P, i, j, k) : FT(0.5) * (dz_b_at(P, i, j, k) + dz_b_at(P, i, j, k + 1));
    const FT N2p = N2 > FT(0) ? N2 : FT(0);            // max(zero(FT), N²)
    const FT ratio = J.Cb * N2p / S2;
    const FT sig2 = FT(1) - (ratio < FT(1) ? ratio : (ratio != ratio ? ratio : FT(1)));   // min(1, NaN) is NaN in Julia
    const FT sig = S2 == FT(0) ? FT(0) : sqrt(sig2);
    const FT cs2 = sig * (J.C * J.C);
    nu = cs2 * df2 * root;
  } else {
    nu = (FT)((J.C64 * J.C64) * (double)df2 * (double)root);
  }
  J.nue[g.idx(i, j, k)] = nu;
}

// ---- the tendencies ---------------------------------------------------------------------------------------
// `div` = div_𝐯u_a at the cell (evaluated by the caller: per cell, or from face fluxes staged in shared memory)
template <typename FT, int a>
__device__ __forceinline__ FT velocity_tendency(const TendParams<FT>& P, int i, int j, int k, FT div, FT closure_div) {
  const GridDev<FT>& g = P.g;
  FT G = -div;
  if (a == 2 && P.w_buoyancy) {
    FT bz = g.topo[2] == NH_FLAT ? buoyancy_at(P, i, j, k)
                                 : FT(0.5) * (buoyancy_at(P, i, j, k - 1) + buoyancy_at(P, i, j, k));
    G = G + bz;
  } else {
    G = G + FT(0);
  }
  G = G - f_cross_U<FT, a>(P, i, j, k);
  if (a < 2) {
    if (P.pHY) G = G - deriv<FT, a>(g, P.pHY, 0, 1, i, j, k);
    else G = G - FT(0);
  }
  if (P.has_closure) G = G - closure_div;        // ∂ⱼτ_aⱼ, evaluated by the caller
  else G = G - FT(0);
  return G;
}

template <typename FT>
__device__ __forceinline__ FT tracer_tendency(const TendParams<FT>& P, int n, int i, int j, int k, FT div, FT closure_div) {
  FT G = -div;
  if (P.has_closure) G = G - closure_div;        // ∇·q_c, evaluated by the caller
  else G = G - FT(0);
  return G;
}

// One thread per cell of 1:Nx × 1:Ny × 1:Nz; Face fields skip index 1 in Bounded dims (exclude_periphery,
// src/Utils/kernel_launching.jl:90-111; compute_nonhydrostatic_tendencies.jl:104-136).
template <typename FT, bool ZT = false>
__global__ void __launch_bounds__(128) tendency_generic_kernel(const __grid_constant__ TendParams<FT> P, int what) {
  const GridDev<FT>& g = P.g;
  int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  long long n = g.idx(i, j, k);
  if (what == 0) {
    if (!(g.topo[0] == NH_BOUNDED && g.N[0] > 1 && i == 1)) P.GU[0][n] = velocity_tendency<FT, 0>(P, i, j, k, div_Uu<FT, 0, ZT>(P, i, j, k), P.has_closure ? div_viscous_flux<FT, 0>(P, i, j, k) : FT(0));
  } else if (what == 1) {
    if (!(g.topo[1] == NH_BOUNDED && g.N[1] > 1 && j == 1)) P.GU[1][n] = velocity_tendency<FT, 1>(P, i, j, k, div_Uu<FT, 1, ZT>(P, i, j, k), P.has_closure ? div_viscous_flux<FT, 1>(P, i, j, k) : FT(0));
  } else if (what == 2) {
    if (!(g.topo[2] == NH_BOUNDED && g.N[2] > 1 && k == 1)) P.GU[2][n] = velocity_tendency<FT, 2>(P, i, j, k, div_Uu<FT, 2, ZT>(P, i, j, k), P.has_closure ? div_viscous_flux<FT, 2>(P, i, j, k) : FT(0));
  } else {
    int t = what - 3;
    P.Gc[t][n] = tracer_tendency<FT>(P, t, i, j, k, div_Uc<FT, ZT>(P, P.c[t], i, j, k), P.has_closure ? div_diffusive_flux<FT>(P, P.c[t], t, i, j, k) : FT(0));
  }
}


// ---- tiled variant: every advective face flux of a TG_TX×TG_TY×TG_TZ tile is evaluated ONCE into shared memory --------
// Same flux functions, same differences and the same association of the sums as the per-cell kernel above (which evaluates
// each face twice, once from either neighbour, like the reference's kernels), so the results are bit-identical.  A second
// pass through the same shared memory stages the closure's viscous / diffusive face fluxes; the remaining terms (buoyancy,
// Coriolis, hydrostatic pressure gradient) are evaluated per cell as before.  Block = TG_TX×TG_TY threads, each marching the tile's TG_TZ planes.
#define TG_TX 32
#ifndef TG_TY
#define TG_TY 8          // 32×8×4 tile, 256 threads: 256³ (P,P,B) tendencies 3.50 → 3.24 ms per substep against 32×4×4 (32×4×8: 3.51)
#endif
#ifndef TG_TZ
#define TG_TZ 8          // = TG_TY: every thread owns one line of faces in the y- and in the z-marching of the fast flux stage
#endif
#define TG_SMEM_ELEMS (TG_TZ * TG_TY * (TG_TX + 1) + TG_TZ * (TG_TY + 1) * TG_TX + (TG_TZ + 1) * TG_TY * TG_TX)
#ifndef TG_MINBLOCKS
#define TG_MINBLOCKS 4      // 64 registers, four CTAs per SM: 256³ LES tendencies 6.15 → 4.76 ms per substep against 128 registers / two CTAs
#endif

// ---- fast flux stage of the tiled kernel ---------------------------------------------------------------------------------
// Interior tiles of a WENO(order = 5) model on a grid that is Periodic and regular in x, y, with a regular or stretched z (Periodic,
// or Bounded with the tile at least four cells away from either wall, where topologically_conditional_interpolation keeps the full order):
// the advective face fluxes of the tile are evaluated from a six-point window — read by pointer stepping along x, carried in
// registers along y and z (one new value per face) — with 32-bit offsets and no kind / level dispatch.  The arithmetic is that
// of momentum_flux / tracer_flux above (symmetric_interp's four-point sum with the area folded in per point, biased_interp's
// upwind selection, weno5_from_q), so the staged values are the same bits.
// KIND 0: momentum flux across a direction other than the component's (advecting velocity interpolated along the component's
// direction); 1: along the component's own direction (flux at Centres, advecting velocity from the window itself); 2: tracer.
// A0 … A3: the face area at each of the four advecting-velocity points (they differ only when that interpolation runs along a
// stretched z: Ax = Δy·Δzᶜ[k′]); KIND 1 and 2 use A0 alone.
template <typename FT, int KIND, int VARIANT = 0>
__device__ __forceinline__ FT tg_window_flux(const SchemeTables<FT>& T, FT S0, FT S1, FT S2, FT S3, FT S4, FT S5, FT v0, FT v1,
                                             FT v2, FT v3, FT A0, FT A1, FT A2, FT A3) {
  FT ut;
  bool left;
  if (KIND == 2) { ut = A0 * v0; left = v0 > FT(0); }
  else {
    const FT u0 = KIND == 1 ? S1 : v0, u1 = KIND == 1 ? S2 : v1, u2 = KIND == 1 ? S3 : v2, u3 = KIND == 1 ? S4 : v3;
    ut = T.sym[2][3] * ((KIND == 1 ? A0 : A0) * u0);
    ut = ut + T.sym[2][2] * ((KIND == 1 ? A0 : A1) * u1);
    ut = ut + T.sym[2][1] * ((KIND == 1 ? A0 : A2) * u2);
    ut = ut + T.sym[2][0] * ((KIND == 1 ? A0 : A3) * u3);
    left = ut > FT(0);
  }
  const FT q0 = left ? S0 : S5, q1 = left ? S1 : S4, q2 = left ? S2 : S3, q3 = left ? S3 : S2, q4 = left ? S4 : S1;
  // VARIANT 1 (exact_arithmetic = 0; instantiated in tendency_fma.cu only): the regrouped, quotient-free weights of recon.cuh
  return ut * (VARIANT == 0 ? weno5_from_q<FT>(T, q0, q1, q2, q3, q4) : weno5_contracted<FT>(T, q0, q1, q2, q3, q4));
}

// `count` consecutive fluxes along a direction with stride sb, the first at flux location n: window carried in registers.
// pb: advecting velocity U_b, sa: stride of the component's direction (KIND 0)
template <typename FT, int KIND, int VARIANT = 0>
__device__ __forceinline__ void tg_march(const SchemeTables<FT>& T, const FT* __restrict__ fa, const FT* __restrict__ ub, int n,
                                         int sb, int sa, FT A0, FT A1, FT A2, FT A3, int count, FT* out, int ostride) {
  const FT* p = fa + (n + ((KIND == 1 ? 1 : 0) - 3) * sb);
  FT S0 = p[0], S1 = p[sb], S2 = p[2 * sb], S3 = p[3 * sb], S4 = p[4 * sb];
  p += 5 * sb;
  const FT* pb = ub + (KIND == 0 ? n - 2 * sa : n);
#pragma unroll 1
  for (int c = 0; c < count; ++c) {
    const FT S5 = *p;
    FT v0 = FT(0), v1 = FT(0), v2 = FT(0), v3 = FT(0);
    if (KIND == 0) { v0 = pb[0]; v1 = pb[sa]; v2 = pb[2 * sa]; v3 = pb[3 * sa]; }
    if (KIND == 2) v0 = pb[0];
    *out = tg_window_flux<FT, KIND, VARIANT>(T, S0, S1, S2, S3, S4, S5, v0, v1, v2, v3, A0, A1, A2, A3);
    out += ostride; p += sb; pb += sb;
    S0 = S1; S1 = S2; S2 = S3; S3 = S4; S4 = S5;
  }
}

template <typename FT, int a, int VARIANT = 0>    // a = 0,1,2: velocity component; a = 3: tracer c
__device__ __forceinline__ void tg_stage_fast(const TendParams<FT>& P, const FT* __restrict__ c, FT* Fx, FT* Fy, FT* Fz, int i0,
                                              int j0, int k0, int nx, int ny, int nz, int tid) {
  static_assert(TG_TX * TG_TY == TG_TX * TG_TZ, "one thread per line of faces in both marching directions");
  const GridDev<FT>& g = P.g;
  const int sy = (int)g.sy, sz = (int)g.sz;
  const FT* __restrict__ fa = a == 3 ? c : P.U[a];
  const int sa = a == 0 ? 1 : (a == 1 ? sy : sz);
  const FT Az = g.dx * g.dy;
  constexpr int NT = TG_TX * TG_TY;
  // face areas Ax = Δy·Δzᶜ[k], Ay = Δx·Δzᶜ[k] of the advecting velocities u, v (they live at z-Centres); the x / y fluxes of w
  // interpolate them along z, i.e. over the four levels k-2 … k+1
  // x-direction: flux slot (li, lj, lk), li = 0 … TX
  for (int t = tid; t < (TG_TX + 1) * TG_TY * TG_TZ; t += NT) {
    const int li = t % (TG_TX + 1), r = t / (TG_TX + 1), lj = r % TG_TY, lk = r / TG_TY, k = k0 + min(lk, nz - 1);
    const int n = g.idx32(i0 + min(li, nx) - (a == 0 ? 1 : 0), j0 + min(lj, ny - 1), k);
    const FT* w = fa + n + (a == 0 ? -2 : -3);
    const FT S0 = w[0], S1 = w[1], S2 = w[2], S3 = w[3], S4 = w[4], S5 = w[5];
    const FT Ax = g.dy * g.Dzc(k);
    FT f;
    if (a == 0) f = tg_window_flux<FT, 1, VARIANT>(P.tab, S0, S1, S2, S3, S4, S5, FT(0), FT(0), FT(0), FT(0), Ax, Ax, Ax, Ax);
    else if (a == 3) f = tg_window_flux<FT, 2, VARIANT>(P.tab, S0, S1, S2, S3, S4, S5, P.U[0][n], FT(0), FT(0), FT(0), Ax, Ax, Ax, Ax);
    else {
      const FT* q = P.U[0] + n - 2 * sa;
      if (a == 2) f = tg_window_flux<FT, 0, VARIANT>(P.tab, S0, S1, S2, S3, S4, S5, q[0], q[sa], q[2 * sa], q[3 * sa], g.dy * g.Dzc(k - 2),
                                            g.dy * g.Dzc(k - 1), Ax, g.dy * g.Dzc(k + 1));
      else f = tg_window_flux<FT, 0, VARIANT>(P.tab, S0, S1, S2, S3, S4, S5, q[0], q[sa], q[2 * sa], q[3 * sa], Ax, Ax, Ax, Ax);
    }
    Fx[(lk * TG_TY + lj) * (TG_TX + 1) + li] = f;
  }
  // y- and z-direction: one thread per line of faces
  const int li = tid & (TG_TX - 1), lo = tid / TG_TX, ci = i0 + min(li, nx - 1);
  {
    const int k = k0 + min(lo, nz - 1);
    const int n0 = g.idx32(ci, j0, k);
    const FT Ay = g.dx * g.Dzc(k);
    FT* out = Fy + lo * (TG_TY + 1) * TG_TX + li;
    if (a == 1) tg_march<FT, 1, VARIANT>(P.tab, fa, fa, n0 - sy, sy, sa, Ay, Ay, Ay, Ay, ny + 1, out, TG_TX);
    else if (a == 3) tg_march<FT, 2, VARIANT>(P.tab, fa, P.U[1], n0, sy, sa, Ay, Ay, Ay, Ay, ny + 1, out, TG_TX);
    else if (a == 2) tg_march<FT, 0, VARIANT>(P.tab, fa, P.U[1], n0, sy, sa, g.dx * g.Dzc(k - 2), g.dx * g.Dzc(k - 1), Ay, g.dx * g.Dzc(k + 1), ny + 1, out, TG_TX);
    else tg_march<FT, 0, VARIANT>(P.tab, fa, P.U[1], n0, sy, sa, Ay, Ay, Ay, Ay, ny + 1, out, TG_TX);
  }
  {
    const int n0 = g.idx32(ci, j0 + min(lo, ny - 1), k0);
    FT* out = Fz + lo * TG_TX + li;
    if (a == 2) tg_march<FT, 1, VARIANT>(P.tab, fa, fa, n0 - sz, sz, sa, Az, Az, Az, Az, nz + 1, out, TG_TY * TG_TX);
    else if (a == 3) tg_march<FT, 2, VARIANT>(P.tab, fa, P.U[2], n0, sz, sa, Az, Az, Az, Az, nz + 1, out, TG_TY * TG_TX);
    else tg_march<FT, 0, VARIANT>(P.tab, fa, P.U[2], n0, sz, sa, Az, Az, Az, Az, nz + 1, out, TG_TY * TG_TX);
  }
}

// The same for the closure pass of a constant-coefficient ScalarDiffusivity (no eddy-viscosity field): viscous_Aflux /
// diffusive_Aflux of every face of an interior tile with 32-bit offsets, the regular spacings as constants and the IEEE
// divisions by Δx, Δy, Δz done with one refined reciprocal per spacing (recon.cuh: the same bits as `/`).
template <typename FT>
struct TgSpacing { FT d[3]; FT y[3]; bool ok[3]; };
template <typename FT>
__device__ __forceinline__ FT tg_div(FT a, const TgSpacing<FT>& S, int d) { return a / S.d[d]; }
template <>
__device__ __forceinline__ double tg_div<double>(double a, const TgSpacing<double>& S, int d) { return nh_div_checked(a, S.d[d], S.y[d], S.ok[d]); }

template <typename FT, int a, int b>
__device__ __forceinline__ FT tg_closure_flux(const TendParams<FT>& P, const FT* __restrict__ c, int t, const TgSpacing<FT>& S, int n,
                                              const int (&st)[3], FT Ab) {
  FT flux;
  if (a == 3) {
    if (b == 2 && P.vitd) flux = FT(0);
    else flux = -P.kappa[t] * tg_div<FT>(c[n] - c[n - st[b]], S, b);
  } else if (b == 2 && P.vitd) {
    if (a == 2) flux = FT(0);
    else flux = -(P.nu * tg_div<FT>(P.U[2][n] - P.U[2][n - st[a == 3 ? 0 : a]], S, a == 3 ? 0 : a));
  } else if (a == b) {
    flux = FT(-2) * (P.nu * tg_div<FT>(P.U[a == 3 ? 0 : a][n + st[b]] - P.U[a == 3 ? 0 : a][n], S, b));
  } else {
    constexpr int av = a == 3 ? 0 : a, lo = av < b ? av : b, hi = av < b ? b : av;
    const FT d1 = tg_div<FT>(P.U[lo][n] - P.U[lo][n - st[hi]], S, hi), d2 = tg_div<FT>(P.U[hi][n] - P.U[hi][n - st[lo]], S, lo);
    flux = FT(-2) * (P.nu * (FT(0.5) * (d1 + d2)));
  }
  return Ab * flux;
}

template <typename FT, int a>
__device__ __forceinline__ void tg_stage_closure_fast(const TendParams<FT>& P, const FT* __restrict__ c, int t, FT* Fx, FT* Fy, FT* Fz,
                                                      int i0, int j0, int k0, int nx, int ny, int nz, int tid) {
  const GridDev<FT>& g = P.g;
  const int st[3] = {1, (int)g.sy, (int)g.sz};
  TgSpacing<FT> S;
  S.d[0] = g.dx; S.d[1] = g.dy; S.d[2] = g.dz;
  for (int e = 0; e < 3; ++e) { S.y[e] = nh_rcp_refined(S.d[e]); S.ok[e] = sizeof(FT) == 8 ? nh_den_ok((double)S.d[e]) : false; }
  const FT Ax = g.dy * g.dz, Ay = g.dx * g.dz, Az = g.dx * g.dy;
  constexpr int NT = TG_TX * TG_TY;
  for (int s = tid; s < (TG_TX + 1) * TG_TY * TG_TZ; s += NT) {
    const int li = s % (TG_TX + 1), r = s / (TG_TX + 1), lj = r % TG_TY, lk = r / TG_TY;
    const int n = g.idx32(i0 + min(li, nx) - (a == 0 ? 1 : 0), j0 + min(lj, ny - 1), k0 + min(lk, nz - 1));
    Fx[s] = tg_closure_flux<FT, a, 0>(P, c, t, S, n, st, Ax);
  }
  for (int s = tid; s < TG_TX * (TG_TY + 1) * TG_TZ; s += NT) {
    const int li = s % TG_TX, r = s / TG_TX, lj = r % (TG_TY + 1), lk = r / (TG_TY + 1);
    const int n = g.idx32(i0 + min(li, nx - 1), j0 + min(lj, ny) - (a == 1 ? 1 : 0), k0 + min(lk, nz - 1));
    Fy[s] = tg_closure_flux<FT, a, 1>(P, c, t, S, n, st, Ay);
  }
  for (int s = tid; s < TG_TX * TG_TY * (TG_TZ + 1); s += NT) {
    const int li = s % TG_TX, r = s / TG_TX, lj = r % TG_TY, lk = r / TG_TY;
    const int n = g.idx32(i0 + min(li, nx - 1), j0 + min(lj, ny - 1), k0 + min(lk, nz) - (a == 2 ? 1 : 0));
    Fz[s] = tg_closure_flux<FT, a, 2>(P, c, t, S, n, st, Az);
  }
}

template <typename FT, int a, int VARIANT = 0>    // a = 0,1,2: velocity component; a = 3: tracer P.c[t]
__device__ __forceinline__ void tiled_tendency(const TendParams<FT>& P, int t, FT* sm, bool fast) {
  const GridDev<FT>& g = P.g;
  typedef FT (*FxT)[TG_TY][TG_TX + 1];
  typedef FT (*FyT)[TG_TY + 1][TG_TX];
  typedef FT (*FzT)[TG_TY][TG_TX];
  FxT Fx = reinterpret_cast<FxT>(sm);
  FyT Fy = reinterpret_cast<FyT>(sm + TG_TZ * TG_TY * (TG_TX + 1));
  FzT Fz = reinterpret_cast<FzT>(sm + TG_TZ * TG_TY * (TG_TX + 1) + TG_TZ * (TG_TY + 1) * TG_TX);
  const int i0 = blockIdx.x * TG_TX + 1, j0 = blockIdx.y * TG_TY + 1, k0 = blockIdx.z * TG_TZ + 1;
  const int nx = min(TG_TX, g.N[0] - i0 + 1), ny = min(TG_TY, g.N[1] - j0 + 1), nz = min(TG_TZ, g.N[2] - k0 + 1);
  const int tid = threadIdx.y * TG_TX + threadIdx.x;
  constexpr int NT = TG_TX * TG_TY;
  const FT* c = a == 3 ? P.c[t] : nullptr;
  // a momentum / viscous flux along its own component's direction lives at Centers (difference F[i] - F[i-1]): slots
  // i0-1 … i0+nx-1; every other flux lives at Faces (F[i+1] - F[i]): slots i0 … i0+nx
  const int ox = a == 0 ? -1 : 0, oy = a == 1 ? -1 : 0, oz = a == 2 ? -1 : 0;
  constexpr int av = a == 3 ? 0 : a;           // keeps the momentum templates well-formed in the tracer instantiation
  const int li = threadIdx.x, lj = threadIdx.y;
  const bool mine = li < nx && lj < ny;
  FT divs[2][TG_TZ];                           // [0]: advective divergence, [1]: closure divergence, per plane of the tile
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    if (pass == 1 && !P.has_closure) break;
    if (pass == 1) __syncthreads();            // everyone has read the advective fluxes
    if (pass == 0 && fast) {
      tg_stage_fast<FT, a, VARIANT>(P, c, &Fx[0][0][0], &Fy[0][0][0], &Fz[0][0][0], i0, j0, k0, nx, ny, nz, tid);
    } else if (pass == 1 && fast && !P.nue && g.z_regular) {
      tg_stage_closure_fast<FT, a>(P, c, t, &Fx[0][0][0], &Fy[0][0][0], &Fz[0][0][0], i0, j0, k0, nx, ny, nz, tid);
    } else {
    if (g.topo[0] != NH_FLAT)
      for (int s = tid; s < (TG_TX + 1) * TG_TY * TG_TZ; s += NT) {
        const int fi = s % (TG_TX + 1), r = s / (TG_TX + 1), fj = r % TG_TY, fk = r / TG_TY;
        if (fi > nx || fj >= ny || fk >= nz) continue;
        const int i = i0 + fi + ox, j = j0 + fj, k = k0 + fk;
        if (a == 0 && g.topo[0] == NH_BOUNDED && g.N[0] > 1 && i < 1) continue;   // only the (skipped) periphery cell would read it
        if (pass == 0) Fx[fk][fj][fi] = a == 3 ? tracer_flux<FT, 0>(P, c, i, j, k) : momentum_flux<FT, av, 0>(P, i, j, k);
        else Fx[fk][fj][fi] = a == 3 ? diffusive_Aflux<FT, 0>(P, c, t, i, j, k) : viscous_Aflux<FT, av, 0>(P, i, j, k);
      }
    if (g.topo[1] != NH_FLAT)
      for (int s = tid; s < TG_TX * (TG_TY + 1) * TG_TZ; s += NT) {
        const int fi = s % TG_TX, r = s / TG_TX, fj = r % (TG_TY + 1), fk = r / (TG_TY + 1);
        if (fi >= nx || fj > ny || fk >= nz) continue;
        const int i = i0 + fi, j = j0 + fj + oy, k = k0 + fk;
        if (a == 1 && g.topo[1] == NH_BOUNDED && g.N[1] > 1 && j < 1) continue;
        if (pass == 0) Fy[fk][fj][fi] = a == 3 ? tracer_flux<FT, 1>(P, c, i, j, k) : momentum_flux<FT, av, 1>(P, i, j, k);
        else Fy[fk][fj][fi] = a == 3 ? diffusive_Aflux<FT, 1>(P, c, t, i, j, k) : viscous_Aflux<FT, av, 1>(P, i, j, k);
      }
    if (g.topo[2] != NH_FLAT)
      for (int s = tid; s < TG_TX * TG_TY * (TG_TZ + 1); s += NT) {
        const int fi = s % TG_TX, r = s / TG_TX, fj = r % TG_TY, fk = r / TG_TY;
        if (fi >= nx || fj >= ny || fk > nz) continue;
        const int i = i0 + fi, j = j0 + fj, k = k0 + fk + oz;
        if (a == 2 && g.topo[2] == NH_BOUNDED && g.N[2] > 1 && k < 1) continue;
        if (pass == 0) Fz[fk][fj][fi] = a == 3 ? tracer_flux<FT, 2>(P, c, i, j, k) : momentum_flux<FT, av, 2>(P, i, j, k);
        else Fz[fk][fj][fi] = a == 3 ? diffusive_Aflux<FT, 2>(P, c, t, i, j, k) : viscous_Aflux<FT, av, 2>(P, i, j, k);
      }
    }
    __syncthreads();
    if (mine) {
#pragma unroll
      for (int lk = 0; lk < TG_TZ; ++lk) {
        if (lk >= nz) break;
        const FT tx = g.topo[0] == NH_FLAT ? FT(0) : Fx[lk][lj][li + 1] - Fx[lk][lj][li];
        const FT ty = g.topo[1] == NH_FLAT ? FT(0) : Fy[lk][lj + 1][li] - Fy[lk][lj][li];
        const FT tz = g.topo[2] == NH_FLAT ? FT(0) : Fz[lk + 1][lj][li] - Fz[lk][lj][li];
        FT total = tx + ty;
        total = total + tz;
        divs[pass][lk] = FT(1) / volume_at(g, a == 2 ? 0 : 1, k0 + lk) * total;
      }
    }
  }
  if (!mine) return;
  const int i = i0 + li, j = j0 + lj;
#pragma unroll
  for (int lk = 0; lk < TG_TZ; ++lk) {
    if (lk >= nz) break;
    const int k = k0 + lk;
    if (a == 0 && g.topo[0] == NH_BOUNDED && g.N[0] > 1 && i == 1) continue;      // exclude_periphery
    if (a == 1 && g.topo[1] == NH_BOUNDED && g.N[1] > 1 && j == 1) continue;
    if (a == 2 && g.topo[2] == NH_BOUNDED && g.N[2] > 1 && k == 1) continue;
    const FT div = divs[0][lk], cdiv = P.has_closure ? divs[1][lk] : FT(0);
    const long long n = g.idx(i, j, k);
    if (a == 0) P.GU[0][n] = velocity_tendency<FT, 0>(P, i, j, k, div, cdiv);
    else if (a == 1) P.GU[1][n] = velocity_tendency<FT, 1>(P, i, j, k, div, cdiv);
    else if (a == 2) P.GU[2][n] = velocity_tendency<FT, 2>(P, i, j, k, div, cdiv);
    else P.Gc[t][n] = tracer_tendency<FT>(P, t, i, j, k, div, cdiv);
  }
}

// One launch for every prognostic field (what = -1: u, v, w, then the tracers, the tile's shared memory reused field after field)
// or for one of them (what = 0, 1, 2 or 3 + tracer index).  fast_ok: the model qualifies for the fast flux stage (set by the
// host: WENO5 in x, y, z; x, y Periodic; z Periodic or Bounded; regular spacing; parent volume below 2³¹ elements).
template <typename FT, int VARIANT = 0>
#if TG_MINBLOCKS > 1
#define TG_BOUNDS __launch_bounds__(TG_TX * TG_TY, TG_MINBLOCKS)
#else
#define TG_BOUNDS __launch_bounds__(TG_TX * TG_TY)
#endif
__global__ void TG_BOUNDS tendency_tiled_kernel(const __grid_constant__ TendParams<FT> P, int what, int fast_ok) {
  extern __shared__ __align__(16) unsigned char tg_smem[];
  FT* sm = reinterpret_cast<FT*>(tg_smem);
  const GridDev<FT>& g = P.g;
  // full-order region of a Bounded z for every flux location the tile touches (faces and centres k0-1 … k0+TZ)
  const int k0 = blockIdx.z * TG_TZ + 1, nz = min(TG_TZ, g.N[2] - k0 + 1);
  const bool fast = fast_ok && (g.topo[2] == NH_PERIODIC || (k0 - 1 >= 4 && k0 + nz <= g.N[2] - 2));
  const int first = what < 0 ? 0 : what, last = what < 0 ? 2 + P.ntr : what;
  for (int f = first; f <= last; ++f) {
    if (f == 0) tiled_tendency<FT, 0, VARIANT>(P, 0, sm, fast);
    else if (f == 1) tiled_tendency<FT, 1, VARIANT>(P, 0, sm, fast);
    else if (f == 2) tiled_tendency<FT, 2, VARIANT>(P, 0, sm, fast);
    else tiled_tendency<FT, 3, VARIANT>(P, f - 3, sm, fast);
    __syncthreads();
  }
}

// the contracted build of the tiled kernel lives in tendency_fma.cu (compiled with FMA contraction)
template <typename FT> void launch_tiled_tendency_fma(const TendParams<FT>& P, dim3 grid, int fast_ok, cudaStream_t stream);
