// closure_amd.cuh — AnisotropicMinimumDissipation eddy viscosity νₑ and per-tracer eddy diffusivities κₑ.
//   src/TurbulenceClosures/turbulence_closure_implementations/anisotropic_minimum_dissipation.jl:143-208 (kernels),
//   :216-227 (filter widths Δᶠ = 2Δ at the index asked for), :233-313, 333-365 (the 30 terms),
//   src/TurbulenceClosures/velocity_tracer_gradients.jl:8-23,129-262 (normalised gradients and strains),
//   src/Operators/interpolation_operators.jl:45-53 (ℑ_ab = ℑ_b(ℑ_a ·)).
// Sums and products keep the reference's association; the ℑxzᶜᵃᶜ(norm_∂y_w) of cy_uy (:347) is reproduced as written.
// Cb (buoyancy modification) is `nothing`, the reference's default.  One thread per cell of :xyz.
#pragma once
#include "tendency_generic.cuh"

template <typename FT>
struct AmdOps {
  const TendParams<FT>& P;
  const GridDev<FT>& g;
  __device__ __forceinline__ AmdOps(const TendParams<FT>& P_) : P(P_), g(P_.g) {}
  __device__ __forceinline__ FT Dx() const { return FT(2) * g.dx; }
  __device__ __forceinline__ FT Dy() const { return FT(2) * g.dy; }
  __device__ __forceinline__ FT Dz(int k) const { return FT(2) * g.Dzc(k); }
  // ccc
  __device__ __forceinline__ FT dx_u(int i, int j, int k) const { return deriv<FT, 0>(g, P.U[0], 1, 1, i, j, k); }
  __device__ __forceinline__ FT dy_v(int i, int j, int k) const { return deriv<FT, 1>(g, P.U[1], 1, 1, i, j, k); }
  __device__ __forceinline__ FT dz_w(int i, int j, int k) const { return deriv<FT, 2>(g, P.U[2], 1, 1, i, j, k); }
  // ffc
  __device__ __forceinline__ FT dx_v(int i, int j, int k) const { return Dx() / Dy() * deriv<FT, 0>(g, P.U[1], 0, 1, i, j, k); }
  __device__ __forceinline__ FT dy_u(int i, int j, int k) const { return Dy() / Dx() * deriv<FT, 1>(g, P.U[0], 0, 1, i, j, k); }
  // fcf
  __device__ __forceinline__ FT dx_w(int i, int j, int k) const { return Dx() / Dz(k) * deriv<FT, 0>(g, P.U[2], 0, 0, i, j, k); }
  __device__ __forceinline__ FT dz_u(int i, int j, int k) const { return Dz(k) / Dx() * deriv<FT, 2>(g, P.U[0], 0, 0, i, j, k); }
  // cff
  __device__ __forceinline__ FT dy_w(int i, int j, int k) const { return Dy() / Dz(k) * deriv<FT, 1>(g, P.U[2], 0, 0, i, j, k); }
  __device__ __forceinline__ FT dz_v(int i, int j, int k) const { return Dz(k) / Dy() * deriv<FT, 2>(g, P.U[1], 0, 0, i, j, k); }
  __device__ __forceinline__ FT S12(int i, int j, int k) const { return FT(0.5) * (dy_u(i, j, k) + dx_v(i, j, k)); }
  __device__ __forceinline__ FT S13(int i, int j, int k) const { return FT(0.5) * (dz_u(i, j, k) + dx_w(i, j, k)); }
  __device__ __forceinline__ FT S23(int i, int j, int k) const { return FT(0.5) * (dz_v(i, j, k) + dy_w(i, j, k)); }
  // tracer gradients (normalised)
  __device__ __forceinline__ FT cx(const FT* c, int i, int j, int k) const { return Dx() * deriv<FT, 0>(g, c, 0, 1, i, j, k); }
  __device__ __forceinline__ FT cy(const FT* c, int i, int j, int k) const { return Dy() * deriv<FT, 1>(g, c, 0, 1, i, j, k); }
  __device__ __forceinline__ FT cz(const FT* c, int i, int j, int k) const { return Dz(k) * deriv<FT, 2>(g, c, 0, 0, i, j, k); }
  __device__ __forceinline__ FT delta2(int k) const {
    const FT a = Dx(), b = Dy(), c = Dz(k);
    return FT(3) / (FT(1) / (a * a) + FT(1) / (b * b) + FT(1) / (c * c));
  }
};

// ℑ onto Centers along d (identity in a Flat direction), and along lo then hi
template <typename FT, int d, typename F>
__device__ __forceinline__ FT amd_c1(const GridDev<FT>& g, const F& f, int i, int j, int k) {
  if (g.topo[d] == NH_FLAT) return f(i, j, k);
  constexpr int di = d == 0, dj = d == 1, dk = d == 2;
  return FT(0.5) * (f(i, j, k) + f(i + di, j + dj, k + dk));
}
template <typename FT, int lo, int hi, typename F>
__device__ __forceinline__ FT amd_c2(const GridDev<FT>& g, const F& f, int i, int j, int k) {
  auto inner = [&](int a, int b, int c) { return amd_c1<FT, lo>(g, f, a, b, c); };
  return amd_c1<FT, hi>(g, inner, i, j, k);
}

// The viscosity kernel evaluates the off-diagonal normalised gradients and strains ONCE at the four corners of each of the
// three edge planes around the cell (ffc, fcf, cff) and forms every ℑ_ab of the 30 terms from those twelve corner triples:
// the same products and the same ℑ_hi(ℑ_lo ·) sums as interpolating each term separately, a quarter of the derivative work.
template <typename FT, int lo, int hi>
struct AmdCorners {
  FT p[4], q[4], s[4];      // ∂_lo u_hi, ∂_hi u_lo (normalised) and Σ_lohi at corners (0,0), (1,0), (0,1), (1,1) in (lo, hi)
  bool flat_lo, flat_hi;
  template <typename F>
  __device__ __forceinline__ FT I(const F& f) const {     // ℑ_hi(ℑ_lo f) onto the Center; identity in Flat directions
    const FT x0 = flat_lo ? f(0) : FT(0.5) * (f(0) + f(1));
    if (flat_hi) return x0;
    const FT x1 = flat_lo ? f(2) : FT(0.5) * (f(2) + f(3));
    return FT(0.5) * (x0 + x1);
  }
};

template <typename FT>
__global__ void __launch_bounds__(128) amd_viscosity_kernel(const __grid_constant__ TendParams<FT> P, FT* nue, FT Cnu) {
  const GridDev<FT>& g = P.g;
  int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  const AmdOps<FT> o(P);
  AmdCorners<FT, 0, 1> xy; AmdCorners<FT, 0, 2> xz; AmdCorners<FT, 1, 2> yz;
  xy.flat_lo = xz.flat_lo = g.topo[0] == NH_FLAT;
  xy.flat_hi = yz.flat_lo = g.topo[1] == NH_FLAT;
  xz.flat_hi = yz.flat_hi = g.topo[2] == NH_FLAT;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int a = c & 1, b = c >> 1;
    // in a Flat direction only the corners with offset 0 are used (and exist)
    if (!((a && xy.flat_lo) || (b && xy.flat_hi))) {
      xy.p[c] = o.dx_v(i + a, j + b, k); xy.q[c] = o.dy_u(i + a, j + b, k); xy.s[c] = FT(0.5) * (xy.q[c] + xy.p[c]);
    }
    if (!((a && xz.flat_lo) || (b && xz.flat_hi))) {
      xz.p[c] = o.dx_w(i + a, j, k + b); xz.q[c] = o.dz_u(i + a, j, k + b); xz.s[c] = FT(0.5) * (xz.q[c] + xz.p[c]);
    }
    if (!((a && yz.flat_lo) || (b && yz.flat_hi))) {
      yz.p[c] = o.dy_w(i, j + a, k + b); yz.q[c] = o.dz_v(i, j + a, k + b); yz.s[c] = FT(0.5) * (yz.q[c] + yz.p[c]);
    }
  }
  const FT ux = o.dx_u(i, j, k), vy = o.dy_v(i, j, k), wz = o.dz_w(i, j, k);
  const FT xv2 = xy.I([&](int c) { return xy.p[c] * xy.p[c]; }), yu2 = xy.I([&](int c) { return xy.q[c] * xy.q[c]; });
  const FT xw2 = xz.I([&](int c) { return xz.p[c] * xz.p[c]; }), zu2 = xz.I([&](int c) { return xz.q[c] * xz.q[c]; });
  const FT yw2 = yz.I([&](int c) { return yz.p[c] * yz.p[c]; }), zv2 = yz.I([&](int c) { return yz.q[c] * yz.q[c]; });
  FT q = ux * ux + vy * vy + wz * wz;
  q = q + xv2; q = q + yu2; q = q + xw2; q = q + zu2; q = q + yw2; q = q + zv2;
  FT nu = FT(0);
  if (q != FT(0)) {
    const FT two = FT(2);
    FT g1 = ux * (ux * ux);
    g1 = g1 + vy * xv2;
    g1 = g1 + wz * xw2;
    g1 = g1 + two * ux * xy.I([&](int c) { return xy.p[c] * xy.s[c]; });
    g1 = g1 + two * ux * xz.I([&](int c) { return xz.p[c] * xz.s[c]; });
    g1 = g1 + two * xy.I([&](int c) { return xy.p[c]; }) * xz.I([&](int c) { return xz.p[c]; }) * yz.I([&](int c) { return yz.s[c]; });
    FT g2 = ux * yu2;
    g2 = g2 + vy * (vy * vy);
    g2 = g2 + wz * yw2;
    g2 = g2 + two * vy * xy.I([&](int c) { return xy.q[c] * xy.s[c]; });
    g2 = g2 + two * xy.I([&](int c) { return xy.q[c]; }) * yz.I([&](int c) { return yz.p[c]; }) * xz.I([&](int c) { return xz.s[c]; });
    g2 = g2 + two * vy * yz.I([&](int c) { return yz.p[c] * yz.s[c]; });
    FT g3 = ux * zu2;
    g3 = g3 + vy * zv2;
    g3 = g3 + wz * (wz * wz);
    g3 = g3 + two * xz.I([&](int c) { return xz.q[c]; }) * yz.I([&](int c) { return yz.q[c]; }) * xy.I([&](int c) { return xy.s[c]; });
    g3 = g3 + two * wz * xz.I([&](int c) { return xz.q[c] * xz.s[c]; });
    g3 = g3 + two * wz * yz.I([&](int c) { return yz.q[c] * yz.s[c]; });
    const FT r = (g1 + g2) + g3;
    nu = -Cnu * o.delta2(k) * (r - FT(0)) / q;
    nu = (nu > FT(0) || nu != nu) ? nu : FT(0);            // max(zero(FT), νˢᵍˢ), NaN-propagating like Julia's max
  }
  nue[g.idx(i, j, k)] = nu;
}

#define AMD_F(expr) [&](int a, int b, int c) { return (expr); }
#define XY(expr) amd_c2<FT, 0, 1>(g, AMD_F(expr), i, j, k)
#define XZ(expr) amd_c2<FT, 0, 2>(g, AMD_F(expr), i, j, k)
#define YZ(expr) amd_c2<FT, 1, 2>(g, AMD_F(expr), i, j, k)
template <typename FT>
__global__ void __launch_bounds__(128) amd_diffusivity_kernel(const __grid_constant__ TendParams<FT> P, const FT* tr, FT* kape,
                                                              FT Ckappa) {
  const GridDev<FT>& g = P.g;
  int i = blockIdx.x * blockDim.x + threadIdx.x + 1;
  int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  const AmdOps<FT> o(P);
  auto fx = [&](int a, int b, int d) { return o.cx(tr, a, b, d); };
  auto fy = [&](int a, int b, int d) { return o.cy(tr, a, b, d); };
  auto fz = [&](int a, int b, int d) { return o.cz(tr, a, b, d); };
  auto fx2 = [&](int a, int b, int d) { const FT x_ = o.cx(tr, a, b, d); return x_ * x_; };
  auto fy2 = [&](int a, int b, int d) { const FT x_ = o.cy(tr, a, b, d); return x_ * x_; };
  auto fz2 = [&](int a, int b, int d) { const FT x_ = o.cz(tr, a, b, d); return x_ * x_; };
  const FT cx2 = amd_c1<FT, 0>(g, fx2, i, j, k), cy2 = amd_c1<FT, 1>(g, fy2, i, j, k), cz2 = amd_c1<FT, 2>(g, fz2, i, j, k);
  const FT sigma = (cx2 + cy2) + cz2;
  FT kappa = FT(0);
  if (sigma != FT(0)) {
    const FT Icx = amd_c1<FT, 0>(g, fx, i, j, k), Icy = amd_c1<FT, 1>(g, fy, i, j, k), Icz = amd_c1<FT, 2>(g, fz, i, j, k);
    const FT ux = o.dx_u(i, j, k), vy = o.dy_v(i, j, k), wz = o.dz_w(i, j, k);
    FT cx_ux = ux * cx2;
    cx_ux = cx_ux + XY(o.dx_v(a, b, c)) * Icx * Icy;
    cx_ux = cx_ux + XZ(o.dx_w(a, b, c)) * Icx * Icz;
    FT cy_uy = XY(o.dy_u(a, b, c)) * Icy * Icx;
    cy_uy = cy_uy + vy * cy2;
    cy_uy = cy_uy + XZ(o.dy_w(a, b, c)) * Icy * Icz;      // ℑxzᶜᵃᶜ of a cff quantity, as the reference writes it
    FT cz_uz = XZ(o.dz_u(a, b, c)) * Icz * Icx;
    cz_uz = cz_uz + YZ(o.dz_v(a, b, c)) * Icz * Icy;
    cz_uz = cz_uz + wz * cz2;
    const FT theta = (cx_ux + cy_uy) + cz_uz;
    kappa = -Ckappa * o.delta2(k) * theta / sigma;
    kappa = (kappa > FT(0) || kappa != kappa) ? kappa : FT(0);
  }
  kape[g.idx(i, j, k)] = kappa;
}
#undef AMD_F
#undef XY
#undef XZ
#undef YZ
