// solver_generic.cu — GridHost::build and the complete Poisson solvers (host orchestration of fft_generic.cuh)
#include "model.cuh"

#define LAUNCH_COUNT(lc) do { if (lc) (lc)->n++; } while (0)

static int round_up(int a, int b) { return (a + b - 1) / b * b; }

template <typename FT>
void GridHost<FT>::build(const nh_grid_desc& d, const std::vector<double>& zfaces, const int Hin[3]) {
  for (int e = 0; e < 3; ++e) {
    g.N[e] = d.N[e];
    g.topo[e] = d.topology[e];
    g.H[e] = d.topology[e] == NH_FLAT ? 0 : Hin[e];
    NH_REQUIRE(g.N[e] >= 1, "grid size must be >= 1");
    NH_REQUIRE(g.topo[e] >= 0 && g.topo[e] <= 2, "bad topology");
  }
  // x: first interior element on a 128-byte line, pitch a multiple of 16 elements
  g.ox = g.H[0] == 0 ? 0 : round_up(g.H[0], 16);
  g.px = round_up(g.ox + g.N[0] + 1 + g.H[0], 16);
  g.oy = g.H[1]; g.py = g.N[1] + 2 * g.H[1] + 1;
  g.oz = g.H[2]; g.pz = g.N[2] + 2 * g.H[2] + 1;
  g.sy = g.px; g.sz = (long long)g.px * g.py;
  // metrics
  const double* iv[3] = {d.x, d.y, d.z};
  FT delta[3];
  for (int e = 0; e < 3; ++e) {
    if (g.topo[e] == NH_FLAT) { delta[e] = FT(1); L[e] = FT(1); continue; }
    if (e == 2 && !zfaces.empty()) continue;
    nhhost::regular_spacing<FT>(iv[e][0], iv[e][1], g.N[e], delta[e], L[e]);
  }
  g.dx = delta[0]; g.dy = delta[1];
  const int Nz = g.N[2], Hz = g.H[2];
  g.zoff = Hz + 2;
  size_t len = (size_t)Nz + 2 * Hz + 6;
  dzc_h.assign(len, FT(0)); dzf_h.assign(len, FT(0));
  z_stretched = (g.topo[2] != NH_FLAT && !zfaces.empty());
  if (z_stretched) {
    NH_REQUIRE((int)zfaces.size() == Nz + 1, "z_faces must have Nz+1 entries");
    std::vector<FT> dc, df;
    nhhost::stretched_spacings<FT>(zfaces.data(), Nz, Hz, g.topo[2] == NH_BOUNDED, dc, df, L[2]);
    for (size_t q = 0; q < dc.size(); ++q) dzc_h[(1 - Hz) + (int)q + g.zoff] = dc[q];
    for (size_t q = 0; q < df.size(); ++q) dzf_h[(-Hz) + (int)q + g.zoff] = df[q];
    g.z_regular = 0; g.dz = FT(0);
  } else {
    for (auto& v : dzc_h) v = delta[2];
    for (auto& v : dzf_h) v = delta[2];
    g.z_regular = 1; g.dz = delta[2];
  }
  dzc_d.upload(dzc_h); dzf_d.upload(dzf_h);
  g.dzc = dzc_d.p; g.dzf = dzf_d.p;
}
template struct GridHost<float>;
template struct GridHost<double>;

// ---- PoissonGeneric ---------------------------------------------------------------------------------------------
template <typename FT>
void PoissonGeneric<FT>::create(const GridHost<FT>& gh, int kind_) {
  kind = kind_;
  if (!lc) lc = &own_lc;
  for (int e = 0; e < 3; ++e) { N[e] = gh.g.N[e]; topo[e] = gh.g.topo[e]; }
  if (kind == NH_SOLVER_FFT)
    NH_REQUIRE(!gh.z_stretched, "FFTBasedPoissonSolver requires a grid that is regular in x, y and z");
  if (kind == NH_SOLVER_FOURIER_TRIDIAGONAL)
    NH_REQUIRE(topo[2] == NH_BOUNDED, "`FourierTridiagonalPoissonSolver` can only be used when the stretched direction's topology is `Bounded`.");
  A.alloc(count());
  // per device, so once per solver (the current device is the solver's)
  CUDA_CHECK(cudaFuncSetAttribute(line_transform_kernel<FT, CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  int ndims = kind == NH_SOLVER_FFT ? 3 : 2;
  for (int e = 0; e < 3; ++e) {
    // h = L/N evaluated in FT like the reference's `L / N` (poisson_eigenvalues.jl:10)
    double h = (double)(FT)(gh.L[e] / (FT)N[e]);
    std::vector<double> l = (e < ndims) ? nhhost::poisson_eigenvalues(N[e], h, topo[e]) : std::vector<double>(N[e], 0.0);
    lam[e].upload(l);
    if (e < ndims && topo[e] != NH_FLAT && N[e] > 1) {
      std::vector<double> a, b;
      nhhost::twiddles(N[e], a, b);
      std::vector<CT> ta(N[e]), tb(N[e]);
      for (int q = 0; q < N[e]; ++q) { ta[q].x = (FT)a[2 * q]; ta[q].y = (FT)a[2 * q + 1]; tb[q].x = (FT)b[2 * q]; tb[q].y = (FT)b[2 * q + 1]; }
      tw[e].upload(ta); tw4[e].upload(tb);
    }
  }
  if (kind == NH_SOLVER_FOURIER_TRIDIAGONAL) {
    Phi.alloc(count());
    D.alloc(count(), false);
    tt.alloc(count());
    std::vector<FT> low;
    for (int q = 2; q <= N[2]; ++q) low.push_back(FT(1) / gh.dzf_h[q + gh.g.zoff]);
    if (low.empty()) low.push_back(FT(0));
    lower.upload(low);
    std::vector<FT> dz1(N[2]);
    for (int k = 1; k <= N[2]; ++k) dz1[k - 1] = gh.dzc_h[k + gh.g.zoff];
    dzc1.upload(dz1);
    dim3 blk(32, 4), grd(ceil_div(N[0], 32), ceil_div(N[1], 4));
    main_diagonal_kernel<FT><<<grd, blk, 0, stream>>>(gh.g, lam[0].p, lam[1].p, D.p);
    LAUNCH_COUNT(lc);
    CUDA_CHECK(cudaGetLastError());
    partial.alloc(1024);
  }
}

template <typename FT>
void PoissonGeneric<FT>::transform(CT* a, int d, int mode) {
  if (topo[d] == NH_FLAT || N[d] == 1) {
    // a length-1 FFT is the identity; a length-1 DCT-II is ×2 and its normalised inverse ×(1/2): both cancel,
    // but the reference applies them, so do the same through the kernel for N == 1 Bounded dims.
    if (!(topo[d] == NH_BOUNDED && N[d] == 1)) return;
  }
  LineJob J;
  J.N = N[d]; J.mode = mode;
  J.pow2 = (N[d] & (N[d] - 1)) == 0 && N[d] > 1;
  long long Nx = N[0], Ny = N[1], Nz = N[2];
  if (d == 0) { J.S = 1; J.nlines = Ny * Nz; J.P = (int)std::min<long long>(J.nlines, 1 << 30); J.bp = Nx; J.bq = 0; NH_REQUIRE(J.nlines < (1LL << 30), "too many lines"); }
  else if (d == 1) { J.S = Nx; J.nlines = Nx * Nz; J.P = (int)Nx; J.bp = 1; J.bq = Nx * Ny; }
  else { J.S = Nx * Ny; J.nlines = Nx * Ny; J.P = (int)(Nx * Ny); J.bp = 1; J.bq = 0; }
  size_t per_line = 2 * (size_t)N[d] * sizeof(CT);
  const size_t budget = 96 * 1024;
  int lpb = (int)std::max<size_t>(1, std::min<size_t>(8, budget / per_line));
  NH_REQUIRE(per_line <= 200 * 1024, "line too long for the shared-memory transform");
  J.lpb = lpb;
  size_t smem = per_line * lpb;
  int blocks = (int)((J.nlines + lpb - 1) / lpb);
  if (N[d] == 1) {
    // dummy tables are never read for N == 1 except tw4[0] = 1 in the DCT modes
    if (!tw4[d].p) { std::vector<CT> one(1); one[0].x = 1; one[0].y = 0; tw4[d].upload(one); tw[d].upload(one); }
  }
  line_transform_kernel<FT, CT><<<blocks, 256, smem, stream>>>(a, J, tw[d].p, tw4[d].p);
  LAUNCH_COUNT(lc);
  CUDA_CHECK(cudaGetLastError());
}

template <typename FT>
void PoissonGeneric<FT>::forward(CT* a) {
  int nd = kind == NH_SOLVER_FFT ? 3 : 2;
  for (int d = 0; d < nd; ++d) if (topo[d] == NH_BOUNDED) transform(a, d, LT_FWD_DCT);
  for (int d = 0; d < nd; ++d) if (topo[d] == NH_PERIODIC) transform(a, d, LT_FWD_FFT);
}
template <typename FT>
void PoissonGeneric<FT>::backward(CT* a) {
  int nd = kind == NH_SOLVER_FFT ? 3 : 2;
  for (int d = 0; d < nd; ++d) if (topo[d] == NH_PERIODIC) transform(a, d, LT_INV_FFT);
  for (int d = 0; d < nd; ++d) if (topo[d] == NH_BOUNDED) transform(a, d, LT_INV_DCT);
}

template <typename FT>
typename PoissonGeneric<FT>::CT* PoissonGeneric<FT>::solve(double m) {
  long long n = count();
  int blocks = (int)((n + 255) / 256);
  forward(A.p);
  if (kind == NH_SOLVER_FFT) {
    spectral_divide_kernel<FT, CT><<<blocks, 256, 0, stream>>>(A.p, N[0], N[1], N[2], lam[0].p, lam[1].p, lam[2].p, m, m == 0.0);
    LAUNCH_COUNT(lc);
    backward(A.p);
    CUDA_CHECK(cudaGetLastError());
    return A.p;
  }
  long long cols = (long long)N[0] * N[1];
  thomas_z_kernel<FT, CT><<<(int)((cols + 127) / 128), 128, 0, stream>>>(N[0], N[1], N[2], lower.p, D.p, A.p, Phi.p, tt.p);
  LAUNCH_COUNT(lc);
  backward(Phi.p);
  int rb = (int)std::min<long long>(1024, (n + 255) / 256);
  csum_kernel<CT><<<rb, 256, 0, stream>>>(Phi.p, n, partial.p);
  csum_final_kernel<<<1, 256, 0, stream>>>(partial.p, rb);
  subtract_mean_kernel<FT, CT><<<blocks, 256, 0, stream>>>(Phi.p, n, partial.p);
  if (lc) lc->n += 3;
  CUDA_CHECK(cudaGetLastError());
  return Phi.p;
}

template struct PoissonGeneric<float>;
template struct PoissonGeneric<double>;
