// dist_host.cuh — host orchestration of the y-slab multi-GPU pieces (kernels in dist.cuh); included by model.cuh
#pragma once
#include "dist.cuh"

#define PD_SWITCH_N(N_, CALL)                                          \
  switch (N_) {                                                         \
    case 16: { constexpr int NN = 16; CALL; } break;                    \
    case 32: { constexpr int NN = 32; CALL; } break;                    \
    case 64: { constexpr int NN = 64; CALL; } break;                    \
    case 128: { constexpr int NN = 128; CALL; } break;                  \
    case 256: { constexpr int NN = 256; CALL; } break;                  \
    case 512: { constexpr int NN = 512; CALL; } break;                  \
    case 1024: { constexpr int NN = 1024; CALL; } break;                \
    default: throw NhError("unsupported FFT size on the distributed fused path");  \
  }

template <typename FT>
struct DistState {
  typedef typename cx_of<FT>::type CT;
  bool on = false;
  int rank = 0, R = 1, Ny = 0;          // Ny: global extent in y
  PFJob<FT, CT> J;                       // local sizes; S = buffer A
  DevBuf<CT> A, B, twx, twy, twz;
  DevBuf<double> lx, ly, lz;
  DevBuf<FT> send_lo, send_hi, recv_lo, recv_hi;
  size_t halo_elems = 0, spec_elems = 0;
  double t_next = 0;
  // peer-to-peer mode (nh_dist_ipc_*): NVLink peer pointers of every rank's spectral buffers and of the neighbours' fields
  bool p2p = false;
  CT* peerA[PF_MAXR];
  CT* peerB[PF_MAXR];
  FT *prevF[3], *prevF2[3], *nextF[3], *nextF2[3], *next_p = nullptr;
  // flag barrier (peer_barrier_kernel): a 2 MB allocation of its own so that its IPC handle covers nothing else
  DevBuf<unsigned int> flags;
  unsigned int* peer_flags[PF_MAXR];
  unsigned int epoch = 0;
  std::vector<void*> opened;
  ~DistState() { for (void* q : opened) cudaIpcCloseMemHandle(q); }
  // complete-path models on a slab (any physics, z Periodic or Bounded): the model's own kernels run on the local grid, the
  // y halo fill pulls the neighbours' rows over NVLink and calculate_pressure_correction! calls solve_generic below
  bool generic = false;
  bool tri = false;                      // z Bounded: Fourier-tridiagonal (x, y transformed; Thomas sweep along z on buffer B)
  GridDev<FT> gdense;                    // dense local N³ layout of the real source term
  DevBuf<FT> tt, lower;
  DevBuf<double> dA, dC;
  DevBuf<PFStale<FT>> stale;
  FT* peer_prev[32];                     // the previous / next rank's field of each halo-filled field id (nullptr: none)
  FT* peer_next[32];

  void create(const GridHost<FT>& gh, int rank_, int R_, double y0, double y1, int Ny_global, FT& dy_out, bool generic_ = false) {
    const GridDev<FT>& g = gh.g;
    rank = rank_; R = R_; Ny = Ny_global;
    generic = generic_;
    tri = generic && g.topo[2] == NH_BOUNDED;
    for (int q = 0; q < 32; ++q) peer_prev[q] = peer_next[q] = nullptr;
    NH_REQUIRE(R >= 1 && rank >= 0 && rank < R, "bad rank / nranks");
    NH_REQUIRE(g.N[1] * R == Ny, "global Ny must equal nranks x local Ny");
    NH_REQUIRE(g.N[1] % 2 == 0 && g.N[1] >= 2 * g.H[1], "local Ny must be even and >= 2 halos");
    for (int N : {g.N[0], Ny}) NH_REQUIRE(N >= 16 && N <= 1024 && !(N & (N - 1)), "distributed path needs power-of-two sizes in 16..1024 in x and y");
    if (!tri) NH_REQUIRE(g.N[2] >= 16 && g.N[2] <= 1024 && !(g.N[2] & (g.N[2] - 1)), "distributed path needs a power-of-two size in 16..1024 in a Periodic z");
    else NH_REQUIRE(g.N[2] >= 2, "distributed Fourier-tridiagonal solve needs Nz >= 2");
    FT Ly;
    nhhost::regular_spacing<FT>(y0, y1, Ny, dy_out, Ly);
    J.g = g; J.g.dy = dy_out;
    J.Nx = g.N[0]; J.Ny = g.N[1]; J.Nz = g.N[2]; J.Mx = J.Nx / 2 + 2; J.ywrap = 0;
    int cols = J.Nx / 2 + 1;
    J.Cx = ((cols + R - 1) / R + 7) / 8 * 8;                        // columns per rank, a multiple of the 8-column tile
    spec_elems = (size_t)R * J.Cx * J.Ny * J.Nz;
    A.alloc(spec_elems); B.alloc(spec_elems);
    J.S = A.p;
    J.px = ff_make_plan(J.Nx); J.py = ff_make_plan(Ny); J.pz = ff_make_plan(tri ? 16 : J.Nz);
    gdense = g;
    for (int e = 0; e < 3; ++e) gdense.H[e] = 0;
    gdense.ox = gdense.oy = gdense.oz = 0; gdense.px = g.N[0]; gdense.py = g.N[1]; gdense.pz = g.N[2];
    gdense.sy = g.N[0]; gdense.sz = (long long)g.N[0] * g.N[1];
    gdense.dy = dy_out;
    if (tri) {
      std::vector<double> a, c;
      std::vector<FT> low, dz1;
      PoissonFast<FT>::tri_tables(gh, a, c, low, dz1);
      dA.upload(a); dC.upload(c); lower.upload(low);
      tt.alloc(spec_elems);
      stale.alloc(1);
    }
    int Ns[3] = {J.Nx, Ny, J.Nz};
    FT Ls[3] = {gh.L[0], Ly, gh.L[2]};
    DevBuf<CT>* tws[3] = {&twx, &twy, &twz};
    DevBuf<double>* ls[3] = {&lx, &ly, &lz};
    for (int e = 0; e < (tri ? 2 : 3); ++e) {
      std::vector<double> a, b;
      nhhost::twiddles(Ns[e], a, b);
      std::vector<CT> t(Ns[e]);
      for (int q = 0; q < Ns[e]; ++q) { t[q].x = (FT)a[2 * q]; t[q].y = (FT)a[2 * q + 1]; }
      tws[e]->upload(t);
      double h = (double)(FT)(Ls[e] / (FT)Ns[e]);
      ls[e]->upload(nhhost::poisson_eigenvalues(Ns[e], h, NH_PERIODIC));
    }
    J.twx = twx.p; J.twy = twy.p; J.twz = twz.p; J.lx = lx.p; J.ly = ly.p; J.lz = lz.p;
    for (int d = 0; d < PF_MAXR; ++d) { J.peer[d] = nullptr; peerA[d] = peerB[d] = nullptr; }
    J.p_next = nullptr; J.tw4z = nullptr; J.dzscale = nullptr;
    halo_elems = (size_t)(g.N[0] + 2 * g.H[0]) * g.H[1] * (g.N[2] + 2 * g.H[2]) * 3;
    send_lo.alloc(halo_elems); send_hi.alloc(halo_elems); recv_lo.alloc(halo_elems); recv_hi.alloc(halo_elems);
    PD_SWITCH_N(J.Nx, (set_attr_x<NN>()));
    if (!tri) PD_SWITCH_N(J.Nz, (set_attr_z<NN>()));
    PD_SWITCH_N(Ny, (set_attr_y<NN>()));
    on = true;
  }
  template <int N> void set_attr_x() {
    const int smemx = PF_L * ((N + N / 8 + N / 64 + 1) | 1) * (int)sizeof(CT);
    CUDA_CHECK(cudaFuncSetAttribute(pf_x_forward_kernel<FT, CT, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemx));
    CUDA_CHECK(cudaFuncSetAttribute(pf_x_forward_kernel<FT, CT, N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemx));
    CUDA_CHECK(cudaFuncSetAttribute(pf_x_inverse_kernel<FT, CT, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemx));
  }
  template <int N> void set_attr_z() {
    const int smem = PF_L * FFPitch8<N>::value * (int)sizeof(CT);
    CUDA_CHECK(cudaFuncSetAttribute(pd_z_kernel<FT, CT, N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    CUDA_CHECK(cudaFuncSetAttribute(pd_z_kernel<FT, CT, N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    CUDA_CHECK(cudaFuncSetAttribute(pd_z_scatter_kernel<FT, CT, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  }
  template <int N> void set_attr_y() {
    const int smem = PF_L * FFPitch8<N>::value * (int)sizeof(CT);
    CUDA_CHECK(cudaFuncSetAttribute(pd_y_solve_kernel<FT, CT, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    CUDA_CHECK(cudaFuncSetAttribute(pd_y_solve_kernel<FT, CT, N, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    CUDA_CHECK(cudaFuncSetAttribute(pd_y_solve_kernel<FT, CT, N, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  }
  template <int N> void launch_x_rhs(const FT* rhs, cudaStream_t s) {
    const int smemx = PF_L * ((N + N / 8 + N / 64 + 1) | 1) * (int)sizeof(CT);
    dim3 gx(ceil_div(J.Ny, 2 * PF_L), J.Nz);
    pf_x_forward_kernel<FT, CT, N, true><<<gx, N, smemx, s>>>(J, rhs, nullptr, nullptr);
  }
  template <int N> void launch_y_mode(int mode, cudaStream_t s) {
    const int smem = PF_L * FFPitch8<N>::value * (int)sizeof(CT);
    dim3 gy(J.Cx / PF_L, J.Nz);
    if (mode == 1) pd_y_solve_kernel<FT, CT, N, 1><<<gy, N, smem, s>>>(J, B.p, rank);
    else pd_y_solve_kernel<FT, CT, N, 2><<<gy, N, smem, s>>>(J, B.p, rank);
  }
  // ---- complete-path solve (peer-to-peer only): rhs = dense real source term (already × Δzᶜ when tri) ----
  // phase 0: x-FFT of rhs → (z-FFT) → scatter into the peers' y-local buffers
  void generic_forward(const FT* rhs, cudaStream_t s, LaunchCounter& lc) {
    const GridDev<FT> keep = J.g;
    J.g = gdense; J.dt = 1.0; J.m = 0.0; J.S = A.p; J.p_next = nullptr; J.dzscale = nullptr;
    for (int d = 0; d < PF_MAXR; ++d) J.peer[d] = peerB[d];
    PD_SWITCH_N(J.Nx, (launch_x_rhs<NN>(rhs, s)));
    J.g = keep;
    if (tri) {
      ScatterJob<CT> Sj;
      Sj.src = A.p; Sj.chunk = (long long)J.Cx * J.Ny * J.Nz; Sj.R = R; Sj.rank = rank;
      for (int d = 0; d < PF_MAXR; ++d) Sj.peer[d] = peerB[d];
      pd_scatter_kernel<CT><<<(unsigned)ceil_div(Sj.chunk * R, 256), 256, 0, s>>>(Sj);
    } else {
      PD_SWITCH_N(J.Nz, (launch_z<NN>(false, s)));
    }
    lc.n += 2;
    CUDA_CHECK(cudaGetLastError());
  }
  // phase 1: y-FFT, division (or Thomas sweep along z), y-IFFT, scatter back into the peers' z-local buffers
  void generic_solve_y(cudaStream_t s, LaunchCounter& lc) {
    for (int d = 0; d < PF_MAXR; ++d) J.peer[d] = peerA[d];
    if (!tri) { PD_SWITCH_N(Ny, (launch_y<NN>(s))); lc.n += 1; CUDA_CHECK(cudaGetLastError()); return; }
    PD_SWITCH_N(Ny, (launch_y_mode<NN>(1, s)));
    J.S = B.p;
    dim3 gt(ceil_div(J.Cx, 128), Ny);
    pf_thomas_kernel<FT, CT, true><<<gt, 128, 0, s>>>(J, lower.p, dA.p, dC.p, B.p, tt.p, stale.p, rank);
    J.S = A.p;
    PD_SWITCH_N(Ny, (launch_y_mode<NN>(2, s)));
    lc.n += 3;
    CUDA_CHECK(cudaGetLastError());
  }
  // phase 2: (z-IFFT,) x-IFFT, real store into the interior of p
  void generic_backward(FT* p, cudaStream_t s, LaunchCounter& lc) {
    J.S = A.p; J.p_next = nullptr;
    for (int d = 0; d < PF_MAXR; ++d) J.peer[d] = nullptr;
    if (!tri) { PD_SWITCH_N(J.Nz, (launch_z<NN>(true, s))); lc.n += 1; }
    PD_SWITCH_N(J.Nx, (launch_x<NN>(false, nullptr, nullptr, nullptr, p, s)));
    lc.n += 1;
    if (tri) {
      // the entries the Thomas guard skipped keep what the previous solve left in real space (pf_stale_gather_kernel); with
      // slabs only the singular (kx, ky) = (0, 0) column matters and both its spectral and real-space homes are rank 0
      if (rank == 0) { pf_stale_gather_kernel<FT><<<1, PF_STALE_MAX, 0, s>>>(stale.p, p, J.g); lc.n += 1; }
    }
    CUDA_CHECK(cudaGetLastError());
  }
  template <int N> void launch_x(bool fwd, const FT* u, const FT* v, const FT* w, FT* p, cudaStream_t s) {
    const int smemx = PF_L * ((N + N / 8 + N / 64 + 1) | 1) * (int)sizeof(CT);
    dim3 gx(ceil_div(J.Ny, 2 * PF_L), J.Nz);
    if (fwd) pf_x_forward_kernel<FT, CT, N><<<gx, N, smemx, s>>>(J, u, v, w);
    else pf_x_inverse_kernel<FT, CT, N><<<gx, N, smemx, s>>>(J, p);
  }
  template <int N> void launch_z(bool inv, cudaStream_t s) {
    const int smem = PF_L * FFPitch8<N>::value * (int)sizeof(CT);
    dim3 gz(R * J.Cx / PF_L, J.Ny);
    if (inv) pd_z_kernel<FT, CT, N, true><<<gz, N, smem, s>>>(J, R);
    else if (p2p) pd_z_scatter_kernel<FT, CT, N><<<gz, N, smem, s>>>(J, rank);
    else pd_z_kernel<FT, CT, N, false><<<gz, N, smem, s>>>(J, R);
  }
  template <int N> void launch_y(cudaStream_t s) {
    const int smem = PF_L * FFPitch8<N>::value * (int)sizeof(CT);
    dim3 gy(J.Cx / PF_L, J.Nz);
    pd_y_solve_kernel<FT, CT, N><<<gy, N, smem, s>>>(J, B.p, rank);
  }
  // div(U*)/Δt → x-FFT → z-FFT, result in A (the all-to-all send buffer)
  void forward(const FT* u, const FT* v, const FT* w, double dt, cudaStream_t s, LaunchCounter& lc) {
    J.dt = dt; J.m = 0.0; J.S = A.p; J.p_next = nullptr;
    for (int d = 0; d < PF_MAXR; ++d) J.peer[d] = p2p ? peerB[d] : nullptr;      // the scatter targets the y-local buffers
    PD_SWITCH_N(J.Nx, (launch_x<NN>(true, u, v, w, nullptr, s)));
    PD_SWITCH_N(J.Nz, (launch_z<NN>(false, s)));
    lc.n += 2;
    CUDA_CHECK(cudaGetLastError());
  }
  // after all_to_all(A → B): y-FFT, division, y-IFFT in B
  void solve_y(cudaStream_t s, LaunchCounter& lc) {
    for (int d = 0; d < PF_MAXR; ++d) J.peer[d] = p2p ? peerA[d] : nullptr;      // … and back into the z-local ones
    PD_SWITCH_N(Ny, (launch_y<NN>(s)));
    lc.n += 1;
    CUDA_CHECK(cudaGetLastError());
  }
  // after all_to_all(B → A): z-IFFT, x-IFFT, real store into p
  void backward(FT* p, cudaStream_t s, LaunchCounter& lc) {
    J.S = A.p; J.p_next = p2p ? next_p : nullptr;
    for (int d = 0; d < PF_MAXR; ++d) J.peer[d] = nullptr;
    PD_SWITCH_N(J.Nz, (launch_z<NN>(true, s)));
    PD_SWITCH_N(J.Nx, (launch_x<NN>(false, nullptr, nullptr, nullptr, p, s)));
    lc.n += 2;
    CUDA_CHECK(cudaGetLastError());
  }
};
