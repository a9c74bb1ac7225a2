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

  void create(const GridHost<FT>& gh, int rank_, int R_, double y0, double y1, int Ny_global, FT& dy_out) {
    const GridDev<FT>& g = gh.g;
    rank = rank_; R = R_; Ny = Ny_global;
    NH_REQUIRE(R >= 1 && rank >= 0 && rank < R, "bad rank / nranks");
    NH_REQUIRE(g.N[1] * R == Ny, "global Ny must equal nranks x local Ny");
    NH_REQUIRE(g.N[1] % 2 == 0 && g.N[1] >= 2 * g.H[1], "local Ny must be even and >= 2 halos");
    for (int N : {g.N[0], Ny, g.N[2]}) NH_REQUIRE(N >= 16 && N <= 1024 && !(N & (N - 1)), "distributed fused path needs power-of-two sizes in 16..1024");
    FT Ly;
    nhhost::regular_spacing<FT>(y0, y1, Ny, dy_out, Ly);
    J.g = g; J.g.dy = dy_out;
    J.Nx = g.N[0]; J.Ny = g.N[1]; J.Nz = g.N[2]; J.Mx = J.Nx / 2 + 2; J.ywrap = 0;
    int cols = J.Nx / 2 + 1;
    J.Cx = ((cols + R - 1) / R + 7) / 8 * 8;                        // columns per rank, a multiple of the 8-column tile
    spec_elems = (size_t)R * J.Cx * J.Ny * J.Nz;
    A.alloc(spec_elems); B.alloc(spec_elems);
    J.S = A.p;
    J.px = ff_make_plan(J.Nx); J.py = ff_make_plan(Ny); J.pz = ff_make_plan(J.Nz);
    int Ns[3] = {J.Nx, Ny, J.Nz};
    FT Ls[3] = {gh.L[0], Ly, gh.L[2]};
    DevBuf<CT>* tws[3] = {&twx, &twy, &twz};
    DevBuf<double>* ls[3] = {&lx, &ly, &lz};
    for (int e = 0; e < 3; ++e) {
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
    PD_SWITCH_N(J.Nz, (set_attr_z<NN>()));
    PD_SWITCH_N(Ny, (set_attr_y<NN>()));
    on = true;
  }
  template <int N> void set_attr_x() {
    const int smemx = PF_L * ((N + N / 8 + N / 64 + 1) | 1) * (int)sizeof(CT);
    CUDA_CHECK(cudaFuncSetAttribute(pf_x_forward_kernel<FT, CT, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemx));
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
