// model.cuh — host-side objects behind the opaque handles of include/nhstep.h
#pragma once
#include "kernels_basic.cuh"
#include "fft_generic.cuh"
#include "tendency_fast.cuh"
#include "tendency_plane.cuh"
#include "poisson_fast.cuh"
#include "grid_host.h"
#include <memory>
#include <map>
#include <deque>
#include <cstring>
#include <cmath>

struct nh_grid {
  nh_grid_desc desc;
  std::vector<double> zfaces;
  int device;
};

template <typename T>
struct DevBuf {
  T* p = nullptr; size_t n = 0;
  DevBuf() {}
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { if (p) cudaFree(p); }
  void alloc(size_t count, bool zero = true) {
    if (p) { cudaFree(p); p = nullptr; }
    n = count;
    if (count == 0) return;
    CUDA_CHECK(cudaMalloc(&p, count * sizeof(T)));
    if (zero) CUDA_CHECK(cudaMemset(p, 0, count * sizeof(T)));
  }
  void upload(const std::vector<T>& v) { alloc(v.size(), false); if (n) CUDA_CHECK(cudaMemcpy(p, v.data(), n * sizeof(T), cudaMemcpyHostToDevice)); }
};

// Host description of the grid at a given halo, with the device metric arrays.
template <typename FT>
struct GridHost {
  GridDev<FT> g;
  FT L[3];
  bool z_stretched = false;
  std::vector<FT> dzc_h, dzf_h;          // indexed [k + zoff]
  DevBuf<FT> dzc_d, dzf_d;
  void build(const nh_grid_desc& d, const std::vector<double>& zfaces, const int H[3]);
  long long volume() const { return (long long)g.px * g.py * g.pz; }
};

struct LaunchCounter { long long n = 0; };

// ---- Poisson solvers (complete path) ---------------------------------------------------------------------------
template <typename FT>
struct PoissonGeneric {
  typedef typename cx_of<FT>::type CT;
  int kind = NH_SOLVER_FFT;
  int N[3], topo[3];
  DevBuf<CT> A;            // FFT: storage; tridiagonal: source_term
  DevBuf<CT> Phi;          // tridiagonal: storage (keeps its content between solves, as the reference's does)
  DevBuf<double> lam[3];
  DevBuf<CT> tw[3], tw4[3];
  DevBuf<double> D;        // tridiagonal main diagonal (Float64, 3-D)
  DevBuf<FT> tt, lower, dzc1;
  DevBuf<double2> partial;
  cudaStream_t stream = 0;
  LaunchCounter* lc = nullptr;
  LaunchCounter own_lc;
  void create(const GridHost<FT>& gh, int kind_);
  void transform(CT* a, int d, int mode);
  void forward(CT* a);
  void backward(CT* a);
  CT* solve(double m);     // transforms whatever is in A; returns the array whose real part is the solution
  long long count() const { return (long long)N[0] * N[1] * N[2]; }
};

// ---- fused FFT solve for (Periodic, Periodic, Periodic), power-of-two sizes (poisson_fast.cuh) -------------------
template <typename FT>
struct PoissonFast {
  typedef typename cx_of<FT>::type CT;
  bool ok = false;
  PFJob<FT, CT> J;
  DevBuf<CT> S, twx, twy, twz;
  DevBuf<double> lx, ly, lz;
  size_t smem_x = 0, smem_y = 0, smem_z = 0;
  bool zdct = false;                 // z is Bounded: DCT-II/III along z inside the z-solve kernel
  DevBuf<CT> tw4z;
  GridDev<FT> gdense;                // the reference's dense (Nx, Ny, Nz) array layout, for standalone solves
  // x, y Periodic; z Periodic (model fast path and standalone) or Bounded (standalone); power-of-two sizes 16 … 1024
  static bool supported(const GridHost<FT>& gh, bool allow_bounded_z = false) {
    if (gh.z_stretched) return false;
    for (int e = 0; e < 3; ++e) {
      int N = gh.g.N[e];
      bool topo_ok = gh.g.topo[e] == NH_PERIODIC || (e == 2 && allow_bounded_z && gh.g.topo[e] == NH_BOUNDED);
      if (!topo_ok || N < 16 || N > 1024 || (N & (N - 1))) return false;
      if ((size_t)PF_L * ff_line_pitch(N) * sizeof(CT) > 200 * 1024) return false;
    }
    return true;
  }
  // Fourier-tridiagonal variant (standalone): x, y Periodic power-of-two; z Bounded with any spacing, any Nz ≥ 2 —
  // FFT passes in x and y around a Thomas sweep along z on the half spectrum (pf_thomas_kernel)
  bool tri = false;
  DevBuf<PFStale<FT>> stale;
  DevBuf<FT> tt, lower, dzc1;
  DevBuf<double> dA, dC;
  static bool supported_tridiagonal(const GridHost<FT>& gh) {
    for (int e = 0; e < 2; ++e) {
      int N = gh.g.N[e];
      if (gh.g.topo[e] != NH_PERIODIC || N < 16 || N > 1024 || (N & (N - 1))) return false;
    }
    return gh.g.topo[2] == NH_BOUNDED && gh.g.N[2] >= 2;
  }
  static void tri_tables(const GridHost<FT>& gh, std::vector<double>& a, std::vector<double>& c, std::vector<FT>& low,
                         std::vector<FT>& dz1) {
    const GridDev<FT>& g = gh.g;
    const int Nz = g.N[2];
    auto Dzf = [&](int k) { return gh.dzf_h[k + g.zoff]; };
    auto Dzc = [&](int k) { return gh.dzc_h[k + g.zoff]; };
    a.assign(Nz, 0.0); c.assign(Nz, 0.0); dz1.assign(Nz, FT(0)); low.clear();
    a[0] = (double)(FT(-1) / Dzf(2));
    for (int k = 2; k <= Nz - 1; ++k) a[k - 1] = (double)(-(FT(1) / Dzf(k + 1) + FT(1) / Dzf(k)));
    a[Nz - 1] = (double)(FT(-1) / Dzf(Nz));
    for (int k = 1; k <= Nz; ++k) { c[k - 1] = (double)Dzc(k); dz1[k - 1] = Dzc(k); }
    for (int q = 2; q <= Nz; ++q) low.push_back(FT(1) / Dzf(q));
  }
  void create(const GridHost<FT>& gh, bool tridiagonal = false) {
    const GridDev<FT>& g = gh.g;
    tri = tridiagonal;
    zdct = !tri && g.topo[2] == NH_BOUNDED;
    gdense = g;
    for (int e = 0; e < 3; ++e) gdense.H[e] = 0;
    gdense.ox = gdense.oy = gdense.oz = 0; gdense.px = g.N[0]; gdense.py = g.N[1]; gdense.pz = g.N[2];
    gdense.sy = g.N[0]; gdense.sz = (long long)g.N[0] * g.N[1];
    J.g = g; J.Nx = g.N[0]; J.Ny = g.N[1]; J.Nz = g.N[2]; J.Mx = J.Nx / 2 + 2; J.Cx = J.Mx; J.ywrap = 1;
    for (int d = 0; d < PF_MAXR; ++d) J.peer[d] = nullptr;
    J.p_next = nullptr;
    J.px = ff_make_plan(J.Nx); J.py = ff_make_plan(J.Ny); J.pz = ff_make_plan(tri ? 16 : J.Nz);
    S.alloc((size_t)J.Mx * J.Ny * J.Nz);
    DevBuf<CT>* tws[3] = {&twx, &twy, &twz};
    DevBuf<double>* ls[3] = {&lx, &ly, &lz};
    if (tri) {
      // grid-only parts of the main diagonal and the off-diagonals, in the reference's expression and precision
      // (compute_main_diagonal!, fourier_tridiagonal_poisson_solver.jl:41-51,88-96)
      std::vector<double> a, c;
      std::vector<FT> low, dz1;
      tri_tables(gh, a, c, low, dz1);
      dA.upload(a); dC.upload(c); lower.upload(low); dzc1.upload(dz1);
      tt.alloc((size_t)J.Mx * J.Ny * J.Nz);
      stale.alloc(1);
    }
    for (int e = 0; e < (tri ? 2 : 3); ++e) {
      int N = g.N[e];
      std::vector<double> a, b;
      nhhost::twiddles(N, a, b);
      std::vector<CT> t(N);
      for (int q = 0; q < N; ++q) { t[q].x = (FT)a[2 * q]; t[q].y = (FT)a[2 * q + 1]; }
      tws[e]->upload(t);
      double h = (double)(FT)(gh.L[e] / (FT)N);
      ls[e]->upload(nhhost::poisson_eigenvalues(N, h, g.topo[e]));
      if (e == 2 && zdct) {
        std::vector<CT> t4(N);
        for (int q = 0; q < N; ++q) { t4[q].x = (FT)b[2 * q]; t4[q].y = (FT)b[2 * q + 1]; }
        tw4z.upload(t4);
      }
    }
    J.tw4z = tw4z.p; J.dzscale = nullptr;
    J.twx = twx.p; J.twy = twy.p; J.twz = twz.p; J.lx = lx.p; J.ly = ly.p; J.lz = lz.p; J.S = S.p;
    smem_x = (size_t)PF_L * ff_line_pitch(J.Nx) * sizeof(CT);
    smem_y = (size_t)PF_L * ff_line_pitch(J.Ny) * sizeof(CT);
    smem_z = (size_t)PF_L * ff_line_pitch(J.Nz) * sizeof(CT);
    yz_dispatch(0, 0, 0);
    ok = true;
  }
  // size-templated kernels: what = 0 set attributes, 1 y-forward, 2 z-solve, 3 y-inverse, 4 x-forward, 5 x-inverse
  const FT *xu = nullptr, *xv = nullptr, *xw = nullptr; FT* xp = nullptr;
  template <int N> void yz_launch(int what, dim3 grid, cudaStream_t stream) {
    const int smem = PF_L * FFPitch8<N>::value * (int)sizeof(CT);
    const int smemx = PF_L * ((N + N / 8 + N / 64 + 1) | 1) * (int)sizeof(CT);
    if (what == 0) {
      if (J.Ny == N) {
        CUDA_CHECK(cudaFuncSetAttribute(pf_y_kernel<FT, CT, N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CUDA_CHECK(cudaFuncSetAttribute(pf_y_kernel<FT, CT, N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      }
      if (J.Nz == N) {
        CUDA_CHECK(cudaFuncSetAttribute(pf_z_solve_kernel<FT, CT, N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CUDA_CHECK(cudaFuncSetAttribute(pf_z_solve_kernel<FT, CT, N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      }
      if (J.Nx == N) {
        CUDA_CHECK(cudaFuncSetAttribute(pf_x_forward_kernel<FT, CT, N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemx));
        CUDA_CHECK(cudaFuncSetAttribute(pf_x_forward_kernel<FT, CT, N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemx));
        CUDA_CHECK(cudaFuncSetAttribute(pf_x_inverse_kernel<FT, CT, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemx));
      }
    } else if (what == 4) pf_x_forward_kernel<FT, CT, N, false><<<grid, N, smemx, stream>>>(J, xu, xv, xw);
    else if (what == 6) pf_x_forward_kernel<FT, CT, N, true><<<grid, N, smemx, stream>>>(J, xu, nullptr, nullptr);
    else if (what == 5) pf_x_inverse_kernel<FT, CT, N><<<grid, N, smemx, stream>>>(J, xp);
    else if (what == 1) pf_y_kernel<FT, CT, N, false><<<grid, N, smem, stream>>>(J);
    else if (what == 2 && zdct) pf_z_solve_kernel<FT, CT, N, true><<<grid, N, smem, stream>>>(J);
    else if (what == 2) pf_z_solve_kernel<FT, CT, N, false><<<grid, N, smem, stream>>>(J);
    else pf_y_kernel<FT, CT, N, true><<<grid, N, smem, stream>>>(J);
  }
  void yz_size(int N, int what, dim3 grid, cudaStream_t stream) {
    switch (N) {
      case 16: yz_launch<16>(what, grid, stream); break;
      case 32: yz_launch<32>(what, grid, stream); break;
      case 64: yz_launch<64>(what, grid, stream); break;
      case 128: yz_launch<128>(what, grid, stream); break;
      case 256: yz_launch<256>(what, grid, stream); break;
      case 512: yz_launch<512>(what, grid, stream); break;
      case 1024: yz_launch<1024>(what, grid, stream); break;
      default: throw NhError("unsupported FFT size on the fused path");
    }
  }
  void yz_dispatch(int what, dim3 grid, cudaStream_t stream) {
    if (what == 0) {
      yz_size(J.Ny, 0, grid, stream);
      if (J.Nz != J.Ny && !tri) yz_size(J.Nz, 0, grid, stream);
      if (J.Nx != J.Ny && (tri || J.Nx != J.Nz)) yz_size(J.Nx, 0, grid, stream);
    } else yz_size(what >= 4 ? J.Nx : (what == 2 ? J.Nz : J.Ny), what, grid, stream);   // 4, 5, 6 are x passes
  }
  // rhs = div(u,v,w)/dt → p (interior of the halo'd field); 5 launches on `stream`
  void solve(const FT* u, const FT* v, const FT* w, FT* p, double dt, cudaStream_t stream, LaunchCounter& lc) {
    J.dt = dt; J.m = 0.0;
    dim3 gx(ceil_div(J.Ny, 2 * PF_L), J.Nz), gy(ceil_div(J.Nx / 2 + 1, PF_L), J.Nz), gz(ceil_div(J.Nx / 2 + 1, PF_L), J.Ny);
    xu = u; xv = v; xw = w; xp = p;
    yz_dispatch(4, gx, stream);
    yz_dispatch(1, gy, stream);
    yz_dispatch(2, gz, stream);
    yz_dispatch(3, gy, stream);
    yz_dispatch(5, gx, stream);
    lc.n += 5;
    CUDA_CHECK(cudaGetLastError());
  }
  // solve!(ϕ, solver, b, m) on dense device arrays (fft_based_poisson_solver.jl:95-125): ϕ = Re(𝓕⁻¹[-𝓕b / (λ - m)])
  // `out` = layout of ϕ (nullptr: dense like b; a model passes its halo'd field geometry and gets the interior written)
  void solve_rhs(const FT* rhs, FT* phi, double m, cudaStream_t stream, LaunchCounter& lc, const GridDev<FT>* out = nullptr) {
    const GridDev<FT> keep = J.g;
    J.g = gdense; J.dt = 1.0; J.m = m;
    dim3 gx(ceil_div(J.Ny, 2 * PF_L), J.Nz), gy(ceil_div(J.Nx / 2 + 1, PF_L), J.Nz), gz(ceil_div(J.Nx / 2 + 1, PF_L), J.Ny);
    xu = rhs; xv = xw = nullptr; xp = phi;
    yz_dispatch(6, gx, stream);
    yz_dispatch(1, gy, stream);
    yz_dispatch(2, gz, stream);
    yz_dispatch(3, gy, stream);
    if (out) J.g = *out;
    yz_dispatch(5, gx, stream);
    J.g = keep;
    lc.n += 5;
    CUDA_CHECK(cudaGetLastError());
  }
  // FourierTridiagonalPoissonSolver: set_source_term! (b·Δz) + solve! (fourier_tridiagonal_poisson_solver.jl:119-147)
  // times_dz = false: b already carries the Δz factor (the model's _fourier_tridiagonal_source_term!)
  void solve_tridiagonal(const FT* rhs, FT* phi, cudaStream_t stream, LaunchCounter& lc, bool times_dz = true,
                         const GridDev<FT>* out = nullptr) {
    const GridDev<FT> keep = J.g;
    J.g = gdense; J.dt = 1.0; J.m = 0.0; J.dzscale = times_dz ? dzc1.p : nullptr; J.S = S.p;
    dim3 gx(ceil_div(J.Ny, 2 * PF_L), J.Nz), gy(ceil_div(J.Nx / 2 + 1, PF_L), J.Nz);
    xu = rhs; xv = xw = nullptr; xp = phi;
    yz_dispatch(6, gx, stream);
    yz_dispatch(1, gy, stream);
    dim3 gt(ceil_div(J.Nx / 2 + 1, 128), J.Ny);
    pf_thomas_kernel<FT, CT><<<gt, 128, 0, stream>>>(J, lower.p, dA.p, dC.p, S.p, tt.p, stale.p);
    yz_dispatch(3, gy, stream);
    if (out) J.g = *out;
    yz_dispatch(5, gx, stream);
    pf_stale_gather_kernel<FT><<<1, PF_STALE_MAX, 0, stream>>>(stale.p, phi, J.g);
    J.dzscale = nullptr; J.g = keep;
    lc.n += 6;
    CUDA_CHECK(cudaGetLastError());
  }
};

#include "dist_host.cuh"

// ---- model -----------------------------------------------------------------------------------------------------
struct nh_model {
  virtual ~nh_model() {}
  int float_type = NH_F64;
  int device = 0;
  nh_model_desc desc;
  nh_grid_desc gdesc;      // with the inflated halo
  nh_clock clock;
  LaunchCounter lc;
  cudaStream_t stream = 0;
  virtual void field_size(int field, int region, int64_t size[3]) = 0;
  virtual void set_field(int field, int region, const void* data, int64_t count, int where) = 0;
  virtual void get_field(int field, int region, void* data, int64_t count, int where) = 0;
  virtual void fill_halo_regions(uint32_t mask, int fill_normal) = 0;
  virtual void set_boundary_condition(int field, int dim, int side, int kind, double value) = 0;
  virtual void set_boundary_condition_array(int field, int dim, int side, int kind, const void* data, long long count, int where) = 0;
  virtual void set_weno_z_coefficients(int buffer, int location, const void* data, long long count, int where) = 0;
  virtual int dist_mode() = 0;
  virtual void update_state(int compute_tendencies) = 0;
  virtual void compute_tendencies() = 0;
  virtual void rk3_substep(double dt, double gamma, double zeta, int has_zeta) = 0;
  virtual void ab2_step(double dt, double chi) = 0;
  virtual void cache_previous_tendencies() = 0;
  virtual void calculate_pressure_correction(double dt) = 0;
  virtual void pressure_correct_velocities(double dt) = 0;
  virtual void time_step(double dt, int euler) = 0;
  virtual void set_state_finalize(int enforce) = 0;
  virtual double max_abs_divergence() = 0;
  virtual void field_device_ptr(int field, void** ptr, int64_t* off, int64_t strides[3]) = 0;
  virtual void set_stream(cudaStream_t s) { stream = s; }
  // y-slab multi-GPU pieces (dist.cu); the messages between them are issued by the host driver
  virtual void dist_configure(int rank, int nranks, double y0, double y1, int Ny) = 0;
  virtual void dist_buffer(int which, void** ptr, int64_t* bytes) = 0;
  virtual int64_t dist_halo(uint32_t mask, int w, int unpack) = 0;
  virtual double dist_tendency_substep(double dt, int stage, int euler, int part = 0) = 0;
  virtual void dist_poisson(int phase, double stage_dt) = 0;
  virtual void dist_pressure_correct(double stage_dt) = 0;
  virtual int64_t dist_ipc_export(unsigned char* blob, int64_t capacity) = 0;
  virtual void dist_ipc_import(const unsigned char* blobs, int64_t nbytes) = 0;
  virtual void dist_barrier() = 0;
  virtual void dist_transpose(int direction) = 0;
  // phase timing (nh_time_step_profiled): CUDA events on the model's stream between the phases of one step
  bool profiling = false;
  std::vector<std::pair<int, cudaEvent_t>> marks;
  void mark(int phase) {
    if (!profiling) return;
    cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, stream); marks.emplace_back(phase, e);
  }
  void profiled_step(double dt, int euler, double ms[4]) {
    profiling = true; marks.clear();
    mark(-1);
    time_step(dt, euler);
    mark(3);
    cudaStreamSynchronize(stream);
    for (int q = 0; q < 4; ++q) ms[q] = 0;
    for (size_t q = 1; q < marks.size(); ++q) {
      float t = 0; cudaEventElapsedTime(&t, marks[q - 1].second, marks[q].second);
      ms[marks[q].first] += t;
    }
    for (auto& m : marks) cudaEventDestroy(m.second);
    marks.clear(); profiling = false;
  }
};

template <typename FT>
struct Model : nh_model {
  typedef typename cx_of<FT>::type CT;
  GridHost<FT> gh;
  int ntr = 0, nf = 3;
  std::vector<std::unique_ptr<DevBuf<FT>>> bufs;
  FT* F[NH_MAXF];          // prognostic fields u,v,w,tracers
  FT* Gn[NH_MAXF];
  FT* Gm[NH_MAXF];
  FT* pNHS = nullptr;
  FT* pHY = nullptr;
  FT* nue = nullptr;       // νₑ of the Smagorinsky / AMD closures (diffusivity_fields.νₑ)
  FT* kape[NH_MAXTR] = {}; // AMD: κₑ of each tracer (diffusivity_fields.κₑ)
  bool vitd = false;       // ScalarDiffusivity with VerticallyImplicitTimeDiscretization
  struct ImplicitTable { double dt = 0, kappa = 0; int w_field = 0, ct = 0; DevBuf<double> tab; };
  std::deque<ImplicitTable> vitd_tabs;   // per-level Thomas coefficients, cached per (Δt, κ, location, coefficient type)
  size_t vitd_next = 0;
  DevBuf<FT> df2;          // Δᶠ² = cbrt(ΔxΔyΔz)² per k
  void compute_diffusivities();
  SchemeDev adv[3];
  SchemeTables<FT> tab;
  PoissonGeneric<FT> solver;
  PoissonFast<FT> pfast;   // fused solver passes where the grid allows (else the complete `solver` does the work)
  DevBuf<FT> prhs;         // dense real source term for them on the complete path
  DevBuf<FT> stage;        // staging for host <-> device field copies
  DevBuf<double> red;      // reduction scratch
  FT rk_g1, rk_g2, rk_g3, rk_z2, rk_z3, chi;
  // fused fast path (triply periodic, regular, WENO5, no tracers): ping-pong state, lazily refreshed Gⁿ and p halos
  bool fast = false, gn_stale = false, p_halo_stale = false;
  bool tiled = true;       // complete path: tiled tendency kernel (each face flux once) instead of the per-cell one
  FT* F2[3] = {nullptr, nullptr, nullptr};
  void fast_stage(int mode, FT dt, FT c1, FT c2, double stage_dt);
  void fast_pressure(double stage_dt);
  void ensure_tendencies();
  void ensure_solver();
  // fused tendency + substep sweep: plane-marching TMA kernel (tendency_plane.cuh) or the tile kernel (tendency_fast.cuh)
  void launch_fused_sweep(const FastTendJob<FT>& J, int zseg = 0);      // zseg > 0: shorter z segments (few tile rows: more CTAs)
  bool plane_sweep = true;
  bool tiled_fast_ok = false;        // the tiled tendency kernel may use its fast flux stage on interior tiles
  int plane_zseg = 32;
  std::map<const void*, CUtensorMap> tmaps;
  void build_filter_width();
  bool solver_ready = false;
  void ensure_p_halo();

  Model(const nh_grid& grid, const nh_model_desc& d);
  FT* alloc_field();
  FT* field_ptr(int field);
  int field_face_dim(int field) const;         // 0,1,2 for (Gⁿ/G⁻ of) u,v,w; -1 for cell-centred fields
  void field_extent(int field, int n[3]) const;
  TendParams<FT> tend_params() const;
  dim3 grid3(dim3 block) const { return dim3(ceil_div(gh.g.N[0], block.x), ceil_div(gh.g.N[1], block.y), gh.g.N[2]); }

  void field_size(int field, int region, int64_t size[3]) override;
  void set_field(int field, int region, const void* data, int64_t count, int where) override;
  void get_field(int field, int region, void* data, int64_t count, int where) override;
  void fill_halo_regions(uint32_t mask, int fill_normal) override;
  void set_boundary_condition(int field, int dim, int side, int kind, double value) override;
  void set_boundary_condition_array(int field, int dim, int side, int kind, const void* data, long long count, int where) override;
  void set_weno_z_coefficients(int buffer, int location, const void* data, long long count, int where) override;
  DevBuf<FT> weno_z[2][2];   // WENO(grid) on a stretched z: [location Face / Center][buffer − 2] (TendParams::wz)
  bool weno_stretched = false;
  int dist_mode() override { return !dist.on ? 0 : (dist.generic ? 2 : 1); }
  FT* halo_field_or_null(int id) {       // the field of a halo-fill mask bit, nullptr when the model has none
    if (id >= 0 && id < nf) return F[id];
    if (id == NH_FIELD_PNHS) return pNHS;
    if (id == NH_FIELD_PHY) return pHY;
    if (id == NH_FIELD_NUE) return nue;
    if (id >= NH_FIELD_KAPPAE0 && id < NH_FIELD_KAPPAE0 + NH_MAXTR && id - NH_FIELD_KAPPAE0 < ntr) return kape[id - NH_FIELD_KAPPAE0];
    return nullptr;
  }
  int bc_kind[NH_MAXF][3][2] = {};   // NH_BC_* of each prognostic field's walls (constant conditions)
  FT bc_val[NH_MAXF][3][2] = {};
  DevBuf<FT> bc_arr[NH_MAXF][3][2];  // array conditions (empty: the number above)
  bool has_flux_bcs = false;
  void update_state(int compute_tendencies) override;
  void compute_tendencies() override;
  void rk3_substep(double dt, double gamma, double zeta, int has_zeta) override;
  void ab2_step(double dt, double chi) override;
  void cache_previous_tendencies() override;
  void calculate_pressure_correction(double dt) override;
  void pressure_correct_velocities(double dt) override;
  void time_step(double dt, int euler) override;
  void set_state_finalize(int enforce) override;
  double max_abs_divergence() override;
  void field_device_ptr(int field, void** ptr, int64_t* off, int64_t strides[3]) override;
  void set_stream(cudaStream_t s) override { stream = s; solver.stream = s; }
  void substep_launch(int mode, FT dt, FT c1, FT c2);
  template <typename CT2> void implicit_launch(double stage_dt);
  void implicit_steps(double stage_dt, bool ab2);
  void update_hydrostatic_pressure();
  DistState<FT> dist;
  void dist_configure(int rank, int nranks, double y0, double y1, int Ny) override;
  void dist_buffer(int which, void** ptr, int64_t* bytes) override;
  int64_t dist_halo(uint32_t mask, int w, int unpack) override;
  double dist_tendency_substep(double dt, int stage, int euler, int part = 0) override;
  void dist_poisson(int phase, double stage_dt) override;
  void dist_pressure_correct(double stage_dt) override;
  int64_t dist_ipc_export(unsigned char* blob, int64_t capacity) override;
  void dist_ipc_import(const unsigned char* blobs, int64_t nbytes) override;
  void dist_barrier() override;
  void dist_transpose(int direction) override;
};

struct nh_poisson {
  virtual ~nh_poisson() {}
  virtual void solve(const void* rhs, void* phi, double m, int where) = 0;
  virtual void sync() = 0;
  virtual long long launches() const = 0;
  virtual float time_solves(const void* rhs, void* phi, int n) = 0;
  int device = 0;
};

struct nh_tridiag {
  virtual ~nh_tridiag() {}
  virtual void solve(const void* rhs, void* phi) = 0;
  int device = 0;
};
