// api.cu — the extern "C" surface of libnhstep.so (include/nhstep.h)
#include "model.cuh"

static thread_local std::string g_last_error;
void nh_set_error(const std::string& msg) { g_last_error = msg; }

#define NH_TRY try {
#define NH_CATCH } catch (const std::exception& e) { nh_set_error(e.what()); return 1; } catch (...) { nh_set_error("unknown error"); return 1; } return 0;

template <typename FT>
struct PoissonStandalone : nh_poisson {
  typedef typename cx_of<FT>::type CT;
  GridHost<FT> gh;
  PoissonGeneric<FT> s;
  DevBuf<FT> in, out;
  cudaStream_t stream = 0;
  PoissonStandalone(const nh_grid& grid, int kind) {
    device = grid.device;
    CUDA_CHECK(cudaSetDevice(device));
    gh.build(grid.desc, grid.zfaces, grid.desc.H);
    s.stream = stream;
    s.create(gh, kind);
    // fused passes for (Periodic, Periodic, Periodic | Bounded) power-of-two grids; everything else takes the complete path
    if (getenv("NHSTEP_NO_FAST_POISSON")) return;
    if (kind == NH_SOLVER_FFT && PoissonFast<FT>::supported(gh, true)) fast.create(gh);
    if (kind == NH_SOLVER_FOURIER_TRIDIAGONAL && PoissonFast<FT>::supported_tridiagonal(gh)) fast.create(gh, true);
  }
  PoissonFast<FT> fast;
  void run(const FT* rhs_dev, FT* phi_dev, double m) {
    if (fast.ok && fast.tri) { fast.solve_tridiagonal(rhs_dev, phi_dev, stream, *s.lc); return; }
    if (fast.ok) { fast.solve_rhs(rhs_dev, phi_dev, m, stream, *s.lc); return; }
    long long n = s.count();
    int blocks = (int)((n + 255) / 256);
    real_to_complex_kernel<FT, CT><<<blocks, 256, 0, stream>>>(rhs_dev, s.A.p, n, (long long)s.N[0] * s.N[1], s.dzc1.p,
                                                              s.kind == NH_SOLVER_FOURIER_TRIDIAGONAL);
    CT* sol = s.solve(m);
    complex_real_part_kernel<FT, CT><<<blocks, 256, 0, stream>>>(sol, phi_dev, n);
    s.lc->n += 2;
    CUDA_CHECK(cudaGetLastError());
  }
  void solve(const void* rhs, void* phi, double m, int where) override {
    CUDA_CHECK(cudaSetDevice(device));
    long long n = s.count();
    if (where == NH_DEVICE) { run((const FT*)rhs, (FT*)phi, m); return; }
    if (in.n < (size_t)n) { in.alloc(n, false); out.alloc(n, false); }
    CUDA_CHECK(cudaMemcpyAsync(in.p, rhs, n * sizeof(FT), cudaMemcpyHostToDevice, stream));
    run(in.p, out.p, m);
    CUDA_CHECK(cudaMemcpyAsync(phi, out.p, n * sizeof(FT), cudaMemcpyDeviceToHost, stream));
    CUDA_CHECK(cudaStreamSynchronize(stream));
  }
  void sync() override { CUDA_CHECK(cudaStreamSynchronize(stream)); }
  long long launches() const override { return s.lc->n; }
  float time_solves(const void* rhs, void* phi, int n) override {
    cudaEvent_t a, b;
    CUDA_CHECK(cudaEventCreate(&a)); CUDA_CHECK(cudaEventCreate(&b));
    CUDA_CHECK(cudaEventRecord(a, stream));
    for (int q = 0; q < n; ++q) run((const FT*)rhs, (FT*)phi, 0.0);
    CUDA_CHECK(cudaEventRecord(b, stream));
    CUDA_CHECK(cudaEventSynchronize(b));
    float ms = 0; CUDA_CHECK(cudaEventElapsedTime(&ms, a, b));
    cudaEventDestroy(a); cudaEventDestroy(b);
    return ms;
  }
};

template <typename FT>
struct TridiagStandalone : nh_tridiag {
  int N[3], dir, b3d;
  DevBuf<FT> a, b, c, f, phi, tt;
  TridiagStandalone(const nh_grid& grid, const double* ah, const double* bh, const double* ch, int b_is_3d, int direction) {
    device = grid.device;
    CUDA_CHECK(cudaSetDevice(device));
    for (int e = 0; e < 3; ++e) N[e] = grid.desc.N[e];
    dir = direction; b3d = b_is_3d;
    NH_REQUIRE(dir >= 0 && dir < 3, "bad tridiagonal direction");
    int n = N[dir];
    long long tot = (long long)N[0] * N[1] * N[2];
    std::vector<FT> av(std::max(1, n - 1)), cv(std::max(1, n - 1)), bv(b3d ? tot : n);
    for (int q = 0; q < n - 1; ++q) { av[q] = (FT)ah[q]; cv[q] = (FT)ch[q]; }
    for (size_t q = 0; q < bv.size(); ++q) bv[q] = (FT)bh[q];
    a.upload(av); b.upload(bv); c.upload(cv);
    f.alloc(tot); phi.alloc(tot); tt.alloc(tot);
  }
  void solve(const void* rhs, void* out) override {
    CUDA_CHECK(cudaSetDevice(device));
    long long tot = (long long)N[0] * N[1] * N[2];
    CUDA_CHECK(cudaMemcpy(f.p, rhs, tot * sizeof(FT), cudaMemcpyHostToDevice));
    long long lines = tot / N[dir];
    thomas_generic_kernel<FT><<<(int)((lines + 127) / 128), 128>>>(N[0], N[1], N[2], dir, a.p, b.p, b3d, c.p, f.p, phi.p, tt.p);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpy(out, phi.p, tot * sizeof(FT), cudaMemcpyDeviceToHost));
  }
};

extern "C" {

const char* nh_last_error(void) { return g_last_error.c_str(); }
int32_t nh_version(void) { return 100; }
int32_t nh_device_count(int32_t* count) {
  NH_TRY
  int n = 0; cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) { n = 0; cudaGetLastError(); }
  *count = n;
  NH_CATCH
}

int32_t nh_grid_create(const nh_grid_desc* desc, int32_t device, nh_grid** out) {
  NH_TRY
  NH_REQUIRE(desc && out, "null argument");
  NH_REQUIRE(desc->float_type == NH_F32 || desc->float_type == NH_F64, "float_type must be NH_F32 or NH_F64");
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); throw NhError("no CUDA device available: libnhstep has no CPU fallback"); }
  NH_REQUIRE(device >= 0 && device < n, "bad device index");
  std::unique_ptr<nh_grid> g(new nh_grid());
  g->desc = *desc;
  g->device = device;
  for (int e = 0; e < 3; ++e) {
    NH_REQUIRE(desc->N[e] >= 1, "size must be >= 1");
    NH_REQUIRE(desc->topology[e] >= 0 && desc->topology[e] <= 2, "topology must be Periodic, Bounded or Flat");
    if (desc->topology[e] == NH_FLAT) g->desc.H[e] = 0;
    NH_REQUIRE(g->desc.H[e] >= 0, "halo must be >= 0");
  }
  if (desc->z_faces && desc->topology[2] != NH_FLAT) {
    g->zfaces.assign(desc->z_faces, desc->z_faces + desc->N[2] + 1);
    for (int k = 0; k < desc->N[2]; ++k) NH_REQUIRE(g->zfaces[k + 1] >= g->zfaces[k], "The values of z must increase as the index increases!");
  }
  g->desc.z_faces = nullptr;
  const double* iv[3] = {desc->x, desc->y, desc->z};
  for (int e = 0; e < 3; ++e) {
    if (desc->topology[e] == NH_FLAT || (e == 2 && !g->zfaces.empty())) continue;
    NH_REQUIRE(iv[e][0] < iv[e][1], "coordinate interval must be increasing");
  }
  *out = g.release();
  NH_CATCH
}
int32_t nh_grid_destroy(nh_grid* grid) { delete grid; return 0; }
int32_t nh_grid_info(const nh_grid* grid, nh_grid_desc* out) { NH_TRY NH_REQUIRE(grid && out, "null argument"); *out = grid->desc; NH_CATCH }

int32_t nh_grid_spacings(const nh_grid* grid, double* dxdy, double* dzc, double* dzf, int64_t nz_entries) {
  NH_TRY
  NH_REQUIRE(grid, "null argument");
  CUDA_CHECK(cudaSetDevice(grid->device));
  int Nz = grid->desc.N[2], Hz = grid->desc.H[2];
  NH_REQUIRE(nz_entries == Nz + 2 * Hz, "nz_entries must be Nz + 2Hz");
  auto fill = [&](auto tag) {
    typedef decltype(tag) FT;
    GridHost<FT> gh; gh.build(grid->desc, grid->zfaces, grid->desc.H);
    dxdy[0] = (double)gh.g.dx; dxdy[1] = (double)gh.g.dy;
    for (int k = 1 - Hz; k <= Nz + Hz; ++k) { dzc[k - 1 + Hz] = (double)gh.dzc_h[k + gh.g.zoff]; dzf[k - 1 + Hz] = (double)gh.dzf_h[k + gh.g.zoff]; }
  };
  if (grid->desc.float_type == NH_F32) fill(float()); else fill(double());
  NH_CATCH
}

int32_t nh_reconstruction_coefficients(int32_t float_type, int32_t which, int32_t buffer, int32_t r, double out[8], int32_t* count) {
  NH_TRY
  NH_REQUIRE(out && count, "null argument");
  NH_REQUIRE(float_type == NH_F32 || float_type == NH_F64, "float_type must be NH_F32 or NH_F64");
  auto fill = [&](auto tag) {
    typedef decltype(tag) FT;
    nhhost::Tables<FT> T;
    nhhost::make_scheme_tables<FT>(T);
    int n = 0;
    const FT* src = nullptr;
    if (which == 0) { NH_REQUIRE(buffer >= 1 && buffer <= 2, "symmetric tables exist for buffer 1, 2"); src = T.sym[buffer]; n = 2 * buffer; }
    else if (which == 1 || which == 2) { NH_REQUIRE(buffer >= 1 && buffer <= 3, "upwind tables exist for buffer 1..3"); src = which == 1 ? T.upl[buffer] : T.upr[buffer]; n = 2 * buffer - 1; }
    else if (which == 3) { NH_REQUIRE(buffer >= 2 && buffer <= 3 && r >= 0 && r < buffer, "WENO tables exist for buffer 2, 3 and 0 <= r < buffer"); src = T.wc[buffer][r]; n = buffer; }
    else if (which == 4) { NH_REQUIRE(buffer >= 2 && buffer <= 3, "WENO tables exist for buffer 2, 3"); src = T.cstar[buffer]; n = buffer; }
    else throw NhError("which must be 0..4");
    for (int q = 0; q < n; ++q) out[q] = (double)src[q];
    *count = n;
  };
  if (float_type == NH_F32) fill(float()); else fill(double());
  NH_CATCH
}

int32_t nh_model_create(nh_grid* grid, const nh_model_desc* desc, nh_model** out) {
  NH_TRY
  NH_REQUIRE(grid && desc && out, "null argument");
  if (grid->desc.float_type == NH_F32) *out = new Model<float>(*grid, *desc);
  else *out = new Model<double>(*grid, *desc);
  NH_CATCH
}
int32_t nh_model_destroy(nh_model* m) { if (m) { cudaSetDevice(m->device); cudaDeviceSynchronize(); delete m; } return 0; }
int32_t nh_model_grid_info(const nh_model* m, nh_grid_desc* out) { NH_TRY NH_REQUIRE(m && out, "null argument"); *out = m->gdesc; NH_CATCH }

#define MODEL_CALL(m, body) NH_TRY NH_REQUIRE(m, "null model"); CUDA_CHECK(cudaSetDevice(m->device)); body; NH_CATCH

int32_t nh_field_size(const nh_model* m, int32_t field, int32_t region, int64_t size[3]) { MODEL_CALL(m, const_cast<nh_model*>(m)->field_size(field, region, size)) }
int32_t nh_set_field(nh_model* m, int32_t field, int32_t region, const void* data, int64_t count, int32_t where) { MODEL_CALL(m, m->set_field(field, region, data, count, where)) }
int32_t nh_get_field(nh_model* m, int32_t field, int32_t region, void* data, int64_t count, int32_t where) { MODEL_CALL(m, m->get_field(field, region, data, count, where)) }
int32_t nh_set_state_finalize(nh_model* m, int32_t enforce) { MODEL_CALL(m, m->set_state_finalize(enforce)) }
int32_t nh_fill_halo_regions(nh_model* m, uint32_t mask, int32_t fill_normal) { MODEL_CALL(m, m->fill_halo_regions(mask, fill_normal)) }
int32_t nh_model_set_boundary_condition(nh_model* m, int32_t field, int32_t dim, int32_t side, int32_t kind, double value) { MODEL_CALL(m, m->set_boundary_condition(field, dim, side, kind, value)) }
int32_t nh_model_set_boundary_condition_array(nh_model* m, int32_t field, int32_t dim, int32_t side, int32_t kind, const void* data, int64_t count, int32_t where) { MODEL_CALL(m, m->set_boundary_condition_array(field, dim, side, kind, data, count, where)) }
int32_t nh_model_set_weno_z_coefficients(nh_model* m, int32_t buffer, int32_t location, const void* data, int64_t count, int32_t where) { MODEL_CALL(m, m->set_weno_z_coefficients(buffer, location, data, count, where)) }
int32_t nh_update_state(nh_model* m, int32_t compute) { MODEL_CALL(m, m->update_state(compute)) }
int32_t nh_compute_tendencies(nh_model* m) { MODEL_CALL(m, m->compute_tendencies()) }
int32_t nh_rk3_substep(nh_model* m, double dt, double gamma, double zeta, int32_t has_zeta) { MODEL_CALL(m, m->rk3_substep(dt, gamma, zeta, has_zeta)) }
int32_t nh_ab2_step(nh_model* m, double dt, double chi) { MODEL_CALL(m, m->ab2_step(dt, chi)) }
int32_t nh_cache_previous_tendencies(nh_model* m) { MODEL_CALL(m, m->cache_previous_tendencies()) }
int32_t nh_calculate_pressure_correction(nh_model* m, double dt) { MODEL_CALL(m, m->calculate_pressure_correction(dt)) }
int32_t nh_pressure_correct_velocities(nh_model* m, double dt) { MODEL_CALL(m, m->pressure_correct_velocities(dt)) }
int32_t nh_time_step(nh_model* m, double dt, int32_t euler, int32_t n_steps) { MODEL_CALL(m, for (int q = 0; q < (n_steps < 1 ? 1 : n_steps); ++q) m->time_step(dt, euler)) }
int32_t nh_time_step_profiled(nh_model* m, double dt, int32_t euler, double phase_ms[4]) { MODEL_CALL(m, m->profiled_step(dt, euler, phase_ms)) }
int32_t nh_clock_get(const nh_model* m, nh_clock* out) { NH_TRY NH_REQUIRE(m && out, "null argument"); *out = m->clock; NH_CATCH }
int32_t nh_clock_set(nh_model* m, const nh_clock* in) { NH_TRY NH_REQUIRE(m && in, "null argument"); m->clock = *in; NH_CATCH }
int32_t nh_max_abs_divergence(nh_model* m, double* out) { MODEL_CALL(m, *out = m->max_abs_divergence()) }
int32_t nh_synchronize(nh_model* m) { MODEL_CALL(m, CUDA_CHECK(cudaStreamSynchronize(m->stream))) }
int32_t nh_field_device_ptr(nh_model* m, int32_t field, void** ptr, int64_t off[1], int64_t strides[3]) { MODEL_CALL(m, m->field_device_ptr(field, ptr, off, strides)) }
int32_t nh_model_set_stream(nh_model* m, void* s) { NH_TRY NH_REQUIRE(m, "null model"); m->set_stream((cudaStream_t)s); NH_CATCH }
int32_t nh_dist_mode(nh_model* m, int32_t* mode) { MODEL_CALL(m, *mode = m->dist_mode()) }
int32_t nh_dist_configure(nh_model* m, int32_t rank, int32_t nranks, double y0, double y1, int32_t Ny) { MODEL_CALL(m, m->dist_configure(rank, nranks, y0, y1, Ny)) }
int32_t nh_dist_buffer(nh_model* m, int32_t which, void** ptr, int64_t* bytes) { MODEL_CALL(m, NH_REQUIRE(ptr && bytes, "null argument"); m->dist_buffer(which, ptr, bytes)) }
int32_t nh_dist_halo(nh_model* m, uint32_t mask, int32_t w, int32_t unpack, int64_t* elements) { MODEL_CALL(m, int64_t n = m->dist_halo(mask, w, unpack); if (elements) *elements = n) }
int32_t nh_dist_tendency_substep_part(nh_model* m, double dt, int32_t stage, int32_t euler, int32_t part, double* stage_dt) { MODEL_CALL(m, double s = m->dist_tendency_substep(dt, stage, euler, part); if (stage_dt) *stage_dt = s) }
int32_t nh_dist_tendency_substep(nh_model* m, double dt, int32_t stage, int32_t euler, double* stage_dt) { MODEL_CALL(m, double s = m->dist_tendency_substep(dt, stage, euler); if (stage_dt) *stage_dt = s) }
int32_t nh_dist_poisson(nh_model* m, int32_t phase, double stage_dt) { MODEL_CALL(m, m->dist_poisson(phase, stage_dt)) }
int32_t nh_dist_pressure_correct(nh_model* m, double stage_dt) { MODEL_CALL(m, m->dist_pressure_correct(stage_dt)) }
int32_t nh_dist_ipc_export(nh_model* m, unsigned char* blob, int64_t capacity, int64_t* nbytes) { MODEL_CALL(m, int64_t n = m->dist_ipc_export(blob, capacity); if (nbytes) *nbytes = n) }
int32_t nh_dist_barrier(nh_model* m) { MODEL_CALL(m, m->dist_barrier()) }
int32_t nh_dist_transpose(nh_model* m, int32_t direction) { MODEL_CALL(m, m->dist_transpose(direction)) }
int32_t nh_dist_ipc_import(nh_model* m, const unsigned char* blobs, int64_t nbytes) { MODEL_CALL(m, NH_REQUIRE(blobs, "null argument"); m->dist_ipc_import(blobs, nbytes)) }
int32_t nh_launch_count(const nh_model* m, int64_t* count) { NH_TRY NH_REQUIRE(m && count, "null argument"); *count = m->lc.n; NH_CATCH }

int32_t nh_poisson_create(nh_grid* grid, int32_t kind, nh_poisson** out) {
  NH_TRY
  NH_REQUIRE(grid && out, "null argument");
  NH_REQUIRE(kind == NH_SOLVER_FFT || kind == NH_SOLVER_FOURIER_TRIDIAGONAL, "unknown solver kind");
  if (grid->desc.float_type == NH_F32) *out = new PoissonStandalone<float>(*grid, kind);
  else *out = new PoissonStandalone<double>(*grid, kind);
  NH_CATCH
}
int32_t nh_poisson_destroy(nh_poisson* s) { if (s) { cudaSetDevice(s->device); cudaDeviceSynchronize(); delete s; } return 0; }
int32_t nh_poisson_solve(nh_poisson* s, const void* rhs, void* phi, double m, int32_t where) { NH_TRY NH_REQUIRE(s && rhs && phi, "null argument"); s->solve(rhs, phi, m, where); NH_CATCH }
int32_t nh_poisson_synchronize(nh_poisson* s) { NH_TRY NH_REQUIRE(s, "null argument"); s->sync(); NH_CATCH }
int32_t nh_poisson_launch_count(const nh_poisson* s, int64_t* count) { NH_TRY NH_REQUIRE(s && count, "null argument"); *count = s->launches(); NH_CATCH }
int32_t nh_poisson_time_solves(nh_poisson* s, const void* rhs, void* phi, int32_t n, float* ms) { NH_TRY NH_REQUIRE(s && rhs && phi && ms, "null argument"); CUDA_CHECK(cudaSetDevice(s->device)); *ms = s->time_solves(rhs, phi, n); NH_CATCH }

int32_t nh_tridiag_create(nh_grid* grid, const double* a, const double* b, const double* c, int32_t b_is_3d, int32_t direction, nh_tridiag** out) {
  NH_TRY
  NH_REQUIRE(grid && a && b && c && out, "null argument");
  if (grid->desc.float_type == NH_F32) *out = new TridiagStandalone<float>(*grid, a, b, c, b_is_3d, direction);
  else *out = new TridiagStandalone<double>(*grid, a, b, c, b_is_3d, direction);
  NH_CATCH
}
int32_t nh_tridiag_solve(nh_tridiag* s, const void* rhs, void* phi) { NH_TRY NH_REQUIRE(s && rhs && phi, "null argument"); s->solve(rhs, phi); NH_CATCH }
int32_t nh_tridiag_destroy(nh_tridiag* s) { delete s; return 0; }

}  // extern "C"
