// model.cu — Model<FT>: allocation, field I/O, halo fills, update_state!, time steppers, pressure correction
#include "model.cuh"

#define KLAUNCH(lcp) do { (lcp).n++; } while (0)

template <typename FT>
FT* Model<FT>::alloc_field() {
  bufs.emplace_back(new DevBuf<FT>());
  bufs.back()->alloc((size_t)gh.volume());
  return bufs.back()->p;
}

static void adapt_order(int& kind, int& B, int N, int topo) {
  // adapt_advection_order.jl:18-51,104-110: lower the order when the dimension is shorter than the buffer
  if (topo == NH_FLAT || kind == NH_ADV_NONE) return;
  if (N >= B) return;
  B = N;
  if (kind == NH_ADV_WENO && B < 2) kind = NH_ADV_UPWIND;
}

template <typename FT>
Model<FT>::Model(const nh_grid& grid, const nh_model_desc& d) {
  desc = d;
  device = grid.device;
  float_type = grid.desc.float_type;
  CUDA_CHECK(cudaSetDevice(device));
  ntr = d.n_tracers;
  NH_REQUIRE(ntr >= 0 && ntr <= NH_MAXTR, "too many tracers");
  nf = 3 + ntr;
  for (int e = 0; e < 3; ++e) {
    int kind = d.advection_kind[e], order = d.advection_order[e], B = 1;
    if (kind == NH_ADV_CENTERED) { NH_REQUIRE(order == 2 || order == 4, "Centered order must be 2 or 4"); B = order / 2; }
    else if (kind == NH_ADV_UPWIND) { NH_REQUIRE(order == 1 || order == 3 || order == 5, "UpwindBiased order must be 1, 3 or 5"); B = (order + 1) / 2; }
    else if (kind == NH_ADV_WENO) {
      NH_REQUIRE(order == 1 || order == 3 || order == 5, "WENO order must be 3 or 5");
      B = (order + 1) / 2;
      if (B == 1) kind = NH_ADV_UPWIND;
    } else NH_REQUIRE(kind == NH_ADV_NONE, "unknown advection kind");
    adapt_order(kind, B, grid.desc.N[e], grid.desc.topology[e]);
    adv[e].kind = kind; adv[e].B = B;
  }
  // inflate_grid_halo_size — nonhydrostatic_model.jl:243-257 with inflate_halo_size (automatic_halo_sizing.jl:71-83): every direction
  // is raised to what ITS (adapted) advection scheme needs — required_halo_size_x/y/z of the FluxFormAdvection — and to the 1 a
  // closure and the pressure correction need; never to another direction's buffer
  int H[3];
  for (int e = 0; e < 3; ++e) {
    const int need = adv[e].kind != NH_ADV_NONE ? std::max(1, adv[e].B) : 1;
    H[e] = grid.desc.topology[e] == NH_FLAT ? 0 : std::max(grid.desc.H[e], need);
  }
  gh.build(grid.desc, grid.zfaces, H);
  gdesc = grid.desc; gdesc.z_faces = nullptr;
  for (int e = 0; e < 3; ++e) gdesc.H[e] = gh.g.H[e];
  for (int e = 0; e < 3; ++e)
    if (gh.g.topo[e] == NH_PERIODIC) NH_REQUIRE(gh.g.H[e] <= gh.g.N[e], "halo must be <= size in a Periodic dimension");
  if (d.buoyancy_kind == NH_BUOYANCY_TRACER) NH_REQUIRE(d.buoyancy_tracer[0] >= 0 && d.buoyancy_tracer[0] < ntr, "buoyancy tracer index");
  if (d.buoyancy_kind == NH_BUOYANCY_SEAWATER_LINEAR)
    NH_REQUIRE(d.buoyancy_tracer[0] >= 0 && d.buoyancy_tracer[0] < ntr && d.buoyancy_tracer[1] >= 0 && d.buoyancy_tracer[1] < ntr, "T/S tracer indices");

  nhhost::Tables<FT> T;
  nhhost::make_scheme_tables<FT>(T);
  static_assert(sizeof(T) == sizeof(tab), "table layouts differ");
  std::memcpy(&tab, &T, sizeof(tab));

  for (int f = 0; f < nf; ++f) { F[f] = alloc_field(); Gn[f] = alloc_field(); Gm[f] = alloc_field(); }
  pNHS = alloc_field();
  bool want_phy = d.hydrostatic_pressure_anomaly < 0 ? d.buoyancy_kind != NH_BUOYANCY_NONE : d.hydrostatic_pressure_anomaly != 0;
  if (want_phy) pHY = alloc_field();
  NH_REQUIRE(d.has_closure >= NH_CLOSURE_NONE && d.has_closure <= NH_CLOSURE_AMD, "unknown closure kind");
  vitd = d.has_closure == NH_CLOSURE_SCALAR && d.vertically_implicit != 0;
  NH_REQUIRE(!d.vertically_implicit || d.has_closure == NH_CLOSURE_SCALAR,
             "VerticallyImplicitTimeDiscretization is implemented for ScalarDiffusivity only");
  // implicit_diffusion_solver (vertically_implicit_diffusion_solver.jl:164-169)
  NH_REQUIRE(!vitd || gh.g.topo[2] == NH_BOUNDED,
             "VerticallyImplicitTimeDiscretization can only be specified on grids that are Bounded in the z-direction");
  if (d.has_closure == NH_CLOSURE_AMD) {
    // build_diffusivity_fields (anisotropic_minimum_dissipation.jl:371-385): νₑ and one κₑ per tracer, default conditions
    nue = alloc_field();
    for (int n = 0; n < ntr; ++n) kape[n] = alloc_field();
  } else if (d.has_closure >= NH_CLOSURE_SMAGORINSKY) {
    // build_diffusivity_fields (smagorinsky.jl:123-133); filter width Δᶠ = cbrt(Δx·Δy·Δz), correctly rounded on the host
    nue = alloc_field();
    build_filter_width();
    for (int n = 0; n < ntr; ++n) NH_REQUIRE(d.smagorinsky_Pr[n] != 0, "Smagorinsky: Pr must be non-zero for every tracer");
  }

  solver.lc = &lc;
  solver.stream = stream;               // the complete solver's N³ complex buffers are allocated on first use (ensure_solver)
  red.alloc(1024);

  rk_g1 = (FT)8 / (FT)15; rk_g2 = (FT)5 / (FT)12; rk_g3 = (FT)3 / (FT)4;
  rk_z2 = (FT)-17 / (FT)60; rk_z3 = (FT)-5 / (FT)12;
  chi = (FT)d.ab2_chi;
  clock.time = 0; clock.iteration = 0; clock.stage = 1; clock.last_dt = INFINITY; clock.last_stage_dt = INFINITY;
  fast = ntr == 0 && !gh.z_stretched && d.buoyancy_kind == NH_BUOYANCY_NONE && !d.has_coriolis && !d.has_closure && !pHY;
  for (int e = 0; e < 3; ++e)
    fast = fast && gh.g.topo[e] == NH_PERIODIC && adv[e].kind == NH_ADV_WENO && adv[e].B == 3 && gh.g.N[e] >= 2 * gh.g.H[e] && gh.g.H[e] >= 3;
  fast = fast && gh.volume() < (1ll << 31);        // the fused kernel uses 32-bit element offsets
  if (getenv("NHSTEP_NO_FAST")) fast = false;
  tiled = getenv("NHSTEP_NO_TILED") == nullptr;
  CUDA_CHECK(cudaFuncSetAttribute(tendency_tiled_kernel<FT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(TG_SMEM_ELEMS * sizeof(FT))));
  tiled_fast_ok = gh.volume() < (1ll << 31) && getenv("NHSTEP_NO_FAST_FLUX") == nullptr;
  for (int e = 0; e < 3; ++e) {
    const bool topo_ok = gh.g.topo[e] == NH_PERIODIC || (e == 2 && gh.g.topo[e] == NH_BOUNDED);
    tiled_fast_ok = tiled_fast_ok && topo_ok && adv[e].kind == NH_ADV_WENO && adv[e].B == 3 && gh.g.H[e] >= 3 && gh.g.N[e] >= 2 * gh.g.H[e];
  }
  if (!getenv("NHSTEP_NO_FAST_POISSON")) {
    // fused FFT / DCT / Thomas passes for (Periodic, Periodic, Periodic | Bounded) power-of-two grids, any model physics
    if (!gh.z_stretched && PoissonFast<FT>::supported(gh, true)) pfast.create(gh);
    else if (gh.z_stretched && PoissonFast<FT>::supported_tridiagonal(gh)) pfast.create(gh, true);
  }
  if (fast) {
    for (int a = 0; a < 3; ++a) F2[a] = alloc_field();
    CUDA_CHECK(cudaFuncSetAttribute(tendency_fast_kernel<FT, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(TF_SMEM_ELEMS * sizeof(FT))));
    CUDA_CHECK(cudaFuncSetAttribute(tendency_plane_kernel<FT, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, TPSizes<FT>::total));
    // contracted arithmetic: the plane-marching TMA sweep (0.94 vs 1.11 ms per launch at 256³); exact arithmetic: the tile sweep
    // (its reconstruction with checked divisions needs the tile kernel's 32 warps per SM: 1.75 vs 1.82 ms).  Both produce the
    // same bits in either mode; NHSTEP_TILE_SWEEP=1 / NHSTEP_PLANE_SWEEP=1 force one or the other (A/B, tests).
    plane_sweep = desc.exact_arithmetic == 0;
    if (getenv("NHSTEP_TILE_SWEEP")) plane_sweep = false;
    if (getenv("NHSTEP_PLANE_SWEEP")) plane_sweep = true;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    plane_zseg = getenv("NHSTEP_PLANE_ZSEG") ? std::max(1, atoi(getenv("NHSTEP_PLANE_ZSEG"))) : tp_choose_zseg(gh.g.N[0], gh.g.N[1], gh.g.N[2], sms);
  }
  update_state(0);
}

// Smagorinsky filter width Δᶠ² per level from the CURRENT metrics (dist_configure replaces Δy by the global spacing)
template <typename FT>
void Model<FT>::build_filter_width() {
  std::vector<FT> t(gh.g.N[2]);
  for (int k = 1; k <= gh.g.N[2]; ++k) {
    FT dz = gh.g.topo[2] == NH_FLAT ? FT(1) : gh.dzc_h[k + gh.g.zoff];
    FT d3 = gh.g.dx * gh.g.dy * dz;
    FT df = nhhost::cbrt_correctly_rounded<FT>(d3);
    t[k - 1] = df * df;
  }
  df2.upload(t);
}

// The fused sweep of the fast path.  Tensor maps are cached per field pointer (u, v, w and their ping-pong twins).
template <typename FT>
void Model<FT>::launch_fused_sweep(const FastTendJob<FT>& J, int zseg) {
  const bool exact = desc.exact_arithmetic != 0;
  if (!plane_sweep) { launch_fused_tendency<FT>(J, stream, exact); return; }
  PlaneJob<FT> P;
  P.J = J;
  P.zseg = zseg > 0 ? std::min(zseg, plane_zseg) : plane_zseg;
  for (int a = 0; a < 3; ++a) {
    auto it = tmaps.find((const void*)J.U[a]);
    if (it == tmaps.end()) {
      CUtensorMap m;
      tp_encode_field_map<FT>(&m, J.U[a], J.g);
      it = tmaps.emplace((const void*)J.U[a], m).first;
    }
    P.tm[a] = it->second;
  }
  launch_plane_tendency<FT>(P, stream, exact);
}

// the complete (any topology, any size) Poisson solver: only models that the fused passes do not cover ever need it
template <typename FT>
void Model<FT>::ensure_solver() {
  if (solver_ready) return;
  solver.create(gh, gh.z_stretched ? NH_SOLVER_FOURIER_TRIDIAGONAL : NH_SOLVER_FFT);
  solver_ready = true;
}

template <typename FT>
void Model<FT>::ensure_tendencies() {
  if (!gn_stale) return;
  gn_stale = false;
  compute_tendencies();
}

template <typename FT>
void Model<FT>::ensure_p_halo() {
  if (!p_halo_stale) return;
  p_halo_stale = false;
  fill_halo_regions(1u << NH_FIELD_PNHS, 1);
}

// One fused stage: Gⁿ = G(U), U* = U + Δt(…) into the ping-pong buffer, then the pressure correction.
template <typename FT>
void Model<FT>::fast_stage(int mode, FT dt, FT c1, FT c2, double stage_dt) {
  FastTendJob<FT> J;
  J.g = gh.g; J.tab = tab; J.dt = dt; J.gamma = c1; J.zeta = c2; J.mode = mode;
  for (int a = 0; a < 3; ++a) { J.U[a] = F[a]; J.Unew[a] = F2[a]; J.Gn[a] = Gn[a]; J.Gm[a] = Gm[a]; }
  launch_fused_sweep(J);
  KLAUNCH(lc);
  mark(0);
  for (int a = 0; a < 3; ++a) std::swap(F[a], F2[a]);
  fast_pressure(stage_dt);
}

template <typename FT>
void Model<FT>::fast_pressure(double stage_dt) {
  const GridDev<FT>& g = gh.g;
  if (pfast.ok) {
    // fused solve: source term with periodic wrap inside the first FFT pass, no halo fill, 5 launches
    dim3 blk(64, 4, 1);
    pfast.solve(F[0], F[1], F[2], pNHS, stage_dt, stream, lc);
    mark(1);
    launch_pressure_correct_halo<FT>(g, F[0], F[1], F[2], pNHS, stage_dt, 1, stream);
    KLAUNCH(lc);
    mark(2);
    p_halo_stale = true;
    CUDA_CHECK(cudaGetLastError());
    return;
  }
  FastHaloJob<FT> H; H.g = g; H.nf = 3; for (int a = 0; a < 3; ++a) H.p[a] = F[a];
  dim3 hg(1, g.N[1] + 2 * g.H[1], g.N[2] + 2 * g.H[2]);
  periodic_halo_all_kernel<FT><<<hg, 128, 0, stream>>>(H);
  KLAUNCH(lc);
  dim3 blk(64, 4, 1);
  ensure_solver();
  source_term_kernel<FT, CT><<<grid3(blk), blk, 0, stream>>>(g, F[0], F[1], F[2], stage_dt, 0, solver.A.p);
  KLAUNCH(lc);
  CT* sol = solver.solve(0.0);
  copy_real_kernel<FT, CT><<<grid3(blk), blk, 0, stream>>>(g, sol, pNHS);
  KLAUNCH(lc);
  mark(1);
  launch_pressure_correct_halo<FT>(g, F[0], F[1], F[2], pNHS, stage_dt, 1, stream);
  KLAUNCH(lc);
  mark(2);
  p_halo_stale = true;
  CUDA_CHECK(cudaGetLastError());
}

template <typename FT>
int Model<FT>::field_face_dim(int field) const {
  int f = field;
  if (field >= NH_FIELD_GM0) f = field - NH_FIELD_GM0;
  else if (field >= NH_FIELD_GN0) f = field - NH_FIELD_GN0;
  return (f >= 0 && f < 3) ? f : -1;
}

template <typename FT>
FT* Model<FT>::field_ptr(int field) {
  if (field >= 0 && field < nf) return F[field];
  if (field == NH_FIELD_PNHS) return pNHS;
  if (field == NH_FIELD_PHY) { NH_REQUIRE(pHY, "model has no hydrostatic pressure anomaly"); return pHY; }
  if (field == NH_FIELD_NUE) { NH_REQUIRE(nue, "model has no eddy viscosity (closure is not a Smagorinsky or AMD)"); return nue; }
  if (field >= NH_FIELD_KAPPAE0 && field < NH_FIELD_KAPPAE0 + NH_MAXTR) {
    NH_REQUIRE(field - NH_FIELD_KAPPAE0 < ntr && kape[field - NH_FIELD_KAPPAE0], "model has no such eddy diffusivity field");
    return kape[field - NH_FIELD_KAPPAE0];
  }
  if (field >= NH_FIELD_GN0 && field < NH_FIELD_GN0 + nf) return Gn[field - NH_FIELD_GN0];
  if (field >= NH_FIELD_GM0 && field < NH_FIELD_GM0 + nf) return Gm[field - NH_FIELD_GM0];
  throw NhError("unknown field id");
}

template <typename FT>
void Model<FT>::field_extent(int field, int n[3]) const {
  int fd = field_face_dim(field);
  for (int e = 0; e < 3; ++e) n[e] = gh.g.N[e] + ((fd == e && gh.g.topo[e] == NH_BOUNDED) ? 1 : 0);
}

template <typename FT>
void Model<FT>::field_size(int field, int region, int64_t size[3]) {
  (void)field_ptr(field);
  int n[3]; field_extent(field, n);
  for (int e = 0; e < 3; ++e) size[e] = n[e] + (region == NH_PARENT ? 2 * gh.g.H[e] : 0);
}

template <typename FT>
void Model<FT>::set_field(int field, int region, const void* data, int64_t count, int where) {
  FT* p = field_ptr(field);
  int64_t sz[3]; field_size(field, region, sz);
  NH_REQUIRE(count == sz[0] * sz[1] * sz[2], "nh_set_field: element count does not match the field's size");
  const FT* src = (const FT*)data;
  if (where == NH_HOST) {
    if (stage.n < (size_t)count) stage.alloc(count, false);
    CUDA_CHECK(cudaMemcpyAsync(stage.p, data, count * sizeof(FT), cudaMemcpyHostToDevice, stream));
    src = stage.p;
  }
  int lo[3]; for (int e = 0; e < 3; ++e) lo[e] = region == NH_PARENT ? 1 - gh.g.H[e] : 1;
  layout_copy_kernel<FT><<<ceil_div(count, 256), 256, 0, stream>>>(gh.g, p, const_cast<FT*>(src), lo[0], lo[1], lo[2], (int)sz[0], (int)sz[1], (int)sz[2], 1);
  KLAUNCH(lc);
  CUDA_CHECK(cudaGetLastError());
  if (where == NH_HOST) CUDA_CHECK(cudaStreamSynchronize(stream));
}

template <typename FT>
void Model<FT>::get_field(int field, int region, void* data, int64_t count, int where) {
  if (field >= NH_FIELD_GN0 && field < NH_FIELD_GM0) ensure_tendencies();
  if (field == NH_FIELD_PNHS && region == NH_PARENT) ensure_p_halo();
  FT* p = field_ptr(field);
  int64_t sz[3]; field_size(field, region, sz);
  NH_REQUIRE(count == sz[0] * sz[1] * sz[2], "nh_get_field: element count does not match the field's size");
  FT* dst = (FT*)data;
  if (where == NH_HOST) { if (stage.n < (size_t)count) stage.alloc(count, false); dst = stage.p; }
  int lo[3]; for (int e = 0; e < 3; ++e) lo[e] = region == NH_PARENT ? 1 - gh.g.H[e] : 1;
  layout_copy_kernel<FT><<<ceil_div(count, 256), 256, 0, stream>>>(gh.g, p, dst, lo[0], lo[1], lo[2], (int)sz[0], (int)sz[1], (int)sz[2], 0);
  KLAUNCH(lc);
  CUDA_CHECK(cudaGetLastError());
  if (where == NH_HOST) {
    CUDA_CHECK(cudaMemcpyAsync(data, stage.p, count * sizeof(FT), cudaMemcpyDeviceToHost, stream));
    CUDA_CHECK(cudaStreamSynchronize(stream));
  }
}

template <typename FT>
void Model<FT>::field_device_ptr(int field, void** ptr, int64_t* off, int64_t strides[3]) {
  *ptr = field_ptr(field);
  *off = gh.g.idx(1, 1, 1);
  strides[0] = 1; strides[1] = gh.g.sy; strides[2] = gh.g.sz;
}

// FieldBoundaryConditions with number conditions (include/nhstep.h: nh_model_set_boundary_condition)
template <typename FT>
void Model<FT>::set_boundary_condition(int field, int dim, int side, int kind, double value) {
  NH_REQUIRE(field >= 0 && field < nf, "boundary conditions can be set on prognostic fields (u, v, w, tracers) only");
  NH_REQUIRE(dim >= 0 && dim < 3 && (side == 0 || side == 1), "bad dim / side");
  NH_REQUIRE(kind >= NH_BC_DEFAULT && kind <= NH_BC_GRADIENT, "unknown boundary condition kind");
  NH_REQUIRE(gh.g.topo[dim] == NH_BOUNDED, "Value / Gradient / Flux boundary conditions need a Bounded direction");
  NH_REQUIRE(field_face_dim(field) != dim, "a wall-normal velocity component keeps the impenetrable (Open) condition");
  bc_kind[field][dim][side] = kind;
  bc_val[field][dim][side] = kind == NH_BC_DEFAULT ? FT(0) : (FT)value;
  bc_arr[field][dim][side].alloc(0);
  has_flux_bcs = false;
  for (int f = 0; f < nf; ++f) for (int d = 0; d < 3; ++d) for (int s2 = 0; s2 < 2; ++s2)
    if (bc_kind[f][d][s2] == NH_BC_FLUX) has_flux_bcs = true;
}

template <typename FT>
void Model<FT>::set_boundary_condition_array(int field, int dim, int side, int kind, const void* data, long long count, int where) {
  NH_REQUIRE(kind >= NH_BC_FLUX && kind <= NH_BC_GRADIENT, "an array condition is a Flux, Value or Gradient condition");
  NH_REQUIRE(data != nullptr, "null boundary condition array");
  NH_REQUIRE(where == NH_HOST || where == NH_DEVICE, "bad `where`");
  set_boundary_condition(field, dim, side, kind, 0.0);
  const int e1 = dim == 0 ? 1 : 0, e2 = dim == 2 ? 1 : 2;
  NH_REQUIRE(count == (long long)gh.g.N[e1] * gh.g.N[e2], "boundary condition array: count must be the wall's transverse grid size");
  DevBuf<FT>& b = bc_arr[field][dim][side];
  b.alloc((size_t)count, false);
  CUDA_CHECK(cudaMemcpyAsync(b.p, data, count * sizeof(FT), where == NH_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, stream));
  CUDA_CHECK(cudaStreamSynchronize(stream));
}

// WENO(grid) (weno_reconstruction.jl:98-124): the per-index reconstruction coefficients of the stretched z direction, generated on
// the host as compute_reconstruction_coefficients does (include/nhstep.h: nh_model_set_weno_z_coefficients)
template <typename FT>
void Model<FT>::set_weno_z_coefficients(int buffer, int location, const void* data, long long count, int where) {
  NH_REQUIRE(buffer == 2 || buffer == 3, "WENO(grid): buffer must be 2 (order 3) or 3 (order 5)");
  NH_REQUIRE(location == 0 || location == 1, "WENO(grid): location must be 0 (Face targets) or 1 (Center targets)");
  NH_REQUIRE(adv[2].kind == NH_ADV_WENO && adv[2].B >= buffer, "WENO(grid): the model's z advection is not a WENO scheme of that order");
  NH_REQUIRE(data != nullptr && (where == NH_HOST || where == NH_DEVICE), "WENO(grid): bad data / where");
  NH_REQUIRE(count == (long long)(gh.g.N[2] + 2) * (buffer + 1) * buffer, "WENO(grid): count must be (Nz + 2) × (buffer + 1) × buffer");
  DevBuf<FT>& b = weno_z[location][buffer - 2];
  b.alloc((size_t)count, false);
  CUDA_CHECK(cudaMemcpyAsync(b.p, data, count * sizeof(FT), where == NH_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, stream));
  CUDA_CHECK(cudaStreamSynchronize(stream));
  weno_stretched = true;             // tendencies through the per-cell kernel instantiated with the tables (tendency_generic_kernel<FT, true>)
  gn_stale = true;
}

// fill_halo_regions! — open fill, then the three directions sorted so that Periodic runs after the others
// (fill_halo_regions.jl:50-67,121-123,159-196; field_tuples.jl:56-101)
template <typename FT>
void Model<FT>::fill_halo_regions(uint32_t mask, int fill_normal) {
  const GridDev<FT>& g = gh.g;
  std::vector<int> ids;
  for (int f = 0; f < 32; ++f) if (mask & (1u << f)) { (void)field_ptr(f); ids.push_back(f); }
  if (ids.empty()) return;
  NH_REQUIRE((int)ids.size() <= NH_MAXF, "too many fields in one halo fill");
  auto launch = [&](HaloJob<FT>& J) {
    const int d = J.d, e1 = d == 0 ? 1 : 0, e2 = d == 2 ? 1 : 2;
    long long n = (long long)(J.hi[e1] - J.lo[e1] + 1) * (J.hi[e2] - J.lo[e2] + 1);
    if (n <= 0) return;
    dim3 grd(ceil_div(n, 128), J.nf);
    halo_fill_kernel<FT><<<grd, 128, 0, stream>>>(J);
    KLAUNCH(lc);
  };
  if (fill_normal) {
    for (int d = 0; d < 3; ++d) {
      if (g.topo[d] != NH_BOUNDED) continue;
      HaloJob<FT> J; J.g = g; J.d = d; J.nf = 0;
      for (int id : ids) if (id < 3 && id == d) { J.p[J.nf] = F[id]; J.mode[J.nf] = 3; J.nf++; }
      if (!J.nf) continue;
      for (int e = 0; e < 3; ++e) { J.lo[e] = 1; J.hi[e] = g.N[e]; }
      launch(J);
    }
  }
  // order: insertion sort with lt(a,b) = !(a periodic && b not periodic)
  bool per[3]; for (int d = 0; d < 3; ++d) per[d] = g.topo[d] == NH_PERIODIC;
  auto lt = [&](int a, int b) { return !(per[a] && !per[b]); };
  int order[3] = {0, 1, 2};
  for (int i = 1; i < 3; ++i) { int x = order[i], j = i; while (j > 0 && lt(x, order[j - 1])) { order[j] = order[j - 1]; --j; } order[j] = x; }
  for (int q = 0; q < 3; ++q) {
    int d = order[q];
    if (g.topo[d] == NH_FLAT || g.H[d] == 0) continue;
    HaloJob<FT> J; J.g = g; J.d = d; J.nf = 0;
    if (per[d]) {
      int minext[3] = {1 << 30, 1 << 30, 1 << 30};
      for (int id : ids) { int n[3]; field_extent(id, n); for (int e = 0; e < 3; ++e) minext[e] = std::min(minext[e], n[e] + 2 * g.H[e]); }
      for (int e = 0; e < 3; ++e) { J.lo[e] = 1 - g.H[e]; J.hi[e] = minext[e] - g.H[e]; }
      const bool pull = dist.on && dist.generic && d == 1;
      if (pull) NH_REQUIRE(dist.p2p, "a distributed complete-path model needs nh_dist_ipc_import before its first halo fill");
      for (int id : ids) {
        J.p[J.nf] = field_ptr(id); J.mode[J.nf] = 1;
        J.pull_prev[J.nf] = pull ? dist.peer_prev[id] : nullptr;
        J.pull_next[J.nf] = pull ? dist.peer_next[id] : nullptr;
        if (pull) NH_REQUIRE(dist.peer_prev[id] && dist.peer_next[id], "field has no peer pointers");
        J.nf++;
      }
      if (pull) {
        // the neighbours' rows (and their x / z images, filled before y in the reference's order) must be complete before
        // they are read, and read before the neighbours write them again
        dist_barrier();
        launch(J);
        dist_barrier();
        continue;
      }
    } else {
      for (int e = 0; e < 3; ++e) { J.lo[e] = 1; J.hi[e] = g.N[e]; }
      for (int id : ids) {
        if (field_face_dim(id) == d) continue;        // Face field on a wall: open / no boundary condition
        J.p[J.nf] = field_ptr(id); J.mode[J.nf] = 2;
        for (int side = 0; side < 2; ++side) {
          const bool prog = id >= 0 && id < nf;
          J.side_kind[J.nf][side] = prog ? bc_kind[id][d][side] : NH_BC_DEFAULT;
          J.side_val[J.nf][side] = prog ? bc_val[id][d][side] : FT(0);
          J.side_arr[J.nf][side] = prog ? bc_arr[id][d][side].p : nullptr;
        }
        J.nf++;
      }
      if (!J.nf) continue;
    }
    launch(J);
  }
  CUDA_CHECK(cudaGetLastError());
}

template <typename FT>
TendParams<FT> Model<FT>::tend_params() const {
  TendParams<FT> P;
  std::memset(&P, 0, sizeof(P));
  P.g = gh.g;
  for (int a = 0; a < 3; ++a) { P.U[a] = F[a]; P.GU[a] = Gn[a]; P.adv[a] = adv[a]; }
  for (int t = 0; t < ntr; ++t) { P.c[t] = F[3 + t]; P.Gc[t] = Gn[3 + t]; P.kappa[t] = (FT)desc.kappa[t]; }
  P.pHY = pHY;
  P.ntr = ntr;
  P.tab = tab;
  P.buoy_kind = desc.buoyancy_kind;
  P.iT = desc.buoyancy_tracer[0]; P.iS = desc.buoyancy_tracer[1];
  P.grav = (FT)desc.gravitational_acceleration; P.alpha = (FT)desc.thermal_expansion; P.beta = (FT)desc.haline_contraction;
  P.w_buoyancy = (desc.buoyancy_kind != NH_BUOYANCY_NONE && !pHY) ? 1 : 0;
  P.has_coriolis = desc.has_coriolis; P.fcor = (FT)desc.coriolis_f;
  P.has_closure = desc.has_closure; P.nu = (FT)desc.nu;
  P.vitd = vitd ? 1 : 0;
  P.nue = nue;
  for (int l = 0; l < 2; ++l) for (int q = 0; q < 2; ++q) P.wz[l][q] = weno_z[l][q].p;
  const bool amd = desc.has_closure == NH_CLOSURE_AMD;
  for (int t = 0; t < ntr; ++t) { P.Pr[t] = amd ? FT(1) : (FT)desc.smagorinsky_Pr[t]; P.kape[t] = amd ? kape[t] : nue; }
  return P;
}

// compute_diffusivities! (smagorinsky.jl:107-119) over :xyz; the halo of νₑ is filled afterwards by update_state
template <typename FT>
void Model<FT>::compute_diffusivities() {
  if (!nue) return;
  if (desc.has_closure == NH_CLOSURE_AMD) {
    // compute_diffusivities!(…, ::AnisotropicMinimumDissipation, …): νₑ, then κₑ of each tracer
    TendParams<FT> P = tend_params();
    dim3 blk(32, 4, 1);
    amd_viscosity_kernel<FT><<<grid3(blk), blk, 0, stream>>>(P, nue, (FT)desc.amd_Cnu);
    KLAUNCH(lc);
    for (int n = 0; n < ntr; ++n) {
      amd_diffusivity_kernel<FT><<<grid3(blk), blk, 0, stream>>>(P, F[3 + n], kape[n], (FT)desc.amd_Ckappa[n]);
      KLAUNCH(lc);
    }
    return;
  }
  SmagJob<FT> J;
  J.P = tend_params(); J.nue = nue; J.df2 = df2.p;
  J.lilly = desc.has_closure == NH_CLOSURE_SMAGORINSKY_LILLY;
  J.C = (FT)desc.smagorinsky_C; J.Cb = (FT)desc.smagorinsky_Cb; J.C64 = desc.smagorinsky_C;
  dim3 blk(32, 4, 1);
  smagorinsky_kernel<FT><<<grid3(blk), blk, 0, stream>>>(J);
  KLAUNCH(lc);
}

template <typename FT>
void Model<FT>::update_hydrostatic_pressure() {
  if (!pHY || desc.buoyancy_kind == NH_BUOYANCY_NONE || gh.g.topo[2] == NH_FLAT) return;
  TendParams<FT> P = tend_params();
  dim3 blk(32, 4), grd(ceil_div(gh.g.N[0] + 2, 32), ceil_div(gh.g.N[1] + 2, 4));
  hydrostatic_pressure_kernel<FT><<<grd, blk, 0, stream>>>(P, pHY);
  KLAUNCH(lc);
}

template <typename FT>
void Model<FT>::compute_tendencies() {
  gn_stale = false;
  TendParams<FT> P = tend_params();
  // tiled kernel (each advective face flux once, staged in shared memory); NHSTEP_NO_TILED selects the per-cell kernel that
  // evaluates every face from both neighbours like the reference (bit-identical results, kept as the cross-check)
  dim3 blk(32, 4, 1);
  if (tiled && !weno_stretched) {
    // one launch for all prognostic fields; interior tiles of WENO5 models on (Periodic, Periodic, Periodic | Bounded) regular
    // grids take the fast flux stage (tendency_generic.cuh)
    dim3 tb(TG_TX, TG_TY, 1), tg(ceil_div(gh.g.N[0], TG_TX), ceil_div(gh.g.N[1], TG_TY), ceil_div(gh.g.N[2], TG_TZ));
    if (desc.exact_arithmetic == 0 && tiled_fast_ok) launch_tiled_tendency_fma<FT>(P, tg, 1, stream);
    else tendency_tiled_kernel<FT><<<tg, tb, TG_SMEM_ELEMS * sizeof(FT), stream>>>(P, -1, tiled_fast_ok ? 1 : 0);
    KLAUNCH(lc);
  } else {
    for (int what = 0; what < nf; ++what) {
      if (weno_stretched) tendency_generic_kernel<FT, true><<<grid3(blk), blk, 0, stream>>>(P, what);      // WENO(grid)
      else tendency_generic_kernel<FT><<<grid3(blk), blk, 0, stream>>>(P, what);
      KLAUNCH(lc);
    }
  }
  // compute_boundary_tendency_contributions! (compute_nonhydrostatic_tendencies.jl:202-207): x, then y, then z walls
  if (has_flux_bcs) {
    for (int d = 0; d < 3; ++d) for (int f = 0; f < nf; ++f) {
      if (bc_kind[f][d][0] != NH_BC_FLUX && bc_kind[f][d][1] != NH_BC_FLUX) continue;
      FluxBcJob<FT> J; J.g = gh.g; J.G = Gn[f]; J.d = d; J.zloc_face = f == 2;
      for (int side = 0; side < 2; ++side) { J.has[side] = bc_kind[f][d][side] == NH_BC_FLUX; J.q[side] = bc_val[f][d][side]; J.qarr[side] = bc_arr[f][d][side].p; }
      const int e1 = d == 0 ? 1 : 0, e2 = d == 2 ? 1 : 2;
      long long n = (long long)gh.g.N[e1] * gh.g.N[e2];
      flux_bc_kernel<FT><<<ceil_div(n, 128), 128, 0, stream>>>(J);
      KLAUNCH(lc);
    }
  }
  CUDA_CHECK(cudaGetLastError());
}

template <typename FT>
void Model<FT>::update_state(int compute) {
  if (compute) gn_stale = false;
  uint32_t mask = 0; for (int f = 0; f < nf; ++f) mask |= 1u << f;
  fill_halo_regions(mask, 0);
  compute_diffusivities();                       // compute_auxiliaries!: diffusivities, then hydrostatic pressure
  update_hydrostatic_pressure();
  if (nue) {
    uint32_t dmask = 1u << NH_FIELD_NUE;
    if (desc.has_closure == NH_CLOSURE_AMD) for (int n = 0; n < ntr; ++n) dmask |= 1u << (NH_FIELD_KAPPAE0 + n);
    fill_halo_regions(dmask, 1);
  }
  if (compute) compute_tendencies();
}

template <typename FT>
void Model<FT>::substep_launch(int mode, FT dt, FT c1, FT c2) {
  ensure_tendencies();
  StepJob<FT> J;
  J.g = gh.g; J.nf = nf; J.dt = dt; J.c1 = c1; J.c2 = c2; J.mode = mode;
  for (int f = 0; f < nf; ++f) { J.U[f] = F[f]; J.Gn[f] = Gn[f]; J.Gm[f] = Gm[f]; J.face_dim[f] = f < 3 ? f : -1; }
  dim3 blk(64, 4, 1);
  substep_kernel<FT><<<grid3(blk), blk, 0, stream>>>(J);
  KLAUNCH(lc);
  CUDA_CHECK(cudaGetLastError());
}

// implicit_step! of every prognostic field after its explicit substep (runge_kutta_3.jl:157-165,
// quasi_adams_bashforth_2.jl:134-142).  ab2 = the Δt was converted to the grid's type before the coefficients were formed.
template <typename FT>
template <typename CT2>
void Model<FT>::implicit_launch(double stage_dt) {
  const int L = gh.g.N[2] + 2;
  dim3 blk(64, 1, 1);
  dim3 grd((gh.g.N[0] + blk.x - 1) / blk.x, gh.g.N[1], 1);
  for (int f = 0; f < nf; ++f) {
    const CT2 dt = (CT2)stage_dt, kap = (CT2)(f < 3 ? (FT)desc.nu : (FT)desc.kappa[f - 3]);
    const int w_field = f == 2, ct = (int)sizeof(CT2);
    // coefficient tables are cached per (Δt, κ, location, coefficient type): with a constant Δt every stage finds its own
    ImplicitTable* T = nullptr;
    for (auto& e : vitd_tabs) if (e.dt == (double)dt && e.kappa == (double)kap && e.w_field == w_field && e.ct == ct) T = &e;
    if (!T) {
      if (vitd_tabs.size() < 24) { vitd_tabs.emplace_back(); T = &vitd_tabs.back(); T->tab.alloc(4 * (size_t)L); }
      else { T = &vitd_tabs[vitd_next]; vitd_next = (vitd_next + 1) % vitd_tabs.size(); }
      T->dt = (double)dt; T->kappa = (double)kap; T->w_field = w_field; T->ct = ct;
      ImplicitCoeffJob<FT, CT2> C;
      C.g = gh.g; C.tab = T->tab.p; C.dt = dt; C.kappa = kap; C.w_field = w_field;
      implicit_coeff_kernel<FT, CT2><<<1, 32, 0, stream>>>(C);
      KLAUNCH(lc);
    }
    ImplicitJob<FT, CT2> J;
    J.g = gh.g; J.phi = F[f]; J.tab = T->tab.p;
    implicit_step_kernel<FT, CT2><<<grd, blk, 0, stream>>>(J);
    KLAUNCH(lc);
  }
  CUDA_CHECK(cudaGetLastError());
}

template <typename FT>
void Model<FT>::implicit_steps(double stage_dt, bool ab2) {
  if (!vitd) return;
  if (ab2) implicit_launch<FT>((double)(FT)stage_dt);
  else implicit_launch<double>(stage_dt);
}

template <typename FT>
void Model<FT>::rk3_substep(double dt, double gamma, double zeta, int has_zeta) {
  substep_launch(has_zeta ? 1 : 0, (FT)dt, (FT)gamma, (FT)zeta);
  implicit_steps(has_zeta ? (double)(FT)((FT)gamma + (FT)zeta) * dt : (double)(FT)gamma * dt, false);
}

template <typename FT>
void Model<FT>::ab2_step(double dt, double chi_) {
  FT c = (FT)chi_;
  substep_launch(c != FT(-0.5) ? 2 : 3, (FT)dt, c, FT(0));
  implicit_steps(dt, true);
}

template <typename FT>
void Model<FT>::cache_previous_tendencies() { substep_launch(4, FT(0), FT(0), FT(0)); }

// calculate_pressure_correction! — pressure_correction.jl:8-20
template <typename FT>
void Model<FT>::calculate_pressure_correction(double dt) {
  fill_halo_regions(0x7u, 1);
  dim3 blk(64, 4, 1);
  if (dist.on && dist.generic) {
    // slab run: the same real source term, then the distributed solve (x-FFT → transpose → y-FFT, division or Thomas sweep,
    // y-IFFT → transpose back → x-IFFT); the transposes are peer stores, separated by stream-ordered barriers
    NH_REQUIRE(dist.p2p, "a distributed complete-path model needs nh_dist_ipc_import first");
    const long long n = (long long)gh.g.N[0] * gh.g.N[1] * gh.g.N[2];
    if (prhs.n < (size_t)n) prhs.alloc(n, false);
    source_term_real_kernel<FT><<<grid3(blk), blk, 0, stream>>>(gh.g, F[0], F[1], F[2], dt, dist.tri ? 1 : 0, prhs.p);
    KLAUNCH(lc);
    dist.generic_forward(prhs.p, stream, lc);
    dist_barrier();
    dist.generic_solve_y(stream, lc);
    dist_barrier();
    dist.generic_backward(pNHS, stream, lc);
    fill_halo_regions(1u << NH_FIELD_PNHS, 1);
    p_halo_stale = false;
    CUDA_CHECK(cudaGetLastError());
    return;
  }
  if (pfast.ok) {
    // fused passes: real source term (same expression as below) → x/y FFT + z FFT | DCT | Thomas → interior of pNHS
    const long long n = (long long)gh.g.N[0] * gh.g.N[1] * gh.g.N[2];
    if (prhs.n < (size_t)n) prhs.alloc(n, false);
    source_term_real_kernel<FT><<<grid3(blk), blk, 0, stream>>>(gh.g, F[0], F[1], F[2], dt, pfast.tri ? 1 : 0, prhs.p);
    KLAUNCH(lc);
    if (pfast.tri) pfast.solve_tridiagonal(prhs.p, pNHS, stream, lc, false, &gh.g);
    else pfast.solve_rhs(prhs.p, pNHS, 0.0, stream, lc, &gh.g);
    fill_halo_regions(1u << NH_FIELD_PNHS, 1);
    p_halo_stale = false;
    CUDA_CHECK(cudaGetLastError());
    return;
  }
  ensure_solver();
  CT* rhs = solver.A.p;
  source_term_kernel<FT, CT><<<grid3(blk), blk, 0, stream>>>(gh.g, F[0], F[1], F[2], dt,
                                                            solver.kind == NH_SOLVER_FOURIER_TRIDIAGONAL ? 1 : 0, rhs);
  KLAUNCH(lc);
  CT* sol = solver.solve(0.0);
  copy_real_kernel<FT, CT><<<grid3(blk), blk, 0, stream>>>(gh.g, sol, pNHS);
  KLAUNCH(lc);
  fill_halo_regions(1u << NH_FIELD_PNHS, 1);
  CUDA_CHECK(cudaGetLastError());
}

template <typename FT>
void Model<FT>::pressure_correct_velocities(double dt) {
  dim3 blk(64, 4, 1);
  pressure_correct_kernel<FT><<<grid3(blk), blk, 0, stream>>>(gh.g, F[0], F[1], F[2], pNHS, dt);
  KLAUNCH(lc);
  CUDA_CHECK(cudaGetLastError());
}

// time_step! — runge_kutta_3.jl:77-151, quasi_adams_bashforth_2.jl:74-115.  The G⁻ ← Gⁿ copy of
// store_tendencies.jl is replaced by a pointer swap (Gⁿ is fully rewritten by the next compute_tendencies!).
template <typename FT>
void Model<FT>::time_step(double dt, int euler) {
  if (dist.on && !dist.generic)
    throw NhError("a distributed model on the fused path is stepped through nh_dist_* (see include/nhstep.h), not nh_time_step");
  if (fast) {
    // same sequence as below with compute_tendencies! + the next substep fused and the halo fills folded into the
    // pressure-correction kernel; Gⁿ of the final state is refreshed lazily (ensure_tendencies).
    // update_state! at iteration 0 (runge_kutta_3.jl:81, quasi_adams_bashforth_2.jl:80): a field set without the set!
    // tail (Field.set → nh_set_field) has stale halos, and the fused sweep reads filled ones
    if (clock.iteration == 0) update_state(0);
    if (desc.timestepper == NH_RK3) {
      double first = (double)rk_g1 * dt, second = (double)(FT)(rk_g2 + rk_z2) * dt, third = (double)(FT)(rk_g3 + rk_z3) * dt;
      double t_next = clock.time + dt;
      auto swapG = [&]() { for (int f = 0; f < nf; ++f) std::swap(Gn[f], Gm[f]); };
      fast_stage(0, (FT)dt, rk_g1, FT(0), first);
      clock.time += first; clock.stage += 1; clock.last_stage_dt = first;
      swapG();
      fast_stage(1, (FT)dt, rk_g2, rk_z2, second);
      clock.time += second; clock.stage += 1; clock.last_stage_dt = second;
      swapG();
      fast_stage(1, (FT)dt, rk_g3, rk_z3, third);
      double corrected = t_next - clock.time;
      clock.time += third; clock.iteration += 1; clock.stage = 1; clock.last_stage_dt = corrected; clock.last_dt = dt;
    } else {
      bool eul = euler || (dt != clock.last_dt);
      fast_stage(eul ? 3 : 2, (FT)dt, eul ? FT(-0.5) : chi, FT(0), dt);
      clock.time += dt; clock.iteration += 1; clock.stage = 1; clock.last_dt = dt; clock.last_stage_dt = dt;
      for (int f = 0; f < nf; ++f) std::swap(Gn[f], Gm[f]);
    }
    gn_stale = true;
    return;
  }
  if (clock.iteration == 0) update_state(1);
  auto swapG = [&]() { for (int f = 0; f < nf; ++f) std::swap(Gn[f], Gm[f]); };
  if (desc.timestepper == NH_RK3) {
    double first = (double)rk_g1 * dt;
    double second = (double)(FT)(rk_g2 + rk_z2) * dt;
    double third = (double)(FT)(rk_g3 + rk_z3) * dt;
    double t_next = clock.time + dt;
    substep_launch(0, (FT)dt, rk_g1, FT(0)); implicit_steps(first, false); mark(0);
    clock.time += first; clock.stage += 1; clock.last_stage_dt = first;
    calculate_pressure_correction(first); mark(1);
    pressure_correct_velocities(first); mark(2);
    swapG();
    update_state(1); mark(0);
    substep_launch(1, (FT)dt, rk_g2, rk_z2); implicit_steps(second, false); mark(0);
    clock.time += second; clock.stage += 1; clock.last_stage_dt = second;
    calculate_pressure_correction(second); mark(1);
    pressure_correct_velocities(second); mark(2);
    swapG();
    update_state(1); mark(0);
    substep_launch(1, (FT)dt, rk_g3, rk_z3); implicit_steps(third, false); mark(0);
    double corrected = t_next - clock.time;
    clock.time += third; clock.iteration += 1; clock.stage = 1;
    clock.last_stage_dt = corrected; clock.last_dt = dt;
    calculate_pressure_correction(third); mark(1);
    pressure_correct_velocities(third); mark(2);
    update_state(1); mark(0);
  } else {
    bool eul = euler || (dt != clock.last_dt);
    FT c = eul ? FT(-0.5) : chi;
    substep_launch(eul ? 3 : 2, (FT)dt, c, FT(0)); implicit_steps(dt, true); mark(0);
    clock.time += dt; clock.iteration += 1; clock.stage = 1; clock.last_dt = dt; clock.last_stage_dt = dt;
    calculate_pressure_correction(dt); mark(1);
    pressure_correct_velocities(dt); mark(2);
    swapG();
    update_state(1); mark(0);
  }
}

// tail of set!(model; …) — set_nonhydrostatic_model.jl:44-58
template <typename FT>
void Model<FT>::set_state_finalize(int enforce) {
  update_state(0);
  if (enforce) {
    calculate_pressure_correction(1.0);
    pressure_correct_velocities(1.0);
    update_state(0);
  }
}

template <typename FT>
double Model<FT>::max_abs_divergence() {
  const int nb = 512;
  if (red.n < (size_t)nb) red.alloc(nb);
  max_div_kernel<FT><<<nb, 256, 0, stream>>>(gh.g, F[0], F[1], F[2], red.p);
  KLAUNCH(lc);
  std::vector<double> h(nb);
  CUDA_CHECK(cudaMemcpyAsync(h.data(), red.p, nb * sizeof(double), cudaMemcpyDeviceToHost, stream));
  CUDA_CHECK(cudaStreamSynchronize(stream));
  double m = 0;
  for (double v : h) if (v > m || v != v) m = v;
  return m;
}

template struct Model<float>;
template struct Model<double>;
