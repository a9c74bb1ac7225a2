// dist.cu — Model<FT> pieces of the y-slab multi-GPU time step (one process per GPU).
//   src/DistributedComputations/halo_communication.jl:114-366, distributed_fft_based_poisson_solver.jl:141-178,
//   src/Models/NonhydrostaticModels/solve_for_pressure.jl:12-17, runge_kutta_3.jl:77-151
// Each call enqueues kernels on the model's stream and returns; the host driver places the NCCL messages between them:
//   tendency_substep → [v: 1 row] → poisson(0) → [all-to-all A→B] → poisson(1) → [all-to-all B→A] → poisson(2)
//                    → [p: 1 row] → pressure_correct → [u,v,w: H rows]
#include "model.cuh"

template <typename FT>
void Model<FT>::dist_configure(int rank, int nranks, double y0, double y1, int Ny) {
  NH_REQUIRE(!dist.on, "model is already configured for a distributed run");
  // anything that is not the fused configuration (triply periodic, regular, WENO5, no tracers / forcing terms) runs the
  // complete path on its slab: x and y Periodic, z Periodic or Bounded (regular or stretched)
  const bool generic = !fast || getenv("NHSTEP_DIST_GENERIC") != nullptr;
  if (generic) {
    NH_REQUIRE(gh.g.topo[0] == NH_PERIODIC && gh.g.topo[1] == NH_PERIODIC && gh.g.topo[2] != NH_FLAT,
               "a distributed model needs Periodic x and y and a Periodic or Bounded z");
    fast = false;
  }
  FT dy;
  dist.create(gh, rank, nranks, y0, y1, Ny, dy, generic);
  gh.g.dy = dy;                       // the global spacing, so every rank computes with the same metric bits
  if (desc.has_closure == NH_CLOSURE_SMAGORINSKY || desc.has_closure == NH_CLOSURE_SMAGORINSKY_LILLY) build_filter_width();
  pfast.ok = false;                   // the single-GPU solve would wrap y locally
}

template <typename FT>
void Model<FT>::dist_buffer(int which, void** ptr, int64_t* bytes) {
  NH_REQUIRE(dist.on, "model is not configured for a distributed run");
  switch (which) {
    case NH_DIST_SEND_LO: *ptr = dist.send_lo.p; *bytes = (int64_t)(dist.halo_elems * sizeof(FT)); break;
    case NH_DIST_SEND_HI: *ptr = dist.send_hi.p; *bytes = (int64_t)(dist.halo_elems * sizeof(FT)); break;
    case NH_DIST_RECV_LO: *ptr = dist.recv_lo.p; *bytes = (int64_t)(dist.halo_elems * sizeof(FT)); break;
    case NH_DIST_RECV_HI: *ptr = dist.recv_hi.p; *bytes = (int64_t)(dist.halo_elems * sizeof(FT)); break;
    case NH_DIST_SPECTRAL_A: *ptr = dist.A.p; *bytes = (int64_t)(dist.spec_elems * sizeof(CT)); break;
    case NH_DIST_SPECTRAL_B: *ptr = dist.B.p; *bytes = (int64_t)(dist.spec_elems * sizeof(CT)); break;
    default: throw NhError("unknown distributed buffer");
  }
}

// pack (unpack = 0): rows next to both y faces → send_lo / send_hi;  unpack = 1: recv_lo / recv_hi → y halo rows.
// Returns the number of elements per side.
template <typename FT>
int64_t Model<FT>::dist_halo(uint32_t mask, int w, int unpack) {
  NH_REQUIRE(dist.on, "model is not configured for a distributed run");
  const GridDev<FT>& g = gh.g;
  NH_REQUIRE(w >= 1 && w <= g.H[1], "halo width");
  HaloPackJob<FT> J; J.g = g; J.w = w; J.nf = 0;
  for (int f = 0; f < 3; ++f) if (mask & (1u << f)) J.f[J.nf++] = F[f];
  if (mask & (1u << NH_FIELD_PNHS)) J.f[J.nf++] = pNHS;
  NH_REQUIRE(J.nf >= 1 && J.nf <= 3, "between one and three of u, v, w, pNHS per exchange");
  J.lo = unpack ? dist.recv_lo.p : dist.send_lo.p;
  J.hi = unpack ? dist.recv_hi.p : dist.send_hi.p;
  long long n = (long long)(g.N[0] + 2 * g.H[0]) * w * (g.N[2] + 2 * g.H[2]) * J.nf;
  y_halo_pack_kernel<FT><<<(unsigned)ceil_div(n, 256), 256, 0, stream>>>(J, unpack);
  lc.n++;
  CUDA_CHECK(cudaGetLastError());
  return n;
}

// Gⁿ = G(U) and U* = U + Δt(…) in one sweep (the velocities' halos must be current); returns the stage's Δt.
template <typename FT>
double Model<FT>::dist_tendency_substep(double dt, int stage, int euler, int part) {
  NH_REQUIRE(part >= 0 && part <= 2, "part must be 0 (whole sweep), 1 (interior tile rows) or 2 (boundary tile rows)");
  NH_REQUIRE(dist.on, "model is not configured for a distributed run");
  int mode; FT c1, c2; double stage_dt;
  if (desc.timestepper == NH_RK3) {
    NH_REQUIRE(stage >= 0 && stage < 3, "RK3 stage");
    if (stage == 0) { mode = 0; c1 = rk_g1; c2 = FT(0); stage_dt = (double)rk_g1 * dt; }
    else if (stage == 1) { mode = 1; c1 = rk_g2; c2 = rk_z2; stage_dt = (double)(FT)(rk_g2 + rk_z2) * dt; }
    else { mode = 1; c1 = rk_g3; c2 = rk_z3; stage_dt = (double)(FT)(rk_g3 + rk_z3) * dt; }
  } else {
    NH_REQUIRE(stage == 0, "AB2 has a single stage");
    bool eul = euler || (dt != clock.last_dt);
    mode = eul ? 3 : 2; c1 = eul ? FT(-0.5) : chi; c2 = FT(0); stage_dt = dt;
  }
  FastTendJob<FT> J;
  J.g = gh.g; J.tab = tab; J.dt = (FT)dt; J.gamma = c1; J.zeta = c2; J.mode = mode;
  for (int a = 0; a < 3; ++a) { J.U[a] = F[a]; J.Unew[a] = F2[a]; J.Gn[a] = Gn[a]; J.Gm[a] = Gm[a]; }
  if (dist.p2p) J.v_prev = dist.prevF2[1];
  // interior / boundary split: tile rows whose stencils (3 rows either side) stay inside this slab's own rows need no halo row
  // of a neighbour and can run before the barrier that publishes those rows
  const int TY = plane_sweep ? TP_TY : TF_TY, nty = ceil_div(gh.g.N[1], TY);
  const int lo = 1, hi = (gh.g.N[1] - TY - 3) / TY;                 // interior = tile rows lo … hi (needs TY >= 3)
  const bool has_interior = part != 0 && gh.g.N[1] >= 2 * TY + 3 && hi >= lo;
  if (part == 1) {
    if (has_interior) {
      J.ty_split = 0; J.ty_gap = lo; J.ny_tiles = hi - lo + 1;
      launch_fused_sweep(J);
      lc.n++;
      CUDA_CHECK(cudaGetLastError());
    }
    return stage_dt;                 // buffers, tendencies' roles and the clock change hands with part 2
  }
  if (part == 2 && has_interior) { J.ty_split = lo; J.ty_gap = hi + 1 - lo; J.ny_tiles = lo + (nty - 1 - hi); }
  launch_fused_sweep(J, part == 2 && has_interior ? 8 : 0);      // two tile rows: 8-plane segments keep every SM busy
  lc.n++;
  CUDA_CHECK(cudaGetLastError());
  for (int a = 0; a < 3; ++a) { std::swap(F[a], F2[a]); std::swap(dist.prevF[a], dist.prevF2[a]); std::swap(dist.nextF[a], dist.nextF2[a]); }
  for (int f = 0; f < nf; ++f) std::swap(Gn[f], Gm[f]);
  if (desc.timestepper == NH_RK3) {
    if (stage == 0) dist.t_next = clock.time + dt;
    if (stage < 2) { clock.time += stage_dt; clock.stage += 1; clock.last_stage_dt = stage_dt; }
    else {
      // runge_kutta_3.jl:128-136: the last stage's Δt is measured from tⁿ + Δt
      double corrected = dist.t_next - clock.time;
      clock.time += stage_dt; clock.iteration += 1; clock.stage = 1; clock.last_stage_dt = corrected; clock.last_dt = dt;
    }
  } else { clock.time += dt; clock.iteration += 1; clock.stage = 1; clock.last_dt = dt; clock.last_stage_dt = dt; }
  gn_stale = true;
  return stage_dt;
}

template <typename FT>
void Model<FT>::dist_poisson(int phase, double stage_dt) {
  NH_REQUIRE(dist.on, "model is not configured for a distributed run");
  dist.J.g = gh.g;
  if (phase == 0) dist.forward(F[0], F[1], F[2], stage_dt, stream, lc);
  else if (phase == 1) dist.solve_y(stream, lc);
  else if (phase == 2) dist.backward(pNHS, stream, lc);
  else throw NhError("poisson phase must be 0 (forward), 1 (solve along y) or 2 (backward)");
}

template <typename FT>
void Model<FT>::dist_pressure_correct(double stage_dt) {
  NH_REQUIRE(dist.on, "model is not configured for a distributed run");
  YPeers<FT> peers;
  if (dist.p2p) for (int a = 0; a < 3; ++a) { peers.prev[a] = dist.prevF[a]; peers.next[a] = dist.nextF[a]; }
  launch_pressure_correct_halo<FT>(gh.g, F[0], F[1], F[2], pNHS, stage_dt, 0, stream, peers);
  lc.n++;
  p_halo_stale = true;
  CUDA_CHECK(cudaGetLastError());
}

// ---- peer-to-peer mode: CUDA IPC handles of the buffers other ranks write into (or, on the complete path, read) --------
// blob = 10 handles: spectral A, spectral B, u, v, w (current state), u, v, w (ping-pong target), pNHS, barrier flags;
// complete-path models append one handle per halo-filled field id 0 … 31 (zero where the model has no such field)
static const int kIpcHandles = 10;
static const int kIpcHandlesGeneric = kIpcHandles + 32;

template <typename FT>
int64_t Model<FT>::dist_ipc_export(unsigned char* blob, int64_t capacity) {
  NH_REQUIRE(dist.on, "model is not configured for a distributed run");
  const int nh = dist.generic ? kIpcHandlesGeneric : kIpcHandles;
  const int64_t need = (int64_t)nh * sizeof(cudaIpcMemHandle_t);
  if (!blob) return need;
  NH_REQUIRE(capacity >= need, "nh_dist_ipc_export: buffer too small");
  if (!dist.flags.p) dist.flags.alloc(1 << 19);       // zero-filled
  std::vector<void*> ptrs = {dist.A.p, dist.B.p, F[0], F[1], F[2], F2[0], F2[1], F2[2], pNHS, dist.flags.p};
  if (dist.generic) for (int id = 0; id < 32; ++id) ptrs.push_back(halo_field_or_null(id));
  std::memset(blob, 0, need);
  for (int q = 0; q < nh; ++q) {
    if (!ptrs[q]) continue;
    cudaIpcMemHandle_t h;
    CUDA_CHECK(cudaIpcGetMemHandle(&h, ptrs[q]));
    std::memcpy(blob + q * sizeof(h), &h, sizeof(h));
  }
  return need;
}

// blobs: the export of every rank, in rank order.  Opens what this rank writes into: every rank's spectral buffers and
// the fields of the previous / next rank; with one rank (or for this rank's own buffers) the local pointers are used.
template <typename FT>
void Model<FT>::dist_ipc_import(const unsigned char* blobs, int64_t nbytes) {
  NH_REQUIRE(dist.on, "model is not configured for a distributed run");
  NH_REQUIRE(!dist.p2p, "peer buffers are already open");
  const int R = dist.R, rank = dist.rank;
  NH_REQUIRE(R <= PF_MAXR, "peer-to-peer mode supports at most 16 ranks");
  const int nh = dist.generic ? kIpcHandlesGeneric : kIpcHandles;
  const int64_t per = (int64_t)nh * sizeof(cudaIpcMemHandle_t);
  NH_REQUIRE(nbytes == per * R, "nh_dist_ipc_import: expected one export blob per rank");
  NH_REQUIRE(dist.flags.p, "nh_dist_ipc_export must be called before nh_dist_ipc_import");
  std::vector<void*> own = {dist.A.p, dist.B.p, F[0], F[1], F[2], F2[0], F2[1], F2[2], pNHS, dist.flags.p};
  if (dist.generic) for (int id = 0; id < 32; ++id) own.push_back(halo_field_or_null(id));
  std::vector<std::vector<void*>> ptr(R, std::vector<void*>(nh, nullptr));
  auto open = [&](int r, int q) -> void* {
    if (r == rank || !own[q]) return own[q];
    if (ptr[r][q]) return ptr[r][q];
    cudaIpcMemHandle_t h;
    std::memcpy(&h, blobs + r * per + q * sizeof(h), sizeof(h));
    void* out = nullptr;
    CUDA_CHECK(cudaIpcOpenMemHandle(&out, h, cudaIpcMemLazyEnablePeerAccess));
    dist.opened.push_back(out);
    ptr[r][q] = out;
    return out;
  };
  for (int r = 0; r < R; ++r) { dist.peerA[r] = (CT*)open(r, 0); dist.peerB[r] = (CT*)open(r, 1); }
  const int prev = (rank + R - 1) % R, next = (rank + 1) % R;
  if (!dist.generic) {
    for (int a = 0; a < 3; ++a) {
      dist.prevF[a] = (FT*)open(prev, 2 + a); dist.prevF2[a] = (FT*)open(prev, 5 + a);
      dist.nextF[a] = (FT*)open(next, 2 + a); dist.nextF2[a] = (FT*)open(next, 5 + a);
    }
    dist.next_p = (FT*)open(next, 8);
  } else {
    for (int id = 0; id < 32; ++id) {
      dist.peer_prev[id] = (FT*)open(prev, kIpcHandles + id);
      dist.peer_next[id] = (FT*)open(next, kIpcHandles + id);
    }
  }
  for (int r = 0; r < PF_MAXR; ++r) dist.peer_flags[r] = r < R ? (unsigned int*)open(r, 9) : nullptr;
  dist.p2p = true;
}

// Stream-ordered barrier over all ranks of the peer-to-peer group: everything the ranks enqueued before it (their stores
// into peer buffers included) is complete and visible before anything enqueued after it starts.  One small kernel.
template <typename FT>
void Model<FT>::dist_barrier() {
  NH_REQUIRE(dist.on && dist.p2p, "nh_dist_barrier needs the peer-to-peer mode (nh_dist_ipc_import)");
  if (dist.R == 1) return;
  PeerBarrierJob J;
  for (int r = 0; r < PF_MAXR; ++r) J.peer_flags[r] = dist.peer_flags[r];
  J.rank = dist.rank; J.R = dist.R; J.epoch = ++dist.epoch;
  static const unsigned long long timeout_ns = [] {
    const char* e = getenv("NHSTEP_BARRIER_TIMEOUT_S");
    const double s = e ? atof(e) : 600.0;
    return s > 0 ? (unsigned long long)(s * 1e9) : 0ull;
  }();
  J.timeout_ns = timeout_ns;
  peer_barrier_kernel<<<1, 32, 0, stream>>>(J);
  lc.n++;
  CUDA_CHECK(cudaGetLastError());
}

// transpose_z_to_y! (direction 0: block d of the z-local buffer A → block `rank` of rank d's y-local buffer B) and
// transpose_y_to_z! (direction 1: block s of B → block `rank` of rank s's A) on their own, as peer stores over NVLink
// (distributed_transpose.jl:25-191).  The step itself folds these copies into the FFT passes on either side; this entry is the
// pure data movement, bit for bit, for callers and for the round-trip test of test_distributed_transpose.jl:13-54.
template <typename FT>
void Model<FT>::dist_transpose(int direction) {
  NH_REQUIRE(dist.on && dist.p2p, "nh_dist_transpose needs the peer-to-peer mode (nh_dist_ipc_import)");
  NH_REQUIRE(direction == 0 || direction == 1, "direction must be 0 (z-local -> y-local) or 1 (y-local -> z-local)");
  ScatterJob<CT> Sj;
  Sj.src = direction == 0 ? dist.A.p : dist.B.p;
  Sj.chunk = (long long)dist.J.Cx * dist.J.Ny * dist.J.Nz;
  Sj.R = dist.R; Sj.rank = dist.rank;
  for (int d = 0; d < PF_MAXR; ++d) Sj.peer[d] = direction == 0 ? dist.peerB[d] : dist.peerA[d];
  pd_scatter_kernel<CT><<<(unsigned)ceil_div(Sj.chunk * dist.R, 256), 256, 0, stream>>>(Sj);
  lc.n++;
  CUDA_CHECK(cudaGetLastError());
}

#define NH_DIST_INSTANTIATE(FT)                                                          \
  template void Model<FT>::dist_configure(int, int, double, double, int);                \
  template void Model<FT>::dist_buffer(int, void**, int64_t*);                           \
  template int64_t Model<FT>::dist_halo(uint32_t, int, int);                             \
  template double Model<FT>::dist_tendency_substep(double, int, int, int);                    \
  template void Model<FT>::dist_poisson(int, double);                                    \
  template void Model<FT>::dist_pressure_correct(double);                                \
  template int64_t Model<FT>::dist_ipc_export(unsigned char*, int64_t);                  \
  template void Model<FT>::dist_ipc_import(const unsigned char*, int64_t);               \
  template void Model<FT>::dist_barrier();                                               \
  template void Model<FT>::dist_transpose(int);
NH_DIST_INSTANTIATE(float)
NH_DIST_INSTANTIATE(double)
