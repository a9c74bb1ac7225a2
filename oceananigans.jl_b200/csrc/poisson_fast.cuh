// poisson_fast.cuh — fused FFT pressure solve for the triply periodic headline configuration (power-of-two sizes).
//
//   solve_for_pressure! + solve! + copy_real_component!     src/Models/NonhydrostaticModels/solve_for_pressure.jl:12-17,78-82
//                                                           src/Solvers/fft_based_poisson_solver.jl:95-137
// Five launches, 12 words of HBM traffic per cell (SURVEY.md §8d phase B) instead of the reference's ≈ 40:
//   K1 x-forward : div(U*)/Δt computed on the fly (periodic wrap, no halo needed) → two real rows packed as one complex
//                  line → FFT → Hermitian split → half spectrum S[kx ≤ Nx/2, j, k]            (read 3w, write w)
//   K2 y-forward : in-place c2c along y on 8-column tiles                                      (2w)
//   K3 z-solve   : c2c along z → ϕ̂ = -b̂/(λx+λy+λz), ϕ̂[0,0,0] = 0 → inverse c2c along z, in shared memory (2w)
//   K4 y-inverse : in-place c2c along y                                                        (2w)
//   K5 x-inverse : Hermitian merge of two rows → inverse FFT → real parts stored into the halo'd pressure field (2w)
// FFT lines are staged in shared memory and transformed with the register-radix butterflies of fft_fast.cuh.
#pragma once
#include "common.cuh"
#include "recon.cuh"
#include "fft_fast.cuh"

#define PF_L 8   // lines per CTA
#define PF_MAXR 16   // ranks of a y-slab run addressable through peer pointers
// Every FFT kernel asks for 1024 resident threads per SM (a 64-register cap): these passes are latency bound, and the
// measured gain from the extra warps outweighs the few spilled values (x-forward 85 → 64 registers: 233 → 170 µs at 256³).
__host__ __device__ constexpr int pf_min_ctas(int n) { return n >= 1024 ? 1 : (1024 / n > 16 ? 16 : 1024 / n); }

template <typename FT, typename CT>
struct PFJob {
  GridDev<FT> g;
  int Nx, Ny, Nz, Mx;            // Mx = Nx/2 + 2: pitch of the half spectrum (complex elements); Ny is the LOCAL extent
  // blocked-kx layout shared with the multi-GPU transposes: element (kx, j, k) lives at
  //   (kx % Cx) + Cx·(j + Ny·(k + Nz·(kx / Cx)));  one GPU: Cx = Mx (a single block)
  int Cx;
  int ywrap;                     // 1: y is locally periodic (index wrap); 0: y halos are filled by neighbour exchange
  FFPlan px, py, pz;
  const CT *twx, *twy, *twz;
  const CT* tw4z;                // exp(-iπk/(2Nz)): DCT-II/III post/pre-twiddles for a Bounded z (nullptr when z is Periodic)
  const FT* dzscale;             // Δzᵃᵃᶜ[k-1]: the Fourier-tridiagonal solver's set_source_term! multiplies b by it (else nullptr)
  const double *lx, *ly, *lz;
  CT* S;
  double dt, m;
  // y-slab runs with peer-to-peer transfers (csrc/dist.cuh): the spectral buffers of every rank and the next rank's
  // pressure field as NVLink peer pointers; all nullptr on one GPU or when the messages go through NCCL
  CT* peer[PF_MAXR];
  FT* p_next;
};

// K1 — grid (Ny/16, Nz), block N = Nx threads.  Thread (l, t): pair-line l (rows j, j+1), points i-1 = t + m·N/8.
// The divergence is evaluated straight into the FFT's input registers; after the transform one shared-memory exchange
// gives every thread the mirrored coefficient Z[N-kx] it needs for the Hermitian split.
// RHS = true: the right-hand side is given (u = a real array laid out by J.g, standalone solve!(ϕ, solver, b)); v, w unused.
template <typename FT, typename CT, int N, bool RHS = false>
__global__ void __launch_bounds__(N, pf_min_ctas(N)) pf_x_forward_kernel(const __grid_constant__ PFJob<FT, CT> J, const FT* __restrict__ u,
                                                         const FT* __restrict__ v, const FT* __restrict__ w) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  const GridDev<FT>& g = J.g;
  constexpr int tpl = N / 8, pitch = ((N + N / 8 + N / 64 + 1) | 1);
  const int tid = threadIdx.x, l = tid / tpl, t = tid - l * tpl;
  const int k = blockIdx.y + 1, j = blockIdx.x * (2 * PF_L) + 2 * l + 1;
  const int kp = k == J.Nz ? 1 : k + 1;
  const FT Ax = g.dy * g.dz, Ay = g.dx * g.dz, Az = g.dx * g.dy;
  const FT Vinv = FT(1) / ((g.dx * g.dy) * g.dz);
  const bool rowa = j <= J.Ny, rowb = j + 1 <= J.Ny;
  const int jb = (J.ywrap && j + 1 > J.Ny) ? j + 1 - J.Ny : j + 1, jc = (J.ywrap && j + 2 > J.Ny) ? j + 2 - J.Ny : j + 2;
  const long long ra = g.idx(1, j, k), rb = g.idx(1, jb, k), rc = g.idx(1, jc, k);
  const long long raz = g.idx(1, j, kp), rbz = g.idx(1, jb, kp);
  CT z[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int i0 = t + m * tpl;                 // zero-based i
    const int ip = i0 + 1 == N ? 0 : i0 + 1;
    FT da = FT(0), db = FT(0);
    if (RHS) {
      if (rowa) da = u[ra + i0];
      if (rowb) db = u[rb + i0];
      if (J.dzscale != nullptr) { const FT dzk = J.dzscale[k - 1]; da = da * dzk; db = db * dzk; }
    } else if (rowa) {
      FT vb = v[rb + i0];
      FT tx = Ax * u[ra + ip] - Ax * u[ra + i0];
      FT ty = Ay * vb - Ay * v[ra + i0];
      FT tz = Az * w[raz + i0] - Az * w[ra + i0];
      FT total = tx + ty; total = total + tz;
      FT d = Vinv * total;
      if (sizeof(FT) == 4) da = (FT)((double)d / J.dt); else da = (FT)(d / (FT)J.dt);
      if (rowb) {
        tx = Ax * u[rb + ip] - Ax * u[rb + i0];
        ty = Ay * v[rc + i0] - Ay * vb;
        tz = Az * w[rbz + i0] - Az * w[rb + i0];
        total = tx + ty; total = total + tz;
        d = Vinv * total;
        if (sizeof(FT) == 4) db = (FT)((double)d / J.dt); else db = (FT)(d / (FT)J.dt);
      }
    }
    z[m].x = da; z[m].y = db;
  }
  CT* Xl = X + l * pitch;
  ff_line_fft_regs<CT, N, false, 1>(z, Xl, t, J.twx);
  // exchange: natural-order spectrum into shared memory, then A = (Z[k] + conj Z[N-k])/2, B = (Z[k] - conj Z[N-k])/(2i)
#pragma unroll
  for (int m = 0; m < 8; ++m) Xl[ff_phys(t + m * tpl)] = z[m];
  __syncthreads();
  if (!rowa) return;
  CT* Sa = J.S + (long long)J.Cx * ((j - 1) + (long long)J.Ny * (k - 1));
  const long long blk = (long long)J.Cx * J.Ny * J.Nz;
#pragma unroll
  for (int m = 0; m < 5; ++m) {
    const int kx = t + m * tpl;
    if (m == 4 && t > 1) break;                 // kx = N/2 (t = 0) and the zero pad slot kx = N/2 + 1 (t = 1)
    CT a, b, oa, ob;
    if (m < 4) a = z[m]; else a = Xl[ff_phys(N / 2)];
    b = Xl[ff_phys((N - kx) & (N - 1))];
    oa.x = FT(0.5) * (a.x + b.x); oa.y = FT(0.5) * (a.y - b.y);
    ob.x = FT(0.5) * (a.y + b.y); ob.y = FT(0.5) * (b.x - a.x);
    if (m == 4 && t == 1) { oa.x = oa.y = ob.x = ob.y = 0; }
    if (m == 4 && t == 1) { if (J.Cx != J.Mx) break; }          // the pad slot exists only in the single-block layout
    const long long o = (kx % J.Cx) + blk * (kx / J.Cx);
    Sa[o] = oa;
    if (rowb) Sa[o + J.Cx] = ob;
  }
}

// K2 / K4 — in-place c2c along y on an 8-column tile.  grid (ceil((Nx/2+1)/8), Nz), block N = Ny threads.
// Lanes run across the 8 adjacent kx columns (128-byte segments); each thread moves its 8 points global → registers,
// the first butterfly stage runs before anything touches shared memory and the last one writes straight back.
template <typename FT, typename CT, int N, bool INV>
__global__ void __launch_bounds__(N, pf_min_ctas(N)) pf_y_kernel(const __grid_constant__ PFJob<FT, CT> J) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  constexpr int pitch = FFPitch8<N>::value;
  const int tid = threadIdx.x, c = tid & (PF_L - 1), t = tid >> 3;
  const int kx = blockIdx.x * PF_L + c;
  const bool active = kx <= J.Nx / 2;
  CT* base = J.S + kx + (long long)J.Mx * N * blockIdx.y + (long long)J.Mx * t;
  const long long step = (long long)J.Mx * (N / 8);
  CT v[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) { if (active) v[m] = base[m * step]; else { v[m].x = 0; v[m].y = 0; } }
  ff_line_fft_regs<CT, N, INV>(v, X + c * pitch, t, J.twy);
  if (active) {
    const FT scale = FT(1) / FT(N);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      if (INV) { v[m].x *= scale; v[m].y *= scale; }
      base[m * step] = v[m];
    }
  }
}

// K3 — forward c2c along z, spectral division in registers, inverse c2c along z.  grid (ceil((Nx/2+1)/8), Ny), block Nz
// DCT = true: z is Bounded.  The reference's REDFT10 / REDFT01 along z (discrete_transforms.jl, index_permutations.jl:18-90)
// become a complex FFT of the even/odd-permuted column (Makhoul) with the twiddles ω_k = exp(-iπk/(2N)):
//   forward  Y[k] = ω_k V[k] + conj(ω_k) V[N-k];   backward  V[k] = conj(ω_k)(Y[k] - i·Y[N-k])/2, Y[N] := 0
// The permutation is folded into the global load/store index (position p of the FFT line ↔ plane 2p or 2(N-1-p)+1); the
// mirrored coefficient comes from the other half of the line through one shared-memory exchange each way.
template <typename FT, typename CT, int N, bool DCT = false>
__global__ void __launch_bounds__(N, pf_min_ctas(N)) pf_z_solve_kernel(const __grid_constant__ PFJob<FT, CT> J) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  constexpr int pitch = FFPitch8<N>::value;
  const int tid = threadIdx.x, c = tid & (PF_L - 1), t = tid >> 3;
  const int kx = blockIdx.x * PF_L + c, ky = blockIdx.y;
  const bool active = kx <= J.Nx / 2;
  const long long sz = (long long)J.Mx * J.Ny;
  CT* base = J.S + kx + (long long)J.Mx * ky;
  CT* Xl = X + c * pitch;
  long long off[8];
  CT v[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int p = t + m * (N / 8);
    off[m] = sz * (DCT ? (p < N / 2 ? 2 * p : 2 * (N - 1 - p) + 1) : p);
    if (active) v[m] = base[off[m]]; else { v[m].x = 0; v[m].y = 0; }
  }
  ff_line_fft_regs<CT, N, false>(v, Xl, t, J.twz);
  if (DCT) {
#pragma unroll
    for (int m = 0; m < 8; ++m) Xl[ff_phys8(t + m * (N / 8))] = v[m];
    __syncthreads();
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const int p = t + m * (N / 8);
      const CT w = J.tw4z[p], b = Xl[ff_phys8((N - p) & (N - 1))];
      CT wc; wc.x = w.x; wc.y = -w.y;
      const CT a1 = ff_mul(w, v[m]), a2 = ff_mul(wc, b);
      v[m].x = a1.x + a2.x; v[m].y = a1.y + a2.y;
    }
    __syncthreads();
  }
  // ϕ̂ = -b̂ / (λx + λy + λz - m)  (λ are Float64; fft_based_poisson_solver.jl:110-115)
  const double lxy = active ? J.lx[kx] + J.ly[ky] : 1.0;
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int n = t + m * (N / 8);
    double den = lxy + J.lz[n] - J.m;
    const bool den_ok = nh_den_ok(den);
    const double yr = nh_rcp_refined(den);          // shared by the two components (recon.cuh: same bits as `/`)
    CT r;
    r.x = (FT)nh_div_checked((double)(-v[m].x), den, yr, den_ok);
    r.y = (FT)nh_div_checked((double)(-v[m].y), den, yr, den_ok);
    if (J.m == 0.0 && kx == 0 && ky == 0 && n == 0) { r.x = 0; r.y = 0; }
    v[m] = r;
  }
  if (DCT) {
#pragma unroll
    for (int m = 0; m < 8; ++m) Xl[ff_phys8(t + m * (N / 8))] = v[m];
    __syncthreads();
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const int p = t + m * (N / 8);
      CT b; b.x = 0; b.y = 0;
      if (p > 0) b = Xl[ff_phys8(N - p)];
      CT s; s.x = v[m].x + b.y; s.y = v[m].y - b.x;               // Y[k] - i·Y[N-k]
      CT wc = J.tw4z[p]; wc.y = -wc.y;
      const CT r = ff_mul(wc, s);
      v[m].x = r.x * FT(0.5); v[m].y = r.y * FT(0.5);
    }
    __syncthreads();
  }
  ff_line_fft_regs<CT, N, true>(v, Xl, t, J.twz);
  if (active) {
    const FT scale = FT(1) / FT(N);
#pragma unroll
    for (int m = 0; m < 8; ++m) { v[m].x *= scale; v[m].y *= scale; base[off[m]] = v[m]; }
  }
}

// K5 — grid (Ny/16, Nz), block N = Nx threads: two spectral rows → one complex line → inverse FFT → two real rows of p
template <typename FT, typename CT, int N>
__global__ void __launch_bounds__(N, pf_min_ctas(N)) pf_x_inverse_kernel(const __grid_constant__ PFJob<FT, CT> J, FT* __restrict__ p) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  const GridDev<FT>& g = J.g;
  constexpr int tpl = N / 8, pitch = ((N + N / 8 + N / 64 + 1) | 1), half = N / 2;
  const int tid = threadIdx.x, l = tid / tpl, t = tid - l * tpl;
  const int k = blockIdx.y + 1, j = blockIdx.x * (2 * PF_L) + 2 * l + 1;
  const bool rowa = j <= J.Ny, rowb = j + 1 <= J.Ny;
  CT* Xl = X + l * pitch;
  // Z[kx] = A[kx] + i·B[kx] (kx ≤ N/2);  Z[N-kx] = conj(A[kx]) + i·conj(B[kx]) (0 < kx < N/2)
  const CT* Sa = J.S + (long long)J.Cx * ((j - 1) + (long long)J.Ny * (k - 1));
  const long long blk = (long long)J.Cx * J.Ny * J.Nz;
#pragma unroll
  for (int m = 0; m < 5; ++m) {
    const int kx = t + m * tpl;
    if (m == 4 && t > 0) break;
    CT a, b; a.x = a.y = b.x = b.y = 0;
    const long long o = (kx % J.Cx) + blk * (kx / J.Cx);
    if (rowa) a = Sa[o];
    if (rowb) b = Sa[o + J.Cx];
    CT zz, zc;
    zz.x = a.x - b.y; zz.y = a.y + b.x;
    zc.x = a.x + b.y; zc.y = b.x - a.y;
    // the kx = 0 and kx = N/2 modes of a real row are real: Z = Re A + i·Re B
    if (kx == 0 || kx == half) { zz.x = a.x; zz.y = b.x; }
    Xl[ff_phys(kx)] = zz;
    if (kx > 0 && kx < half) Xl[ff_phys(N - kx)] = zc;
  }
  __syncthreads();
  CT z[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) z[m] = Xl[ff_phys(t + m * tpl)];
  __syncthreads();
  ff_line_fft_regs<CT, N, true, 1>(z, Xl, t, J.twx);
  if (!rowa) return;
  const FT scale = FT(1) / FT(N);
  const long long ra = g.idx(1, j, k), rb = g.idx(1, rowb ? j + 1 : j, k);
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int i0 = t + m * tpl;
    p[ra + i0] = z[m].x * scale;
    if (rowb) p[rb + i0] = z[m].y * scale;
  }
  // the last local row is row 0 of the next rank's pressure (its v-correction reads p[j-1])
  if (J.p_next != nullptr && (j == J.Ny || (rowb && j + 1 == J.Ny))) {
    const long long r0 = g.idx(1, 0, k);
#pragma unroll
    for (int m = 0; m < 8; ++m) J.p_next[r0 + t + m * tpl] = (j == J.Ny ? z[m].x : z[m].y) * scale;
  }
}

// Fourier-tridiagonal solve on the half spectrum: one thread per (kx, ky) column, lanes across kx (coalesced planes).
//   solve_batched_tridiagonal_system_z! — src/Solvers/batched_tridiagonal_solver.jl:209-235 (Thomas sweep with the
//   |β| > 10·eps guard: where it fails ϕ keeps what the previous solve left there — `Phi` persists between solves);
//   main diagonal D = dA[k] - dC[k]·(λx + λy) in Float64 — compute_main_diagonal!, fourier_tridiagonal_poisson_solver.jl:41-51
//   (dA, dC are the grid-only parts, tabulated on the host with the reference's expression);
//   ϕ .= ϕ .- mean(ϕ) — fourier_tridiagonal_poisson_solver.jl:140-147: a real-space constant lives in the (0,0) column, so
//   the thread that owns it subtracts the column's k-average there, before the inverse transforms.
// The positions where the guard fails depend only on the matrix, not on the right-hand side (in practice the single
// singular entry (kx, ky, k) = (0, 0, Nz)).  The reference's ϕ array holds the previous solve's REAL-SPACE solution at
// that moment, so what it "keeps" there is ϕ_prev[i = kx+1, j = ky+1, k]: PFStale lists those positions (filled by the
// first solve) and pf_stale_gather_kernel refreshes their values from the real-space output after every solve.
// f = ϕ̂ = J.S in place (forward-transformed source in, spectral solution out), grid (ceil((Nx/2+1)/128), Ny), block 128
#define PF_STALE_MAX 64
template <typename FT>
struct PFStale { int n; int pos[PF_STALE_MAX][3]; FT val[PF_STALE_MAX]; };

template <typename FT>
__global__ void pf_stale_gather_kernel(PFStale<FT>* st, const FT* __restrict__ phi, GridDev<FT> g) {
  const int q = threadIdx.x;
  if (q < min(st->n, PF_STALE_MAX) && st->pos[q][1] < g.N[1]) st->val[q] = phi[g.idx(st->pos[q][0] + 1, st->pos[q][1] + 1, st->pos[q][2] + 1)];
}

// DIST: the y-local buffer of a slab run (csrc/dist.cuh): this rank's Cx columns kx = rank·Cx + kxl, all GLOBAL rows ky,
// element (kxl, ky, k) at kxl + Cx·(ky % NyL) + Cx·NyL·(k + Nz·(ky / NyL)); J.S is that buffer, J.Ny = NyL.
template <typename FT, typename CT, bool DIST = false>
__global__ void __launch_bounds__(128) pf_thomas_kernel(const __grid_constant__ PFJob<FT, CT> J, const FT* __restrict__ lower,
                                                        const double* __restrict__ dA, const double* __restrict__ dC,
                                                        CT* __restrict__ Phi, FT* __restrict__ tt, PFStale<FT>* st, int rank = 0) {
  const int kxl = blockIdx.x * blockDim.x + threadIdx.x, ky = blockIdx.y;
  const int kx = DIST ? rank * J.Cx + kxl : kxl;
  if (kx > J.Nx / 2 || (DIST && kxl >= J.Cx)) return;
  const int Nz = J.Nz;
  const long long plane = (long long)J.Cx * J.Ny;
  const long long sz = DIST ? plane : (long long)J.Mx * J.Ny;
  const long long col = DIST ? kxl + (long long)J.Cx * (ky % J.Ny) + plane * Nz * (ky / J.Ny) : kx + (long long)J.Mx * ky;
  const double lam = J.lx[kx] + J.ly[ky];
  const double guard = 10.0 * (sizeof(FT) == 4 ? 1.1920928955078125e-07 : 2.220446049250313e-16);
  double beta = dA[0] - dC[0] * lam;
  const CT f1 = J.S[col];
  CT prev; prev.x = (FT)((double)f1.x / beta); prev.y = (FT)((double)f1.y / beta);
  Phi[col] = prev;
  for (int k = 2; k <= Nz; ++k) {
    const FT ak = lower[k - 2];
    const FT tk = (FT)((double)ak / beta);
    tt[col + sz * (k - 1)] = tk;
    beta = (dA[k - 1] - dC[k - 1] * lam) - (double)(ak * tk);
    const CT fk = J.S[col + sz * (k - 1)];
    CT num; num.x = fk.x - ak * prev.x; num.y = fk.y - ak * prev.y;
    if (fabs(beta) > guard) { prev.x = (FT)((double)num.x / beta); prev.y = (FT)((double)num.y / beta); }
    else {
      // keep what the previous solve left at this index (zero before the first solve)
      prev.x = FT(0); prev.y = FT(0);
      const int n = min(st->n, PF_STALE_MAX);
      int found = -1;
      for (int q = 0; q < n; ++q) if (st->pos[q][0] == kx && st->pos[q][1] == ky && st->pos[q][2] == k - 1) found = q;
      if (found >= 0) prev.x = st->val[found];
      else {
        const int q = atomicAdd(&st->n, 1);
        if (q < PF_STALE_MAX) { st->pos[q][0] = kx; st->pos[q][1] = ky; st->pos[q][2] = k - 1; st->val[q] = FT(0); }
      }
    }
    Phi[col + sz * (k - 1)] = prev;
  }
  double sx = prev.x, sy = prev.y;
  for (int k = Nz - 1; k >= 1; --k) {
    CT here = Phi[col + sz * (k - 1)];
    const FT tk = tt[col + sz * k];
    here.x = here.x - tk * prev.x; here.y = here.y - tk * prev.y;
    Phi[col + sz * (k - 1)] = here;
    prev = here;
    sx += here.x; sy += here.y;
  }
  if (kx == 0 && ky == 0) {
    const FT mx = (FT)(sx / Nz), my = (FT)(sy / Nz);
    for (int k = 0; k < Nz; ++k) { CT h = Phi[col + sz * k]; h.x -= mx; h.y -= my; Phi[col + sz * k] = h; }
  }
}
