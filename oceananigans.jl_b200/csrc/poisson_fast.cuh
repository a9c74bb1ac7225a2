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
#include "fft_fast.cuh"

#define PF_L 8   // lines per CTA

template <typename FT, typename CT>
struct PFJob {
  GridDev<FT> g;
  int Nx, Ny, Nz, Mx;            // Mx = Nx/2 + 2: pitch of the half spectrum (complex elements)
  FFPlan px, py, pz;
  const CT *twx, *twy, *twz;
  const double *lx, *ly, *lz;
  CT* S;
  double dt, m;
};

// K1 — grid (Ny/16, Nz), block Nx threads
template <typename FT, typename CT>
__global__ void pf_x_forward_kernel(const __grid_constant__ PFJob<FT, CT> J, const FT* __restrict__ u, const FT* __restrict__ v,
                                    const FT* __restrict__ w) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  const GridDev<FT>& g = J.g;
  const int Nx = J.Nx, pitch = ff_line_pitch(Nx), tid = threadIdx.x, nth = blockDim.x;
  const int k = blockIdx.y + 1, jbase = blockIdx.x * (2 * PF_L) + 1;
  const int kp = k == J.Nz ? 1 : k + 1;
  const FT Ax = g.dy * g.dz, Ay = g.dx * g.dz, Az = g.dx * g.dy;
  const FT Vinv = FT(1) / ((g.dx * g.dy) * g.dz);
  // ---- source term: rhs = divᶜᶜᶜ(U*) / Δt for 16 rows, packed (row 2p → real, row 2p+1 → imaginary)
  for (int e = tid; e < 2 * PF_L * Nx; e += nth) {
    int row = e / Nx, i = e - row * Nx + 1;
    int j = jbase + row;
    FT r = FT(0);
    if (j <= J.Ny) {
      int ip = i == Nx ? 1 : i + 1, jp = j == J.Ny ? 1 : j + 1;
      FT tx = Ax * u[g.idx(ip, j, k)] - Ax * u[g.idx(i, j, k)];
      FT ty = Ay * v[g.idx(i, jp, k)] - Ay * v[g.idx(i, j, k)];
      FT tz = Az * w[g.idx(i, j, kp)] - Az * w[g.idx(i, j, k)];
      FT total = tx + ty;
      total = total + tz;
      FT d = Vinv * total;
      if (sizeof(FT) == 4) r = (FT)((double)d / J.dt); else r = (FT)(d / (FT)J.dt);
    }
    CT* line = X + (row >> 1) * pitch;
    if (row & 1) line[ff_phys(i - 1)].y = r; else line[ff_phys(i - 1)].x = r;
  }
  __syncthreads();
  const int tpl = Nx / 8, l = tid / tpl, t = tid - l * tpl;
  ff_block_fft<CT>(X + l * pitch, J.px, t, tpl, J.twx, false, l < PF_L);
  // ---- Hermitian split and store: A = (Z[k] + conj Z[N-k])/2, B = (Z[k] - conj Z[N-k])/(2i)
  const int half = Nx / 2;
  for (int e = tid; e < 2 * PF_L * J.Mx; e += nth) {
    int row = e / J.Mx, kx = e - row * J.Mx;
    int j = jbase + row;
    if (j > J.Ny) continue;
    CT out; out.x = 0; out.y = 0;
    if (kx <= half) {
      const CT* line = X + (row >> 1) * pitch;
      CT a = line[ff_phys(kx)], b = line[ff_phys((Nx - kx) & (Nx - 1))];
      if (row & 1) { out.x = FT(0.5) * (a.y + b.y); out.y = FT(0.5) * (b.x - a.x); }
      else         { out.x = FT(0.5) * (a.x + b.x); out.y = FT(0.5) * (a.y - b.y); }
    }
    J.S[kx + (long long)J.Mx * ((j - 1) + (long long)J.Ny * (k - 1))] = out;
  }
}

// K2 / K4 — in-place c2c along y.  grid (ceil((Nx/2+1)/8), Nz), block Ny threads
template <typename FT, typename CT>
__global__ void pf_y_kernel(const __grid_constant__ PFJob<FT, CT> J, int inverse) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  const int Ny = J.Ny, pitch = ff_line_pitch(Ny), tid = threadIdx.x, nth = blockDim.x;
  const int kx0 = blockIdx.x * PF_L, ncol = min(PF_L, J.Nx / 2 + 1 - kx0);
  CT* base = J.S + kx0 + (long long)J.Mx * Ny * blockIdx.y;
  for (int e = tid; e < PF_L * Ny; e += nth) {
    int c = e & (PF_L - 1), n = e >> 3;
    if (c < ncol) X[c * pitch + ff_phys(n)] = base[c + (long long)J.Mx * n];
  }
  __syncthreads();
  const int tpl = Ny / 8, l = tid / tpl, t = tid - l * tpl;
  ff_block_fft<CT>(X + l * pitch, J.py, t, tpl, J.twy, inverse != 0, l < ncol);
  const FT scale = inverse ? FT(1) / FT(Ny) : FT(1);
  for (int e = tid; e < PF_L * Ny; e += nth) {
    int c = e & (PF_L - 1), n = e >> 3;
    if (c < ncol) {
      CT val = X[c * pitch + ff_phys(n)];
      if (inverse) { val.x *= scale; val.y *= scale; }
      base[c + (long long)J.Mx * n] = val;
    }
  }
}

// K3 — forward c2c along z, spectral division, inverse c2c along z.  grid (ceil((Nx/2+1)/8), Ny), block Nz threads
template <typename FT, typename CT>
__global__ void pf_z_solve_kernel(const __grid_constant__ PFJob<FT, CT> J) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  const int Nz = J.Nz, pitch = ff_line_pitch(Nz), tid = threadIdx.x, nth = blockDim.x;
  const int kx0 = blockIdx.x * PF_L, ncol = min(PF_L, J.Nx / 2 + 1 - kx0), ky = blockIdx.y;
  const long long sz = (long long)J.Mx * J.Ny;
  CT* base = J.S + kx0 + (long long)J.Mx * ky;
  for (int e = tid; e < PF_L * Nz; e += nth) {
    int c = e & (PF_L - 1), n = e >> 3;
    if (c < ncol) X[c * pitch + ff_phys(n)] = base[c + sz * n];
  }
  __syncthreads();
  const int tpl = Nz / 8, l = tid / tpl, t = tid - l * tpl;
  ff_block_fft<CT>(X + l * pitch, J.pz, t, tpl, J.twz, false, l < ncol);
  // ϕ̂ = -b̂ / (λx + λy + λz - m)  (λ are Float64; fft_based_poisson_solver.jl:110-115)
  const double lyv = J.ly[ky];
  for (int e = tid; e < PF_L * Nz; e += nth) {
    int c = e & (PF_L - 1), n = e >> 3;
    if (c < ncol) {
      double den = J.lx[kx0 + c] + lyv + J.lz[n] - J.m;
      CT b = X[c * pitch + ff_phys(n)], r;
      r.x = (FT)((double)(-b.x) / den);
      r.y = (FT)((double)(-b.y) / den);
      if (J.m == 0.0 && kx0 + c == 0 && ky == 0 && n == 0) { r.x = 0; r.y = 0; }
      X[c * pitch + ff_phys(n)] = r;
    }
  }
  __syncthreads();
  ff_block_fft<CT>(X + l * pitch, J.pz, t, tpl, J.twz, true, l < ncol);
  const FT scale = FT(1) / FT(Nz);
  for (int e = tid; e < PF_L * Nz; e += nth) {
    int c = e & (PF_L - 1), n = e >> 3;
    if (c < ncol) {
      CT val = X[c * pitch + ff_phys(n)];
      val.x *= scale; val.y *= scale;
      base[c + sz * n] = val;
    }
  }
}

// K5 — grid (Ny/16, Nz), block Nx threads: two spectral rows → one complex line → inverse FFT → two real rows of p
template <typename FT, typename CT>
__global__ void pf_x_inverse_kernel(const __grid_constant__ PFJob<FT, CT> J, FT* __restrict__ p) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  const GridDev<FT>& g = J.g;
  const int Nx = J.Nx, pitch = ff_line_pitch(Nx), tid = threadIdx.x, nth = blockDim.x, half = Nx / 2;
  const int k = blockIdx.y + 1, jbase = blockIdx.x * (2 * PF_L) + 1;
  // Z[kx] = A[kx] + i·B[kx] (kx ≤ N/2);  Z[N-kx] = conj(A[kx]) + i·conj(B[kx]) (0 < kx < N/2)
  for (int e = tid; e < PF_L * (half + 1); e += nth) {
    int pr = e / (half + 1), kx = e - pr * (half + 1);
    int j = jbase + 2 * pr;
    CT z, zc; z.x = z.y = zc.x = zc.y = 0;
    if (j <= J.Ny) {
      long long o = kx + (long long)J.Mx * ((j - 1) + (long long)J.Ny * (k - 1));
      CT a = J.S[o], b; b.x = b.y = 0;
      if (j + 1 <= J.Ny) b = J.S[o + J.Mx];
      z.x = a.x - b.y; z.y = a.y + b.x;
      zc.x = a.x + b.y; zc.y = b.x - a.y;
    }
    // the kx = 0 and kx = N/2 modes of a real row are real: Z = Re A + i·Re B (drops rounding-level imaginary parts)
    if (kx == 0 || kx == half) { z.x = FT(0.5) * (z.x + zc.x); z.y = FT(0.5) * (z.y + zc.y); }
    CT* line = X + pr * pitch;
    line[ff_phys(kx)] = z;
    if (kx > 0 && kx < half) line[ff_phys(Nx - kx)] = zc;
  }
  __syncthreads();
  const int tpl = Nx / 8, l = tid / tpl, t = tid - l * tpl;
  ff_block_fft<CT>(X + l * pitch, J.px, t, tpl, J.twx, true, l < PF_L);
  const FT scale = FT(1) / FT(Nx);
  for (int e = tid; e < 2 * PF_L * Nx; e += nth) {
    int row = e / Nx, i = e - row * Nx + 1;
    int j = jbase + row;
    if (j > J.Ny) continue;
    CT val = X[(row >> 1) * pitch + ff_phys(i - 1)];
    p[g.idx(i, j, k)] = ((row & 1) ? val.y : val.x) * scale;
  }
}
