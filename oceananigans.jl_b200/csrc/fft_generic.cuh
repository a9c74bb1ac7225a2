// fft_generic.cuh — the complete (any size, any topology) spectral machinery: per-axis line transforms staged in
// shared memory (Stockham radix-2 for power-of-two lines, direct DFT otherwise), DCT-II/III through Makhoul's
// permutation folded into the load/store indexing, spectral division, batched Thomas sweep, mean removal.
//   src/Solvers/fft_based_poisson_solver.jl:95-137, plan_transforms.jl:16-146, discrete_transforms.jl:26-184,
//   index_permutations.jl:18-90, batched_tridiagonal_solver.jl:203-235, fourier_tridiagonal_poisson_solver.jl:41-147
// The fused real-to-complex fast path lives in poisson_fast.cu.
#pragma once
#include "common.cuh"

template <typename FT> struct cx_of;
template <> struct cx_of<double> { typedef double2 type; };
template <> struct cx_of<float> { typedef float2 type; };

template <typename CT> __device__ __forceinline__ CT cmul(CT a, CT b) {
  CT r; r.x = a.x * b.x - a.y * b.y; r.y = a.x * b.y + a.y * b.x; return r;
}
template <typename CT> __device__ __forceinline__ CT cadd(CT a, CT b) { CT r; r.x = a.x + b.x; r.y = a.y + b.y; return r; }
template <typename CT> __device__ __forceinline__ CT csub(CT a, CT b) { CT r; r.x = a.x - b.x; r.y = a.y - b.y; return r; }
template <typename CT> __device__ __forceinline__ CT cconj(CT a) { a.y = -a.y; return a; }

enum { LT_FWD_FFT = 0, LT_INV_FFT = 1, LT_FWD_DCT = 2, LT_INV_DCT = 3 };

struct LineJob {
  int N;                 // line length
  long long S;           // element stride along the line
  int P;                 // line id = p + P*q; base offset = p*bp + q*bq
  long long bp, bq;
  long long nlines;
  int mode;
  int pow2;
  int lpb;               // lines per block
};

// tw[m] = exp(-2πi m/N) (m < N), tw4[k] = exp(-iπk/(2N)) (k < N), both at the working precision's complex type
template <typename FT, typename CT>
__global__ void __launch_bounds__(256) line_transform_kernel(CT* __restrict__ A, const __grid_constant__ LineJob J,
                                                             const CT* __restrict__ tw, const CT* __restrict__ tw4) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  CT* X = reinterpret_cast<CT*>(smem_raw);
  const int N = J.N, L = J.lpb;
  CT* Y = X + (size_t)L * N;
  const long long line0 = (long long)blockIdx.x * L;
  const int nl = (int)min((long long)L, J.nlines - line0);
  const int T = blockDim.x, tid = threadIdx.x;
  const bool inverse = (J.mode == LT_INV_FFT || J.mode == LT_INV_DCT);

  // ---- load (lines fastest across threads when the line stride is not 1, elements fastest otherwise) ----
  for (int t = tid; t < nl * N; t += T) {
    int l, m;
    if (J.S == 1) { m = t % N; l = t / N; } else { l = t % nl; m = t / nl; }
    long long lid = line0 + l;
    CT v = A[(lid % J.P) * J.bp + (lid / J.P) * J.bq + m * J.S];
    int dst = m;
    if (J.mode == LT_FWD_DCT) dst = (m & 1) ? N - 1 - (m >> 1) : (m >> 1);
    X[l * N + dst] = v;
  }
  __syncthreads();

  if (J.mode == LT_INV_DCT) {
    // V[k] = conj(ω_k)·(Y[k] - i·Y[N-k])/2, Y[N] := 0
    for (int t = tid; t < nl * N; t += T) {
      int l = t / N, k = t % N;
      CT a = X[l * N + k];
      CT b; if (k == 0) { b.x = 0; b.y = 0; } else b = X[l * N + N - k];
      CT s; s.x = a.x + b.y; s.y = a.y - b.x;                    // a - i·b
      CT r = cmul(cconj(tw4[k]), s);
      r.x *= FT(0.5); r.y *= FT(0.5);
      Y[l * N + k] = r;
    }
    __syncthreads();
    CT* tmp = X; X = Y; Y = tmp;
  }

  // ---- DFT of every staged line ----
  if (J.pow2) {
    const int half = N >> 1;
    for (int Ns = 1; Ns < N; Ns <<= 1) {
      const int tstride = half / Ns;                              // N / (2 Ns)
      for (int t = tid; t < nl * half; t += T) {
        int l = t / half, j = t % half;
        int k = j & (Ns - 1);
        CT w = tw[k * tstride];
        if (inverse) w.y = -w.y;
        CT a = X[l * N + j];
        CT b = cmul(w, X[l * N + j + half]);
        int j0 = ((j - k) << 1) + k;
        Y[l * N + j0] = cadd(a, b);
        Y[l * N + j0 + Ns] = csub(a, b);
      }
      __syncthreads();
      CT* tmp = X; X = Y; Y = tmp;
    }
  } else if (N > 1) {
    for (int t = tid; t < nl * N; t += T) {
      int l = t / N, k = t % N;
      CT acc; acc.x = 0; acc.y = 0;
      int idx = 0;
      for (int n = 0; n < N; ++n) {
        CT w = tw[idx];
        if (inverse) w.y = -w.y;
        acc = cadd(acc, cmul(X[l * N + n], w));
        idx += k; if (idx >= N) idx -= N;
      }
      Y[l * N + k] = acc;
    }
    __syncthreads();
    CT* tmp = X; X = Y; Y = tmp;
  }

  if (J.mode == LT_FWD_DCT) {
    // Y[k] = ω_k V[k] + conj(ω_k) V[(N-k) mod N]
    for (int t = tid; t < nl * N; t += T) {
      int l = t / N, k = t % N;
      CT w = tw4[k];
      CT r = cadd(cmul(w, X[l * N + k]), cmul(cconj(w), X[l * N + (k == 0 ? 0 : N - k)]));
      Y[l * N + k] = r;
    }
    __syncthreads();
    CT* tmp = X; X = Y; Y = tmp;
  }

  // ---- store ----
  const FT scale = inverse ? FT(1) / FT(N) : FT(1);
  for (int t = tid; t < nl * N; t += T) {
    int l, m;
    if (J.S == 1) { m = t % N; l = t / N; } else { l = t % nl; m = t / nl; }
    long long lid = line0 + l;
    int src = m;
    if (J.mode == LT_INV_DCT) src = (m & 1) ? N - 1 - (m >> 1) : (m >> 1);
    CT v = X[l * N + src];
    if (inverse) { v.x *= scale; v.y *= scale; }
    A[(lid % J.P) * J.bp + (lid / J.P) * J.bq + m * J.S] = v;
  }
}

// ϕ̂ = -b̂ / (λx + λy + λz - m); [1,1,1] := 0 when m == 0 — fft_based_poisson_solver.jl:110-115 (λ are Float64)
template <typename FT, typename CT>
__global__ void __launch_bounds__(256) spectral_divide_kernel(CT* A, int Nx, int Ny, int Nz, const double* lx, const double* ly,
                                                              const double* lz, double m, int zero_mean) {
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  long long tot = (long long)Nx * Ny * Nz;
  if (t >= tot) return;
  int i = (int)(t % Nx), j = (int)((t / Nx) % Ny), k = (int)(t / ((long long)Nx * Ny));
  double den = lx[i] + ly[j] + lz[k] - m;
  CT b = A[t];
  CT r;
  r.x = (FT)((double)(-b.x) / den);
  r.y = (FT)((double)(-b.y) / den);
  if (zero_mean && t == 0) { r.x = 0; r.y = 0; }
  A[t] = r;
}

// compute_main_diagonal! (ZDirection) — fourier_tridiagonal_poisson_solver.jl:41-51; D is Float64
template <typename FT>
__global__ void main_diagonal_kernel(GridDev<FT> g, const double* lx, const double* ly, double* D) {
  int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y * blockDim.y + threadIdx.y;
  int Nx = g.N[0], Ny = g.N[1], Nz = g.N[2];
  if (i >= Nx || j >= Ny) return;
  double lam = lx[i] + ly[j];
  long long col = i + (long long)Nx * j, sz = (long long)Nx * Ny;
  D[col] = (double)(FT(-1) / g.Dzf(2)) - (double)g.Dzc(1) * lam;
  for (int k = 2; k <= Nz - 1; ++k)
    D[col + sz * (k - 1)] = (double)(-(FT(1) / g.Dzf(k + 1) + FT(1) / g.Dzf(k))) - (double)g.Dzc(k) * lam;
  D[col + sz * (Nz - 1)] = (double)(FT(-1) / g.Dzf(Nz)) - (double)g.Dzc(Nz) * lam;
}

// solve_batched_tridiagonal_system_z! — batched_tridiagonal_solver.jl:209-235.  a = c = lower (FT, length Nz-1),
// b = D (Float64, 3-D), f and ϕ complex, t real scratch.  ϕ keeps its previous content where the guard fails.
template <typename FT, typename CT>
__global__ void __launch_bounds__(128) thomas_z_kernel(int Nx, int Ny, int Nz, const FT* lower, const double* D, const CT* f,
                                                       CT* phi, FT* tt) {
  long long col = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  long long sz = (long long)Nx * Ny;
  if (col >= sz) return;
  const double guard = 10.0 * (sizeof(FT) == 4 ? 1.1920928955078125e-07 : 2.220446049250313e-16);
  double beta = D[col];
  CT f1 = f[col];
  CT prev; prev.x = (FT)((double)f1.x / beta); prev.y = (FT)((double)f1.y / beta);
  phi[col] = prev;
  for (int k = 2; k <= Nz; ++k) {
    FT ck = lower[k - 2], ak = lower[k - 2];
    FT tk = (FT)((double)ck / beta);
    tt[col + sz * (k - 1)] = tk;
    beta = D[col + sz * (k - 1)] - (double)(ak * tk);
    CT fk = f[col + sz * (k - 1)];
    CT num; num.x = fk.x - ak * prev.x; num.y = fk.y - ak * prev.y;
    CT star; star.x = (FT)((double)num.x / beta); star.y = (FT)((double)num.y / beta);
    CT old = phi[col + sz * (k - 1)];
    prev = fabs(beta) > guard ? star : old;
    phi[col + sz * (k - 1)] = prev;
  }
  for (int k = Nz - 1; k >= 1; --k) {
    CT here = phi[col + sz * (k - 1)];
    FT tk = tt[col + sz * k];
    here.x = here.x - tk * prev.x; here.y = here.y - tk * prev.y;
    phi[col + sz * (k - 1)] = here;
    prev = here;
  }
}

// generic Thomas sweep along any direction for the standalone BatchedTridiagonalSolver (real rhs, 1-D or 3-D b)
template <typename FT>
__global__ void thomas_generic_kernel(int Nx, int Ny, int Nz, int dir, const FT* a, const FT* b, int b3d, const FT* c,
                                      const FT* f, FT* phi, FT* tt) {
  int n[3] = {Nx, Ny, Nz};
  int e1 = dir == 0 ? 1 : 0, e2 = dir == 2 ? 1 : 2;
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= (long long)n[e1] * n[e2]) return;
  int q[3]; q[e1] = (int)(t % n[e1]); q[e2] = (int)(t / n[e1]);
  long long st[3] = {1, Nx, (long long)Nx * Ny};
  long long base = q[e1] * st[e1] + q[e2] * st[e2], s = st[dir];
  const int N = n[dir];
  const FT guard = FT(10) * (sizeof(FT) == 4 ? FT(1.1920928955078125e-07) : FT(2.220446049250313e-16));
  auto B = [&](int k) { return b3d ? b[base + s * k] : b[k]; };
  FT beta = B(0);
  FT prev = f[base] / beta;
  phi[base] = prev;
  for (int k = 1; k < N; ++k) {
    FT tk = c[k - 1] / beta;
    tt[base + s * k] = tk;
    beta = B(k) - a[k - 1] * tk;
    FT star = (f[base + s * k] - a[k - 1] * prev) / beta;
    FT old = phi[base + s * k];
    prev = fabs(beta) > guard ? star : old;
    phi[base + s * k] = prev;
  }
  for (int k = N - 2; k >= 0; --k) {
    FT here = phi[base + s * k] - tt[base + s * (k + 1)] * prev;
    phi[base + s * k] = here;
    prev = here;
  }
}

// Σ ϕ over the whole array → partial sums (finished by a second launch); then ϕ -= mean
template <typename CT>
__global__ void __launch_bounds__(256) csum_kernel(const CT* A, long long n, double2* partial) {
  double sx = 0, sy = 0;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
    sx += (double)A[t].x; sy += (double)A[t].y;
  }
  __shared__ double ax[256], ay[256];
  ax[threadIdx.x] = sx; ay[threadIdx.x] = sy;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) { ax[threadIdx.x] += ax[threadIdx.x + s]; ay[threadIdx.x] += ay[threadIdx.x + s]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { partial[blockIdx.x].x = ax[0]; partial[blockIdx.x].y = ay[0]; }
}
static __global__ void csum_final_kernel(double2* partial, int n) {
  __shared__ double ax[256], ay[256];
  double sx = 0, sy = 0;
  for (int t = threadIdx.x; t < n; t += 256) { sx += partial[t].x; sy += partial[t].y; }
  ax[threadIdx.x] = sx; ay[threadIdx.x] = sy;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) { ax[threadIdx.x] += ax[threadIdx.x + s]; ay[threadIdx.x] += ay[threadIdx.x + s]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { partial[0].x = ax[0]; partial[0].y = ay[0]; }
}
template <typename FT, typename CT>
__global__ void __launch_bounds__(256) subtract_mean_kernel(CT* A, long long n, const double2* sum) {
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  FT mx = (FT)(sum[0].x / (double)n), my = (FT)(sum[0].y / (double)n);
  CT v = A[t]; v.x = v.x - mx; v.y = v.y - my; A[t] = v;
}

// real (N³ dense) → complex storage, optionally × Δzᶜ[k] (set_source_term!, fourier_tridiagonal_poisson_solver.jl:155-177)
template <typename FT, typename CT>
__global__ void __launch_bounds__(256) real_to_complex_kernel(const FT* b, CT* A, long long n, long long nxy, const FT* dzc1, int times_dz) {
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= n) return;
  FT v = b[t];
  if (times_dz) v = v * dzc1[t / nxy];
  CT o; o.x = v; o.y = FT(0); A[t] = o;
}
template <typename FT, typename CT>
__global__ void __launch_bounds__(256) complex_real_part_kernel(const CT* A, FT* out, long long n) {
  long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t < n) out[t] = A[t].x;
}
