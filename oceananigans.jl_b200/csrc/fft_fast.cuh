// fft_fast.cuh — register-radix FFT building blocks for the fused Poisson fast path (poisson_fast.cuh).
// A line of N = 2^m complex points lives in shared memory; every thread owns 8 points per stage and runs radix-8/4/2
// Stockham butterflies in registers, so a 512-point line needs 3 shared-memory round trips instead of 9.
// Twiddles come from a precomputed full-precision table (exp(-2πi m/N)), never from device sincos.
// All butterflies are __host__ __device__ so tests/ can exercise them on the CPU through a small nvcc-built harness.
#pragma once
#include <cuda_runtime.h>

#ifndef FF_HD
#define FF_HD __host__ __device__ __forceinline__
#endif

template <typename CT> FF_HD CT ff_add(CT a, CT b) { CT r; r.x = a.x + b.x; r.y = a.y + b.y; return r; }
template <typename CT> FF_HD CT ff_sub(CT a, CT b) { CT r; r.x = a.x - b.x; r.y = a.y - b.y; return r; }
template <typename CT> FF_HD CT ff_mul(CT a, CT b) { CT r; r.x = a.x * b.x - a.y * b.y; r.y = a.x * b.y + a.y * b.x; return r; }
template <typename CT> FF_HD CT ff_conj(CT a) { a.y = -a.y; return a; }
// multiply by -i (forward) or +i (inverse)
template <typename CT> FF_HD CT ff_rot(CT a, bool inv) { CT r; if (inv) { r.x = -a.y; r.y = a.x; } else { r.x = a.y; r.y = -a.x; } return r; }

template <typename CT> FF_HD void ff_fft2(CT* v) { CT a = v[0], b = v[1]; v[0] = ff_add(a, b); v[1] = ff_sub(a, b); }

template <typename CT> FF_HD void ff_fft4(CT* v, bool inv) {
  CT a0 = ff_add(v[0], v[2]), a1 = ff_sub(v[0], v[2]);
  CT a2 = ff_add(v[1], v[3]), a3 = ff_rot(ff_sub(v[1], v[3]), inv);
  v[0] = ff_add(a0, a2); v[2] = ff_sub(a0, a2);
  v[1] = ff_add(a1, a3); v[3] = ff_sub(a1, a3);
}

template <typename CT> FF_HD void ff_fft8(CT* v, bool inv) {
  typedef decltype(v[0].x) FT;
  const FT h = (FT)0.70710678118654752440;
  CT a0 = ff_add(v[0], v[4]), a1 = ff_sub(v[0], v[4]);
  CT a2 = ff_add(v[2], v[6]), a3 = ff_rot(ff_sub(v[2], v[6]), inv);
  CT a4 = ff_add(v[1], v[5]), a5 = ff_sub(v[1], v[5]);
  CT a6 = ff_add(v[3], v[7]), a7 = ff_rot(ff_sub(v[3], v[7]), inv);
  CT b0 = ff_add(a0, a2), b2 = ff_sub(a0, a2), b1 = ff_add(a1, a3), b3 = ff_sub(a1, a3);
  CT b4 = ff_add(a4, a6), b6 = ff_rot(ff_sub(a4, a6), inv);
  CT s = ff_add(a5, a7), d = ff_sub(a5, a7);
  // s·w8 and d·w8³ with w8 = exp(∓iπ/4)
  CT b5, b7;
  if (!inv) { b5.x = h * (s.x + s.y); b5.y = h * (s.y - s.x); b7.x = h * (d.y - d.x); b7.y = -h * (d.x + d.y); }
  else      { b5.x = h * (s.x - s.y); b5.y = h * (s.x + s.y); b7.x = -h * (d.x + d.y); b7.y = h * (d.x - d.y); }
  v[0] = ff_add(b0, b4); v[4] = ff_sub(b0, b4);
  v[1] = ff_add(b1, b5); v[5] = ff_sub(b1, b5);
  v[2] = ff_add(b2, b6); v[6] = ff_sub(b2, b6);
  v[3] = ff_add(b3, b7); v[7] = ff_sub(b3, b7);
}

// shared-memory index padding: conflict-free for the stride-8 / stride-64 scatter of the Stockham stages
FF_HD int ff_phys(int n) { return n + (n >> 3) + (n >> 6); }
FF_HD int ff_line_pitch(int N) { int p = N + (N >> 3) + (N >> 6) + 1; return (p & 1) ? p : p + 1; }

struct FFPlan { int N, nstages, radix[6]; };

static inline FFPlan ff_make_plan(int N) {
  FFPlan p; p.N = N; p.nstages = 0;
  int rem = N;
  while (rem > 1) {
    int r = (rem % 8 == 0) ? 8 : (rem % 4 == 0 ? 4 : 2);
    p.radix[p.nstages++] = r; rem /= r;
  }
  return p;
}

// One Stockham stage for the 8 points a thread owns: thread `t` of `tpl` (= N/8, or 1 when N < 8) threads per line.
// Phase A (ff_stage_load) gathers and transforms into registers; after a barrier, phase B (ff_stage_store) scatters.
template <typename CT, int R>
FF_HD void ff_stage_load(const CT* X, int N, int Ns, int t, int tpl, const CT* tw, bool inv, CT* v) {
  const int nb = N / R;                      // butterflies per line; tpl == N/8 so each thread runs 8/R of them
  constexpr int PER = 8 / R;
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    int j = t + q * tpl;
    CT* w = v + q * R;
    int k = j & (Ns - 1);
    int tstep = k * (N / (Ns * R));
#pragma unroll
    for (int r = 0; r < R; ++r) {
      CT x = X[ff_phys(j + r * nb)];
      if (r > 0 && k > 0) { CT c = tw[r * tstep]; if (inv) c.y = -c.y; x = ff_mul(x, c); }
      w[r] = x;
    }
    if (R == 8) ff_fft8(w, inv); else if (R == 4) ff_fft4(w, inv); else ff_fft2(w);
  }
}
template <typename CT, int R>
FF_HD void ff_stage_store(CT* X, int N, int Ns, int t, int tpl, const CT* v) {
  constexpr int PER = 8 / R;
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    int j = t + q * tpl;
    int k = j & (Ns - 1);
    int j0 = (j - k) * R + k;
#pragma unroll
    for (int r = 0; r < R; ++r) X[ff_phys(j0 + r * Ns)] = v[q * R + r];
  }
}

#ifdef __CUDACC__
// Transform every line staged in shared memory (line l at X + l*pitch); all threads of the CTA must call this.
// `active` = this thread maps to a real line; tpl = threads per line.
template <typename CT>
__device__ __forceinline__ void ff_block_fft(CT* Xl, const FFPlan& P, int t, int tpl, const CT* tw, bool inv, bool active) {
  CT v[8];
  int Ns = 1;
  for (int s = 0; s < P.nstages; ++s) {
    const int R = P.radix[s];
    if (active) {
      if (R == 8) ff_stage_load<CT, 8>(Xl, P.N, Ns, t, tpl, tw, inv, v);
      else if (R == 4) ff_stage_load<CT, 4>(Xl, P.N, Ns, t, tpl, tw, inv, v);
      else ff_stage_load<CT, 2>(Xl, P.N, Ns, t, tpl, tw, inv, v);
    }
    __syncthreads();
    if (active) {
      if (R == 8) ff_stage_store<CT, 8>(Xl, P.N, Ns, t, tpl, v);
      else if (R == 4) ff_stage_store<CT, 4>(Xl, P.N, Ns, t, tpl, v);
      else ff_stage_store<CT, 2>(Xl, P.N, Ns, t, tpl, v);
    }
    __syncthreads();
    Ns *= R;
  }
}
#endif

// ---- compile-time-size variants: every index, radix and stride is a constant, direction is a template flag ---------
// Radix schedule of an N-point line (product = N); 8 points per thread per stage.
template <int N> struct FFRad;
template <> struct FFRad<16>   { static constexpr int n = 2, r0 = 8, r1 = 2, r2 = 1, r3 = 1; };
template <> struct FFRad<32>   { static constexpr int n = 2, r0 = 8, r1 = 4, r2 = 1, r3 = 1; };
template <> struct FFRad<64>   { static constexpr int n = 2, r0 = 8, r1 = 8, r2 = 1, r3 = 1; };
template <> struct FFRad<128>  { static constexpr int n = 3, r0 = 8, r1 = 4, r2 = 4, r3 = 1; };
template <> struct FFRad<256>  { static constexpr int n = 3, r0 = 8, r1 = 8, r2 = 4, r3 = 1; };
template <> struct FFRad<512>  { static constexpr int n = 3, r0 = 8, r1 = 8, r2 = 8, r3 = 1; };
template <> struct FFRad<1024> { static constexpr int n = 4, r0 = 8, r1 = 8, r2 = 4, r3 = 4; };

// padding for the lines-fastest thread mapping (8 lanes = 8 lines at the same point): one extra slot per 8 points
FF_HD int ff_phys8(int n) { return n + (n >> 3); }
template <int N> struct FFPitch8 { static constexpr int value = ((N + N / 8 + 1) | 1); };

// twiddle + butterflies on the 8 register-resident points of thread t: v[q*R + r] holds input j + r*N/R, j = t + q*N/8
template <typename CT, int N, int R, int Ns, bool INV>
FF_HD void ff_regs_stage(CT* v, int t, const CT* __restrict__ tw) {
  constexpr int PER = 8 / R;
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    CT* w = v + q * R;
    if (Ns > 1) {
      const int j = t + q * (N / 8);
      const int k = j & (Ns - 1);
      const int tstep = k * (N / (Ns * R));
#pragma unroll
      for (int r = 1; r < R; ++r) {
        CT c = tw[r * tstep];
        if (INV) c.y = -c.y;
        w[r] = ff_mul(w[r], c);
      }
    }
    if (R == 8) ff_fft8(w, INV); else if (R == 4) ff_fft4(w, INV); else ff_fft2(w);
  }
}
// scatter the stage's outputs (natural Stockham position j0 + r*Ns) into the padded shared line
// PAD = 0: lines-fastest thread mapping (ff_phys8); PAD = 1: points-fastest mapping (ff_phys)
template <int PAD> FF_HD int ff_pad(int n) { return PAD ? ff_phys(n) : ff_phys8(n); }

template <typename CT, int N, int R, int Ns, int PAD = 0>
FF_HD void ff_regs_scatter(CT* Xl, const CT* v, int t) {
  constexpr int PER = 8 / R;
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    const int j = t + q * (N / 8);
    const int k = j & (Ns - 1);
    const int j0 = (j - k) * R + k;
#pragma unroll
    for (int r = 0; r < R; ++r) Xl[ff_pad<PAD>(j0 + r * Ns)] = v[q * R + r];
  }
}
// gather the next stage's inputs j + r*N/R
template <typename CT, int N, int R, int PAD = 0>
FF_HD void ff_regs_gather(const CT* Xl, CT* v, int t) {
  constexpr int PER = 8 / R;
#pragma unroll
  for (int q = 0; q < PER; ++q) {
    const int j = t + q * (N / 8);
#pragma unroll
    for (int r = 0; r < R; ++r) v[q * R + r] = Xl[ff_pad<PAD>(j + r * (N / R))];
  }
}

#ifdef __CUDACC__
// Full N-point transform of the line owned by this thread group.  On entry v[m] = x[t + m*N/8] (m = 0..7); on exit
// v[m] = X[t + m*N/8] as well (natural order), so a second transform can start from the registers directly.
// Xl: this line's padded shared-memory scratch.  Every thread of the CTA must call (barriers inside).
template <typename CT, int N, bool INV, int PAD = 0>
__device__ __forceinline__ void ff_line_fft_regs(CT* v, CT* Xl, int t, const CT* __restrict__ tw) {
  typedef FFRad<N> RD;
  ff_regs_stage<CT, N, RD::r0, 1, INV>(v, t, tw);
  ff_regs_scatter<CT, N, RD::r0, 1, PAD>(Xl, v, t);
  __syncthreads();
  ff_regs_gather<CT, N, RD::r1, PAD>(Xl, v, t);
  ff_regs_stage<CT, N, RD::r1, RD::r0, INV>(v, t, tw);
  if (RD::n >= 3) {
    __syncthreads();
    ff_regs_scatter<CT, N, RD::r1, RD::r0, PAD>(Xl, v, t);
    __syncthreads();
    ff_regs_gather<CT, N, RD::r2, PAD>(Xl, v, t);
    ff_regs_stage<CT, N, RD::r2, RD::r0 * RD::r1, INV>(v, t, tw);
  }
  if (RD::n >= 4) {
    __syncthreads();
    ff_regs_scatter<CT, N, RD::r2, RD::r0 * RD::r1, PAD>(Xl, v, t);
    __syncthreads();
    ff_regs_gather<CT, N, RD::r3, PAD>(Xl, v, t);
    ff_regs_stage<CT, N, RD::r3, RD::r0 * RD::r1 * RD::r2, INV>(v, t, tw);
  }
  // last stage: output (q, r) sits at t + q*N/8 + r*N/RL = t + (q + r*(8/RL))*N/8 → reorder registers to m = q + r*PER
  constexpr int RL = RD::n == 2 ? RD::r1 : (RD::n == 3 ? RD::r2 : RD::r3);
  constexpr int PER = 8 / RL;
  if (PER > 1) {
    CT w[8];
#pragma unroll
    for (int q = 0; q < PER; ++q)
#pragma unroll
      for (int r = 0; r < RL; ++r) w[q + r * PER] = v[q * RL + r];
#pragma unroll
    for (int m = 0; m < 8; ++m) v[m] = w[m];
  }
  __syncthreads();   // the scratch line may be reused by the caller
}
#endif
