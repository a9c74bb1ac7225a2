// tendency_fast.cuh — fused momentum tendencies + RK3/AB2 substep for the headline configuration: triply periodic,
// regular spacing, WENO5 in all directions (BASELINE.json configs[1]; optional tracers come through the complete path).
//
// One CTA owns a TX×TY×TZ tile of cells.  For each velocity component a, the CTA evaluates every face flux
// F_{b,a} (b = x,y,z) of the tile exactly ONCE into shared memory (the reference evaluates each twice, once per adjacent
// cell, in three separate kernels), then each thread differences the staged fluxes for its cells, finishes
// Gⁿ = -div(𝐯u_a) and applies U* = U + Δt(γGⁿ + ζG⁻) into the ping-pong state buffer in the same pass, so the step's
// traffic is: read u,v,w (+G⁻), write Gⁿ and U* (SURVEY.md §8d phase A: 9w / 12w bytes per cell).
// Arithmetic follows the reference's expression order (recon.cuh), so results are bit-identical to the complete path.
#pragma once
#include "recon.cuh"

#define TF_TX 32
#ifndef TF_TY
#define TF_TY 8
#endif
#ifndef TF_TZ
#define TF_TZ 8
#endif
#define TF_THREADS 256
#ifndef TF_MINBLOCKS
#define TF_MINBLOCKS 4
#endif
#ifndef TF_UNROLL
#define TF_UNROLL 1
#endif
#ifndef TF_ROTATE
#define TF_ROTATE 1      // marching loops written out for the six rotations of the stencil window (no register moves)
#endif
#ifndef TF_ROTATE_EXACT
#define TF_ROTATE_EXACT 0   // the exact build's loop body is ~3× larger (checked divisions): six copies overflow the i-cache
#endif
#ifndef TF_PREFETCH
#define TF_PREFETCH 0    // L2 prefetch of the tiles the next phase reads: measured slower (1.27 vs 1.16 ms at 256³), kept off
#endif
#ifndef TF_XWINDOW
#define TF_XWINDOW 0     // x-direction fluxes load the six-point stencil window unconditionally and select the upwind five in
                         // registers (like the marched lines) instead of loading five values from a sign-dependent address:
                         // one more load (three fewer when a == x), but no load that waits for the advecting velocity
#endif
#ifndef TF_PIPELINE
#define TF_PIPELINE 0    // marching loops load the next face's stencil value and advecting velocities right after the current
                         // face's upwind selection (into the registers that selection just freed) instead of at the top of
                         // the next step: the loads fly during the WENO arithmetic
#endif
#ifndef TF_DIVUNROLL
#define TF_DIVUNROLL 4   // unroll factor of the divergence / substep loop over the tile's planes (loads of U and G⁻ of the next
                         // planes can then be issued ahead of the current plane's stores)
#endif
constexpr int tf_unroll = TF_UNROLL;
constexpr int tf_divunroll = TF_DIVUNROLL;
#define TF_SMEM_ELEMS (TF_TZ * TF_TY * (TF_TX + 1) + TF_TZ * (TF_TY + 1) * TF_TX + (TF_TZ + 1) * TF_TY * TF_TX)

template <typename FT>
struct FastTendJob {
  GridDev<FT> g;
  const FT* U[3];      // state with filled halos
  FT* Unew[3];         // ping-pong target (interior written)
  FT* Gn[3];
  const FT* Gm[3];
  SchemeTables<FT> tab;
  FT dt, gamma, zeta;
  int mode;            // 0: U* = U + Δt·γ·Gⁿ, 1: U* = U + Δt(γGⁿ + ζG⁻), 2: AB2, 3: AB2-Euler, -1: tendencies only
  // y-slab runs with peer-to-peer halos: row j = 1 of v* is also stored into row Ny+1 of the previous rank's v* (NVLink
  // peer pointer), which is all the divergence of that rank needs from here; nullptr otherwise
  FT* v_prev = nullptr;
  // tile rows (in y) this launch covers: row t of the launch is tile row t for t < ty_split, t + ty_gap otherwise; ny_tiles = how
  // many.  Whole grid: ty_split = ny_tiles, ty_gap = 0.  A y-slab run launches the interior tile rows (those whose stencils stay
  // inside the slab's own rows) before the barrier that makes the neighbours' halo rows visible, and the rest after it
  // (interior / boundary overlap, cf. src/Models/interleave_communication_and_computation.jl:9-67).
  int ty_split = 1 << 30, ty_gap = 0, ny_tiles = 0;
};
template <typename FT>
__device__ __forceinline__ int tf_tile_row(const FastTendJob<FT>& J) { return (int)blockIdx.y + ((int)blockIdx.y >= J.ty_split ? J.ty_gap : 0); }

// face flux F_{b,a} at flux-location index n (linear), strides sa (dim a) and sb (dim b)
template <typename FT, bool SAME, int VARIANT = 0>
__device__ __forceinline__ FT fast_flux(const SchemeTables<FT>& T, const FT* __restrict__ ua, const FT* __restrict__ ub,
                                        int n, int sa, int sb, FT Ab) {
  // advecting velocity: Centered(4) interpolation of A_b·U_b along dim b onto a Centre (SAME) or along dim a onto a Face
  const int ss = SAME ? sb : sa;
  const int s0 = SAME ? 1 : 0;
  FT ut;
  if (VARIANT == 0) {
    ut = T.sym[2][3] * (Ab * ub[n + (-2 + s0) * ss]);
    ut = ut + T.sym[2][2] * (Ab * ub[n + (-1 + s0) * ss]);
    ut = ut + T.sym[2][1] * (Ab * ub[n + (0 + s0) * ss]);
    ut = ut + T.sym[2][0] * (Ab * ub[n + (1 + s0) * ss]);
  } else {
    // contracted build: the constant face area factored out of the four-point sum
    ut = T.sym[2][3] * ub[n + (-2 + s0) * ss] + T.sym[2][2] * ub[n + (-1 + s0) * ss];
    ut = (ut + T.sym[2][1] * ub[n + (0 + s0) * ss]) + T.sym[2][0] * ub[n + (1 + s0) * ss];
    ut = Ab * ut;
  }
  const bool left = ut > FT(0);
  // upwind stencil without a branch: left bias reads offsets -3…1, right bias 2…-2 (reversed), from the same pointer
  const int ds = left ? sb : -sb;
  const FT* p = ua + n + s0 * sb + (left ? -3 * sb : 2 * sb);
  FT q0 = p[0], q1 = p[ds], q2 = p[2 * ds], q3 = p[3 * ds], q4 = p[4 * ds];
  if (VARIANT == 0) return ut * weno5_from_q<FT>(T, q0, q1, q2, q3, q4);
  return ut * weno5_contracted<FT>(T, q0, q1, q2, q3, q4);
}

// One face flux of a marched line: the stencil window S0…S5 = U_a at offsets s0-3 … s0+2 (in units of the flux direction's
// stride) around the flux location; left bias uses S0…S4, right bias S5…S1.  Same arithmetic as fast_flux.
template <typename FT, int VARIANT, bool SAME>
__device__ __forceinline__ FT march_one(const SchemeTables<FT>& T, FT S0, FT S1, FT S2, FT S3, FT S4, FT S5,
                                        const FT* __restrict__ pb, int sa, FT Ab) {
  FT u0, u1, u2, u3;
  if (SAME) { u0 = S1; u1 = S2; u2 = S3; u3 = S4; }     // a == b: the advecting velocity reads the same window
  else { u0 = pb[0]; u1 = pb[sa]; u2 = pb[2 * sa]; u3 = pb[3 * sa]; }
  FT ut;
  if (VARIANT == 0) {
    ut = T.sym[2][3] * (Ab * u0);
    ut = ut + T.sym[2][2] * (Ab * u1);
    ut = ut + T.sym[2][1] * (Ab * u2);
    ut = ut + T.sym[2][0] * (Ab * u3);
  } else {
    ut = T.sym[2][3] * u0 + T.sym[2][2] * u1;
    ut = (ut + T.sym[2][1] * u2) + T.sym[2][0] * u3;
    ut = Ab * ut;
  }
  const bool left = ut > FT(0);
  const FT q0 = left ? S0 : S5, q1 = left ? S1 : S4, q2 = left ? S2 : S3, q3 = left ? S3 : S2, q4 = left ? S4 : S1;
  return ut * (VARIANT == 0 ? weno5_from_q<FT>(T, q0, q1, q2, q3, q4) : weno5_contracted<FT>(T, q0, q1, q2, q3, q4));
}

// march_one with the advecting-velocity inputs passed in registers (pipelined loops)
template <typename FT, int VARIANT, bool SAME>
__device__ __forceinline__ FT march_ut(const SchemeTables<FT>& T, FT S1, FT S2, FT S3, FT S4, FT v0, FT v1, FT v2, FT v3, FT Ab) {
  FT u0, u1, u2, u3;
  if (SAME) { u0 = S1; u1 = S2; u2 = S3; u3 = S4; }
  else { u0 = v0; u1 = v1; u2 = v2; u3 = v3; }
  FT ut;
  if (VARIANT == 0) {
    ut = T.sym[2][3] * (Ab * u0);
    ut = ut + T.sym[2][2] * (Ab * u1);
    ut = ut + T.sym[2][1] * (Ab * u2);
    ut = ut + T.sym[2][0] * (Ab * u3);
  } else {
    ut = T.sym[2][3] * u0 + T.sym[2][2] * u1;
    ut = (ut + T.sym[2][1] * u2) + T.sym[2][0] * u3;
    ut = Ab * ut;
  }
  return ut;
}

// `count` consecutive face fluxes F_{b,a} along direction b (stride sb), starting at flux location n, with the stencil
// window carried in registers from one face to the next.  The loop body is written out for the six rotations of the
// window's register names, so advancing one face costs one load and no register moves.
template <typename FT, int VARIANT, bool SAME>
__device__ __forceinline__ void march_fluxes(const SchemeTables<FT>& T, const FT* __restrict__ ua, const FT* __restrict__ ub,
                                             int n, int sb, int sa, FT Ab, int count, FT* out, int ostride) {
  static_assert(TF_THREADS == TF_TX * TF_TY && TF_THREADS == TF_TX * TF_TZ, "one thread per line of faces");
  const FT* p = ua + (n + ((SAME ? 1 : 0) - 3) * sb);
  FT S0 = p[0], S1 = p[sb], S2 = p[2 * sb], S3 = p[3 * sb], S4 = p[4 * sb], S5;
  p += 5 * sb;
  const FT* pb = ub + (n - 2 * sa);
#if TF_PIPELINE
  if ((VARIANT == 0 ? TF_ROTATE_EXACT : TF_ROTATE) != 0) {
    // software-pipelined: window value S5 and the four advecting-velocity inputs of step c are loaded during step c-1
    FT v0 = FT(0), v1 = FT(0), v2 = FT(0), v3 = FT(0);
    S5 = *p;
    if (!SAME) { v0 = pb[0]; v1 = pb[sa]; v2 = pb[2 * sa]; v3 = pb[3 * sa]; }
#define TF_PIPE_STEP(A, B, C, D, E, F, LAST)                                                         \
    {                                                                                                \
      const FT ut = march_ut<FT, VARIANT, SAME>(T, B, C, D, E, v0, v1, v2, v3, Ab);                   \
      const bool left = ut > FT(0);                                                                  \
      const FT q0 = left ? A : F, q1 = left ? B : E, q2 = left ? C : D, q3 = left ? D : C, q4 = left ? E : B; \
      if (!(LAST)) {                                                                                 \
        p += sb; pb += sb;                                                                           \
        A = *p;                                                                                      \
        if (!SAME) { v0 = pb[0]; v1 = pb[sa]; v2 = pb[2 * sa]; v3 = pb[3 * sa]; }                    \
      }                                                                                              \
      *out = ut * (VARIANT == 0 ? weno5_from_q<FT>(T, q0, q1, q2, q3, q4) : weno5_contracted<FT>(T, q0, q1, q2, q3, q4)); \
      out += ostride;                                                                                \
    }
#pragma unroll 1
    for (int c = count; c > 0; c -= 6) {
      TF_PIPE_STEP(S0, S1, S2, S3, S4, S5, c == 1) if (c == 1) break;
      TF_PIPE_STEP(S1, S2, S3, S4, S5, S0, c == 2) if (c == 2) break;
      TF_PIPE_STEP(S2, S3, S4, S5, S0, S1, c == 3) if (c == 3) break;
      TF_PIPE_STEP(S3, S4, S5, S0, S1, S2, c == 4) if (c == 4) break;
      TF_PIPE_STEP(S4, S5, S0, S1, S2, S3, c == 5) if (c == 5) break;
      TF_PIPE_STEP(S5, S0, S1, S2, S3, S4, c == 6)
    }
#undef TF_PIPE_STEP
    return;
  }
#endif
#define TF_MARCH_STEP(A, B, C, D, E, F)                                                \
  F = *p;                                                                              \
  *out = march_one<FT, VARIANT, SAME>(T, A, B, C, D, E, F, pb, sa, Ab);                \
  out += ostride; p += sb; pb += sb;
  if ((VARIANT == 0 ? TF_ROTATE_EXACT : TF_ROTATE) != 0) {
#pragma unroll 1
    for (int c = count; c > 0; c -= 6) {
      TF_MARCH_STEP(S0, S1, S2, S3, S4, S5) if (c == 1) break;
      TF_MARCH_STEP(S1, S2, S3, S4, S5, S0) if (c == 2) break;
      TF_MARCH_STEP(S2, S3, S4, S5, S0, S1) if (c == 3) break;
      TF_MARCH_STEP(S3, S4, S5, S0, S1, S2) if (c == 4) break;
      TF_MARCH_STEP(S4, S5, S0, S1, S2, S3) if (c == 5) break;
      TF_MARCH_STEP(S5, S0, S1, S2, S3, S4)
    }
  } else {
#pragma unroll 1
    for (int c = 0; c < count; ++c) {
      TF_MARCH_STEP(S0, S1, S2, S3, S4, S5)
      S0 = S1; S1 = S2; S2 = S3; S3 = S4; S4 = S5;
    }
  }
#undef TF_MARCH_STEP
}

// prefetch the 128-byte lines of rows [j0-hy, j0+ny+hy) × [k0-hz, k0+nz+hz), x from i0-hx to i0+nx+hx, of one field into L2
template <typename FT>
__device__ __forceinline__ void tf_prefetch_tile(const GridDev<FT>& g, const FT* f, int i0, int j0, int k0, int nx, int ny,
                                                 int nz, int hx, int hy, int hz, int tid) {
  const int rows_y = ny + 2 * hy, rows = rows_y * (nz + 2 * hz);
  const int first = g.idx32(i0 - hx, j0 - hy, k0 - hz);
  const int lines = ((nx + 2 * hx) * (int)sizeof(FT) + 127) / 128 + 1;   // a row may straddle one more line
  for (int t = tid; t < rows * lines; t += TF_THREADS) {
    const int r = t / lines, l = t - r * lines;
    const int ry = r % rows_y, rz = r / rows_y;
    const FT* q = f + first + ry * (int)g.sy + rz * (int)g.sz;
    const char* c = reinterpret_cast<const char*>(q) + l * 128;
    if (c < reinterpret_cast<const char*>(q + nx + 2 * hx)) asm volatile("prefetch.global.L2 [%0];" ::"l"(c));
  }
}

// VARIANT only gives the two builds of this kernel different symbols: 0 is compiled with -fmad=false like the rest of the
// library (bit-identical to the reference's CPU arithmetic), 1 lives in tendency_fma.cu, compiled with FMA contraction
// (nh_model_desc.exact_arithmetic = 0: ≈ ⅓ fewer FP64 instructions, results within rounding of the exact build).
template <typename FT, int VARIANT = 0>
__global__ void __launch_bounds__(TF_THREADS, TF_MINBLOCKS) tendency_fast_kernel(const __grid_constant__ FastTendJob<FT> J) {
  extern __shared__ __align__(16) unsigned char tf_smem[];
  typedef FT (*FxT)[TF_TY][TF_TX + 1];
  typedef FT (*FyT)[TF_TY + 1][TF_TX];
  typedef FT (*FzT)[TF_TY][TF_TX];
  FxT Fx = reinterpret_cast<FxT>(tf_smem);
  FyT Fy = reinterpret_cast<FyT>(tf_smem + sizeof(FT) * TF_TZ * TF_TY * (TF_TX + 1));
  FzT Fz = reinterpret_cast<FzT>(tf_smem + sizeof(FT) * (TF_TZ * TF_TY * (TF_TX + 1) + TF_TZ * (TF_TY + 1) * TF_TX));
  const GridDev<FT>& g = J.g;
  const int i0 = blockIdx.x * TF_TX + 1, j0 = tf_tile_row(J) * TF_TY + 1, k0 = blockIdx.z * TF_TZ + 1;
  const int tid = threadIdx.x;
  const int st[3] = {1, (int)g.sy, (int)g.sz};   // 32-bit element offsets: the parent volume is < 2^31 elements
  const FT Ax = g.dy * g.dz, Ay = g.dx * g.dz, Az = g.dx * g.dy;
  const FT Vinv = FT(1) / ((g.dx * g.dy) * g.dz);
  const int nx = min(TF_TX, g.N[0] - i0 + 1), ny = min(TF_TY, g.N[1] - j0 + 1), nz = min(TF_TZ, g.N[2] - k0 + 1);

  const int mode = J.mode;
  const FT dt = J.dt, gamma = J.gamma, zeta = J.zeta;
#if TF_PREFETCH
  // the advecting velocities of the first component's y- and z-fluxes: on their way to L2 while the x-fluxes are computed
  tf_prefetch_tile(g, J.U[1], i0, j0, k0, nx, ny, nz, 2, 3, 0, tid);
  tf_prefetch_tile(g, J.U[2], i0, j0, k0, nx, ny, nz, 2, 0, 3, tid);
#endif

#pragma unroll 1
  for (int a = 0; a < 3; ++a) {
    const FT* __restrict__ ua = J.U[a];
    const int sa = st[a];
#if TF_PREFETCH
    if (mode == 1 || mode == 2) tf_prefetch_tile(g, J.Gm[a], i0, j0, k0, nx, ny, nz, 0, 0, 0, tid);
    if (a < 2) tf_prefetch_tile(g, J.U[a + 1], i0, j0, k0, nx, ny, nz, 3, 3, 3, tid);
#endif
    // ---- x-direction fluxes F_{x,a}: a == 0 → centres i0-1 … i0+nx-1, else faces i0 … i0+nx
    {
      // flux t of the tile is (li, row) = (t mod (TX+1), t div (TX+1)); advanced incrementally by TF_THREADS per pass
      constexpr int DL = TF_THREADS % (TF_TX + 1), DR = TF_THREADS / (TF_TX + 1);
      int li = tid % (TF_TX + 1), row = tid / (TF_TX + 1);
#pragma unroll tf_unroll
      for (int t = tid; t < (TF_TX + 1) * TF_TY * TF_TZ; t += TF_THREADS) {
        const int lj = row % TF_TY, lk = row / TF_TY;
        // partial tiles: clamp to the last valid point (the value is computed but never used)
        const int n = g.idx32(i0 + min(li, nx) - (a == 0 ? 1 : 0), j0 + min(lj, ny - 1), k0 + min(lk, nz - 1));
#if TF_XWINDOW
        {
          const FT* w = ua + n + (a == 0 ? -2 : -3);
          const FT S0 = w[0], S1 = w[1], S2 = w[2], S3 = w[3], S4 = w[4], S5 = w[5];
          Fx[lk][lj][li] = a == 0 ? march_one<FT, VARIANT, true>(J.tab, S0, S1, S2, S3, S4, S5, nullptr, sa, Ax)
                                  : march_one<FT, VARIANT, false>(J.tab, S0, S1, S2, S3, S4, S5, J.U[0] + n - 2 * sa, sa, Ax);
        }
#else
        Fx[lk][lj][li] = a == 0 ? fast_flux<FT, true, VARIANT>(J.tab, ua, J.U[0], n, sa, 1, Ax)
                                : fast_flux<FT, false, VARIANT>(J.tab, ua, J.U[0], n, sa, 1, Ax);
#endif
        li += DL; row += DR;
        if (li > TF_TX) { li -= TF_TX + 1; ++row; }
      }
    }
    // ---- y- and z-direction fluxes: one thread per line of faces, marching along the flux direction with the six
    // stencil values in registers (each row / plane of U_a is loaded once per line instead of five times)
    {
      const int li = tid & (TF_TX - 1), lo = tid / TF_TX, ci = i0 + min(li, nx - 1);
      // one copy of the marching code serves both directions (instruction-cache footprint)
#pragma unroll 1
      for (int b = 1; b <= 2; ++b) {
        const bool y = b == 1;
        const int sb = y ? st[1] : st[2], count = (y ? ny : nz) + 1, ostride = y ? TF_TX : TF_TY * TF_TX;
        const int n0 = y ? g.idx32(ci, j0, k0 + min(lo, nz - 1)) : g.idx32(ci, j0 + min(lo, ny - 1), k0);
        FT* out = y ? &Fy[lo][0][li] : &Fz[0][lo][li];
        const FT Ab = y ? Ay : Az;
        if (a == b) march_fluxes<FT, VARIANT, true>(J.tab, ua, ua, n0 - sb, sb, sa, Ab, count, out, ostride);
        else march_fluxes<FT, VARIANT, false>(J.tab, ua, y ? J.U[1] : J.U[2], n0, sb, sa, Ab, count, out, ostride);
      }
    }
    __syncthreads();
    // ---- divergence, tendency, substep: thread ↔ (li, lj) column of the tile, marching in k
    {
      const int li = tid & (TF_TX - 1), lj = tid / TF_TX;
      if (li < nx && lj < ny) {
        int n = g.idx32(i0 + li, j0 + lj, k0);
        const FT* __restrict__ gm = J.Gm[a];
        FT* __restrict__ gn = J.Gn[a];
        FT* __restrict__ un = J.Unew[a];
        FT* vp = (a == 1 && j0 + lj == 1) ? J.v_prev : nullptr;
#pragma unroll tf_divunroll
        for (int lk = 0; lk < nz; ++lk, n += st[2]) {
          FT total = (Fx[lk][lj][li + 1] - Fx[lk][lj][li]) + (Fy[lk][lj + 1][li] - Fy[lk][lj][li]);
          total = total + (Fz[lk + 1][lj][li] - Fz[lk][lj][li]);
          FT G = -(Vinv * total);
          G = G + FT(0);
          gn[n] = G;
          if (mode >= 0) {
            FT u = ua[n];
            if (mode == 0) u = u + dt * gamma * G;
            else if (mode == 1) u = u + dt * (gamma * G + zeta * gm[n]);
            else if (mode == 2) u = u + dt * ((FT(1.5) + gamma) * G - (FT(0.5) + gamma) * gm[n]);
            else u = u + dt * ((FT(1.5) + gamma) * G - FT(0));
            un[n] = u;
            if (vp != nullptr) vp[n + g.N[1] * st[1]] = u;
          }
        }
      }
    }
    __syncthreads();
  }
}

// host launch of the fused tendency + substep sweep; the contracted build is defined in tendency_fma.cu
template <typename FT> void launch_fused_tendency_fma(const FastTendJob<FT>& J, cudaStream_t stream);

template <typename FT>
inline void launch_fused_tendency(const FastTendJob<FT>& J, cudaStream_t stream, bool exact = true) {
  if (!exact) { launch_fused_tendency_fma<FT>(J, stream); return; }
  const GridDev<FT>& g = J.g;
  dim3 grd(ceil_div(g.N[0], TF_TX), J.ny_tiles > 0 ? J.ny_tiles : ceil_div(g.N[1], TF_TY), ceil_div(g.N[2], TF_TZ));
  tendency_fast_kernel<FT, 0><<<grd, TF_THREADS, TF_SMEM_ELEMS * sizeof(FT), stream>>>(J);
}

// Periodic halo fill of several fields in ONE launch: every halo cell reads its wrapped interior image directly
// (equivalent to the reference's x → y → z sequence of fill_periodic_*_halo! kernels,
// src/BoundaryConditions/fill_halo_regions_periodic.jl:40-111, corners included).
template <typename FT>
struct FastHaloJob { GridDev<FT> g; FT* p[NH_MAXTR + 3]; int nf; };

template <typename FT>
__global__ void __launch_bounds__(256) periodic_halo_all_kernel(const __grid_constant__ FastHaloJob<FT> J) {
  const GridDev<FT>& g = J.g;
  const int Ex = g.N[0] + 2 * g.H[0], Ey = g.N[1] + 2 * g.H[1], Ez = g.N[2] + 2 * g.H[2];
  // enumerate the parent box plane by plane; skip interior cells quickly (only shells do work)
  int pj = blockIdx.y, pk = blockIdx.z;
  int j = pj + 1 - g.H[1], k = pk + 1 - g.H[2];
  bool row_interior = (j >= 1 && j <= g.N[1] && k >= 1 && k <= g.N[2]);
  int wj = j < 1 ? j + g.N[1] : (j > g.N[1] ? j - g.N[1] : j);
  int wk = k < 1 ? k + g.N[2] : (k > g.N[2] ? k - g.N[2] : k);
  (void)Ey; (void)Ez;
  if (row_interior) {
    // only the 2·Hx cells at the two ends of this row
    int t = threadIdx.x;
    if (t >= 2 * g.H[0]) return;
    int i = t < g.H[0] ? t + 1 - g.H[0] : g.N[0] + 1 + (t - g.H[0]);
    int wi = i < 1 ? i + g.N[0] : i - g.N[0];
    long long dst = g.idx(i, j, k), src = g.idx(wi, j, k);
    for (int f = 0; f < J.nf; ++f) J.p[f][dst] = J.p[f][src];
    return;
  }
  for (int pi = threadIdx.x; pi < Ex; pi += blockDim.x) {
    int i = pi + 1 - g.H[0];
    int wi = i < 1 ? i + g.N[0] : (i > g.N[0] ? i - g.N[0] : i);
    long long dst = g.idx(i, j, k), src = g.idx(wi, wj, wk);
    for (int f = 0; f < J.nf; ++f) J.p[f][dst] = J.p[f][src];
  }
}

// Pressure correction over the interior with periodic index wrap on p (no p halo needed), fused with the periodic halo
// fill of the corrected velocities: a corrected interior value is also stored to each of its periodic images.
// _pressure_correct_velocities! — src/Models/NonhydrostaticModels/pressure_correction.jl:31-37
// Each thread corrects two x-adjacent cells (16-byte loads/stores; i = 1 starts a 128-byte line, pitches are multiples of 16).
template <typename FT>
struct YPeers { FT* prev[3] = {nullptr, nullptr, nullptr}; FT* next[3] = {nullptr, nullptr, nullptr}; };   // peers' u, v, w (y-slab runs)

template <typename FT>
__global__ void __launch_bounds__(256) pressure_correct_halo_kernel(GridDev<FT> g, FT* __restrict__ u, FT* __restrict__ v,
                                                                    FT* __restrict__ w, const FT* __restrict__ p, double dt, int ywrap,
                                                                    YPeers<FT> peers) {
  typedef typename cx_of<FT>::type V2;
  const int i = 2 * (blockIdx.x * blockDim.x + threadIdx.x) + 1;
  const int j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  const int k = blockIdx.z + 1;
  if (i > g.N[0] || j > g.N[1] || k > g.N[2]) return;
  const bool two = i + 1 <= g.N[0];
  const long long n = g.idx(i, j, k);
  const long long nym = g.idx(i, (ywrap && j == 1) ? g.N[1] : j - 1, k), nzm = g.idx(i, j, k == 1 ? g.N[2] : k - 1);
  FT p0[2], pym[2], pzm[2], uu[2], vv[2], ww[2];
  if (two) {
    V2 t;
    t = *reinterpret_cast<const V2*>(p + n); p0[0] = t.x; p0[1] = t.y;
    t = *reinterpret_cast<const V2*>(p + nym); pym[0] = t.x; pym[1] = t.y;
    t = *reinterpret_cast<const V2*>(p + nzm); pzm[0] = t.x; pzm[1] = t.y;
    t = *reinterpret_cast<const V2*>(u + n); uu[0] = t.x; uu[1] = t.y;
    t = *reinterpret_cast<const V2*>(v + n); vv[0] = t.x; vv[1] = t.y;
    t = *reinterpret_cast<const V2*>(w + n); ww[0] = t.x; ww[1] = t.y;
  } else {
    p0[0] = p0[1] = p[n]; pym[0] = pym[1] = p[nym]; pzm[0] = pzm[1] = p[nzm];
    uu[0] = uu[1] = u[n]; vv[0] = vv[1] = v[n]; ww[0] = ww[1] = w[n];
  }
  const FT pxm0 = p[i == 1 ? g.idx(g.N[0], j, k) : n - 1];
  FT un[2], vn[2], wn[2];
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const FT pxm = c == 0 ? pxm0 : p0[0];
    const FT gx = (p0[c] - pxm) / g.dx, gy = (p0[c] - pym[c]) / g.dy, gz = (p0[c] - pzm[c]) / g.dz;
    if (sizeof(FT) == 4) {
      un[c] = (FT)((double)uu[c] - (double)gx * dt); vn[c] = (FT)((double)vv[c] - (double)gy * dt); wn[c] = (FT)((double)ww[c] - (double)gz * dt);
    } else {
      un[c] = uu[c] - gx * (FT)dt; vn[c] = vv[c] - gy * (FT)dt; wn[c] = ww[c] - gz * (FT)dt;
    }
  }
  if (two) {
    V2 t;
    t.x = un[0]; t.y = un[1]; *reinterpret_cast<V2*>(u + n) = t;
    t.x = vn[0]; t.y = vn[1]; *reinterpret_cast<V2*>(v + n) = t;
    t.x = wn[0]; t.y = wn[1]; *reinterpret_cast<V2*>(w + n) = t;
  } else { u[n] = un[0]; v[n] = vn[0]; w[n] = wn[0]; }
  // periodic images: index sets {i, i±N} ∩ parent range, likewise j, k; only cells within a halo width of a face have any
  const bool xedge = i <= g.H[0] || i + 1 > g.N[0] - g.H[0];
  const bool remote = !ywrap && peers.prev[0] != nullptr;
  const bool yedge = (ywrap || remote) && (j <= g.H[1] || j > g.N[1] - g.H[1]);
  const bool zedge = k <= g.H[2] || k > g.N[2] - g.H[2];
  if (!(xedge || yedge || zedge)) return;
  int yj[2], nyj = 1; yj[0] = j;
  if (ywrap) { if (j <= g.H[1]) yj[nyj++] = j + g.N[1]; else if (j > g.N[1] - g.H[1]) yj[nyj++] = j - g.N[1]; }
  int zk[2], nzk = 1; zk[0] = k;
  if (k <= g.H[2]) zk[nzk++] = k + g.N[2]; else if (k > g.N[2] - g.H[2]) zk[nzk++] = k - g.N[2];
  for (int c = 0; c < (two ? 2 : 1); ++c) {
    const int ii = i + c;
    int xi[2], nxi = 1; xi[0] = ii;
    if (ii <= g.H[0]) xi[nxi++] = ii + g.N[0]; else if (ii > g.N[0] - g.H[0]) xi[nxi++] = ii - g.N[0];
    for (int cz = 0; cz < nzk; ++cz)
      for (int b = 0; b < nyj; ++b)
        for (int a = (cz == 0 && b == 0) ? 1 : 0; a < nxi; ++a) {
          const long long m = g.idx(xi[a], yj[b], zk[cz]);
          u[m] = un[c]; v[m] = vn[c]; w[m] = wn[c];
        }
    if (remote && (j <= g.H[1] || j > g.N[1] - g.H[1])) {
      // the neighbour's y halo row that mirrors this row (all x / z images of it too), written over NVLink
      const bool lo = j <= g.H[1];
      FT* const* dst = lo ? peers.prev : peers.next;
      const int jr = lo ? j + g.N[1] : j - g.N[1];
      for (int cz = 0; cz < nzk; ++cz)
        for (int a = 0; a < nxi; ++a) {
          const long long m = g.idx(xi[a], jr, zk[cz]);
          dst[0][m] = un[c]; dst[1][m] = vn[c]; dst[2][m] = wn[c];
        }
    }
  }
}

template <typename FT>
inline void launch_pressure_correct_halo(const GridDev<FT>& g, FT* u, FT* v, FT* w, const FT* p, double dt, int ywrap,
                                         cudaStream_t stream, const YPeers<FT>& peers = YPeers<FT>()) {
  dim3 blk(64, 4, 1), grd(ceil_div(ceil_div(g.N[0], 2), 64), ceil_div(g.N[1], 4), g.N[2]);
  pressure_correct_halo_kernel<FT><<<grd, blk, 0, stream>>>(g, u, v, w, p, dt, ywrap, peers);
}
