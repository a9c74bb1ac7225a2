// tendency_plane.cuh — plane-marching variant of the fused momentum tendencies + RK3/AB2 substep (same configuration and same
// arithmetic as tendency_fast.cuh: triply periodic, regular spacing, WENO5 in all directions).
//
// A CTA owns a TP_TX × TP_TY column of cells and marches along z through a segment of planes.  For every plane k the three
// velocity planes (with their 3-cell x / y halos) are brought into shared memory by TMA (cp.async.bulk.tensor.3d, one box of
// TP_BW × TP_BH × 1 per field, three stages deep, completion on an mbarrier), so
//   * every stencil offset of the x- and y-direction reconstructions is an immediate offset from one per-thread base address
//     (no per-load address arithmetic, no upwind-dependent addresses: the six-point window is read and the upwind five are
//     selected in registers),
//   * the z-direction windows of u, v, w live in registers (each thread keeps planes k-3 … k+2 of its own column; one
//     coalesced load per field and plane, issued one plane ahead),
//   * the bulk copies are asynchronous: plane k+2 is in flight while plane k is being worked on.
// Each face flux is evaluated once: per plane and component the CTA computes (TX+1)·TY x-fluxes, TX·(TY+1) y-fluxes and TX·TY
// z-fluxes.  Warps 0 … TY-1 own the rows of the tile (x-flux slot li, y-flux slot lj, z-flux of their column, the divergence
// and the substep); one extra warp evaluates the tile's east face column and north face row, so nobody runs a second pass.
// The x / y fluxes of a plane are staged in shared memory (four buffers, split-phase mbarrier instead of __syncthreads), the z-flux of the
// lower face is carried in a register.  The flux functions are those of tendency_fast.cuh / recon.cuh, so the results are
// bit-identical to tendency_fast_kernel (and, in the exact build, to the complete path).
#pragma once
#include <cuda.h>
#include "tendency_fast.cuh"

#define TP_TX 32
#ifndef TP_TY
#define TP_TY 7                  // 7 rows + the extra warp = 8 warps = 256 threads: 128 registers per thread at two CTAs per SM (8 rows +
                                 // 1 = 288 threads are allocated as 320 and capped at 96 registers, which spills: 0.98 vs 0.94 ms at 256³)
#endif
#define TP_BW 40                 // box width: columns i0-4 … i0+35 (the box must START on a 16-byte boundary of the row — i0-4 is a
                                 // multiple of four elements from the row's 128-byte aligned first interior cell — and a row must be
                                 // a multiple of 16 bytes in Float32 too)
#define TP_BH (TP_TY + 6)
#define TP_STAGES 3
#define TP_THREADS (TP_TX * (TP_TY + 1))
#ifndef TP_MINBLOCKS
#define TP_MINBLOCKS 2
#endif
#ifndef TP_NO_TMA
#define TP_NO_TMA 0            // 1: A/B variant that stages the planes with ordinary loads and stores (tools/plane_probe.cu)
#endif

template <typename FT>
struct alignas(64) PlaneJob {
  CUtensorMap tm[3];             // u, v, w: dims (px, py, pz), box (TP_BW, TP_BH, 1)
  FastTendJob<FT> J;
  int zseg;                      // planes per CTA
};

template <typename FT> struct TPSizes {
  static constexpr int slot = ((TP_BW * TP_BH * (int)sizeof(FT) + 127) / 128) * 128;        // one field plane, 128-byte aligned
  static constexpr int box_bytes = TP_BW * TP_BH * (int)sizeof(FT);
  static constexpr int planes = TP_STAGES * 3 * slot;
  static constexpr int fx_elems = 3 * TP_TY * (TP_TX + 1);                                    // per buffer
  static constexpr int fy_elems = 3 * (TP_TY + 1) * TP_TX;
  static constexpr int flux_bufs = 4;                                                         // see the hazard notes in the kernel
  static constexpr int flux_bytes = flux_bufs * (fx_elems + fy_elems) * (int)sizeof(FT);
  static constexpr int total = planes + flux_bytes + 64 + 128;                                // + mbarriers + alignment slack
};

__device__ __forceinline__ unsigned tp_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tp_mbar_init(unsigned bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void tp_mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tp_mbar_wait(unsigned bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "TP_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra TP_DONE;\n"
      "bra TP_WAIT;\n"
      "TP_DONE:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tp_tma_load_3d(unsigned dst, const CUtensorMap* tm, unsigned bar, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
               ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

// one face flux from the six-point window S0 … S5 (offsets -3 … +2 of the flux location along the flux direction, or -2 … +3
// around a Centre when the flux runs along the component's own direction) and the four advecting-velocity inputs
#ifndef TP_EXACT_CALL
#define TP_EXACT_CALL 1          // exact build: the reconstruction (≈ 250 instructions with its checked divisions) is a function call, so
                                 // that the plane step's nine evaluations share one body and the loop stays inside the instruction cache
#endif
template <typename FT>
__device__ __noinline__ FT tp_weno5_exact_call(const SchemeTables<FT>* T, FT q0, FT q1, FT q2, FT q3, FT q4) {
  return weno5_from_q<FT>(*T, q0, q1, q2, q3, q4);
}

template <typename FT, int VARIANT, bool SAME>
__device__ __forceinline__ FT tp_flux(const SchemeTables<FT>& T, FT S0, FT S1, FT S2, FT S3, FT S4, FT S5, FT v0, FT v1, FT v2,
                                      FT v3, FT Ab) {
  const FT ut = march_ut<FT, VARIANT, SAME>(T, S1, S2, S3, S4, v0, v1, v2, v3, Ab);
  const bool left = ut > FT(0);
  const FT q0 = left ? S0 : S5, q1 = left ? S1 : S4, q2 = left ? S2 : S3, q3 = left ? S3 : S2, q4 = left ? S4 : S1;
  if (VARIANT != 0) return ut * weno5_contracted<FT>(T, q0, q1, q2, q3, q4);
#if TP_EXACT_CALL
  return ut * tp_weno5_exact_call<FT>(&T, q0, q1, q2, q3, q4);
#else
  return ut * weno5_from_q<FT>(T, q0, q1, q2, q3, q4);
#endif
}

// Row warps, one plane step, component A (0, 1, 2 = u, v, w): its z-flux at face slot k (returned) and — unless this is the
// segment's closing step — its x- and y-fluxes of plane k into the staging buffer.  The uniform `do_xy` branches are deliberate:
// they keep the scheduler from interleaving all nine reconstructions of a step, which costs more in spilled registers (96 per
// thread at two CTAs per SM) than the extra instruction-level parallelism returns (measured: 1.07 → 1.15 ms at 256³).
template <typename FT, int VARIANT, int A>
__device__ __forceinline__ FT tp_row_fluxes(const SchemeTables<FT>& T, const FT* __restrict__ Pu, const FT* __restrict__ Pv,
                                            const FT* __restrict__ Pw, int tb, bool do_xy, const FT (&Su)[6], const FT (&Sv)[6],
                                            const FT (&Sw)[6], FT* __restrict__ cur, int fxo, int fyo, FT Ax, FT Ay, FT Az) {
  constexpr int BW = TP_BW, FXC = TP_TY * (TP_TX + 1), FYC = (TP_TY + 1) * TP_TX, FYB = 3 * FXC;
  const FT* __restrict__ Pa = A == 0 ? Pu : (A == 1 ? Pv : Pw);
  FT fz;
  if (A == 0) fz = tp_flux<FT, VARIANT, false>(T, Su[0], Su[1], Su[2], Su[3], Su[4], Su[5], Pw[tb - 2], Pw[tb - 1], Pw[tb], Pw[tb + 1], Az);
  else if (A == 1) fz = tp_flux<FT, VARIANT, false>(T, Sv[0], Sv[1], Sv[2], Sv[3], Sv[4], Sv[5], Pw[tb - 2 * BW], Pw[tb - BW], Pw[tb], Pw[tb + BW], Az);
  else fz = tp_flux<FT, VARIANT, true>(T, Sw[0], Sw[1], Sw[2], Sw[3], Sw[4], Sw[5], FT(0), FT(0), FT(0), FT(0), Az);
  if (do_xy) {
    const FT* w = Pa + tb - 3;
    const FT S0 = w[0], S1 = w[1], S2 = w[2], S3 = w[3], S4 = w[4], S5 = w[5];
    FT f;
    if (A == 0) f = tp_flux<FT, VARIANT, true>(T, S0, S1, S2, S3, S4, S5, FT(0), FT(0), FT(0), FT(0), Ax);
    else if (A == 1) f = tp_flux<FT, VARIANT, false>(T, S0, S1, S2, S3, S4, S5, Pu[tb - 2 * BW], Pu[tb - BW], Pu[tb], Pu[tb + BW], Ax);
    else f = tp_flux<FT, VARIANT, false>(T, S0, S1, S2, S3, S4, S5, Su[1], Su[2], Su[3], Su[4], Ax);
    cur[A * FXC + fxo] = f;
  }
  if (do_xy) {
    const FT* w = Pa + tb;
    const FT S0 = w[-3 * BW], S1 = w[-2 * BW], S2 = w[-BW], S3 = w[0], S4 = w[BW], S5 = w[2 * BW];
    FT f;
    if (A == 0) f = tp_flux<FT, VARIANT, false>(T, S0, S1, S2, S3, S4, S5, Pv[tb - 2], Pv[tb - 1], Pv[tb], Pv[tb + 1], Ay);
    else if (A == 1) f = tp_flux<FT, VARIANT, true>(T, S0, S1, S2, S3, S4, S5, FT(0), FT(0), FT(0), FT(0), Ay);
    else f = tp_flux<FT, VARIANT, false>(T, S0, S1, S2, S3, S4, S5, Sv[1], Sv[2], Sv[3], Sv[4], Ay);
    cur[FYB + A * FYC + fyo] = f;
  }
  return fz;
}

// The extra warp, one plane step: y-flux slot TY of column `lane` (all lanes) and x-flux slot TX of row `lane` (lanes < TY) of the
// three components.  Kept small (one reconstruction body per direction, a loop over the components): it is a ninth of the
// CTA's threads and must not cost instruction-cache space.  A flux along the component's own direction passes its window's
// inner four values as the advecting inputs, which is exactly what the SAME form computes.
template <typename FT, int VARIANT>
__device__ __forceinline__ void tp_extra_fluxes(const SchemeTables<FT>& T, const unsigned char* __restrict__ planes, int slot, int tbx,
                                                int tby, bool do_x, const FT (&Su)[6], const FT (&Sv)[6], FT* __restrict__ cur, int fxo,
                                                int fyo, FT Ax, FT Ay) {
  constexpr int BW = TP_BW, FXC = TP_TY * (TP_TX + 1), FYC = (TP_TY + 1) * TP_TX, FYB = 3 * FXC;
  const FT* Pu = reinterpret_cast<const FT*>(planes);
  const FT* Pv = reinterpret_cast<const FT*>(planes + slot);
#pragma unroll 1
  for (int A = 0; A < 3; ++A) {
    const FT* Pa = reinterpret_cast<const FT*>(planes + A * slot);
    {
      const FT* w = Pa + tby;
      const FT S0 = w[-3 * BW], S1 = w[-2 * BW], S2 = w[-BW], S3 = w[0], S4 = w[BW], S5 = w[2 * BW];
      FT v0, v1, v2, v3;
      if (A == 0) { v0 = Pv[tby - 2]; v1 = Pv[tby - 1]; v2 = Pv[tby]; v3 = Pv[tby + 1]; }
      else if (A == 1) { v0 = S1; v1 = S2; v2 = S3; v3 = S4; }
      else { v0 = Sv[1]; v1 = Sv[2]; v2 = Sv[3]; v3 = Sv[4]; }
      cur[FYB + A * FYC + fyo] = tp_flux<FT, VARIANT, false>(T, S0, S1, S2, S3, S4, S5, v0, v1, v2, v3, Ay);
    }
    if (do_x) {
      const FT* w = Pa + tbx - 3;
      const FT S0 = w[0], S1 = w[1], S2 = w[2], S3 = w[3], S4 = w[4], S5 = w[5];
      FT v0, v1, v2, v3;
      if (A == 0) { v0 = S1; v1 = S2; v2 = S3; v3 = S4; }
      else if (A == 1) { v0 = Pu[tbx - 2 * BW]; v1 = Pu[tbx - BW]; v2 = Pu[tbx]; v3 = Pu[tbx + BW]; }
      else { v0 = Su[1]; v1 = Su[2]; v2 = Su[3]; v3 = Su[4]; }
      cur[A * FXC + fxo] = tp_flux<FT, VARIANT, false>(T, S0, S1, S2, S3, S4, S5, v0, v1, v2, v3, Ax);
    }
  }
}

// Tendency and substep of plane k-1 for component A from the previous plane's staged x / y fluxes and the two z-fluxes.
// uold = the component's value at the cell (plane k-1 of the own column), gm = G⁻ there (loaded at the top of the step).
template <typename FT, int A>
__device__ __forceinline__ void tp_tendency(const FastTendJob<FT>& J, const FT* __restrict__ fx_prev, const FT* __restrict__ fy_prev,
                                            int fr_x, int fr_y, FT fz_old, FT fz_new, FT uold, FT gm, int n_prev, FT Vinv, int mode,
                                            FT* vp) {
  FT total = (fx_prev[fr_x + 1] - fx_prev[fr_x]) + (fy_prev[fr_y + TP_TX] - fy_prev[fr_y]);
  total = total + (fz_new - fz_old);
  FT G = -(Vinv * total);
  G = G + FT(0);
  J.Gn[A][n_prev] = G;
  if (mode >= 0) {
    const FT dt = J.dt, gamma = J.gamma, zeta = J.zeta;
    FT u = uold;
    if (mode == 0) u = u + dt * gamma * G;
    else if (mode == 1) u = u + dt * (gamma * G + zeta * gm);
    else if (mode == 2) u = u + dt * ((FT(1.5) + gamma) * G - (FT(0.5) + gamma) * gm);
    else u = u + dt * ((FT(1.5) + gamma) * G - FT(0));
    J.Unew[A][n_prev] = u;
    if (A == 1 && vp != nullptr) vp[n_prev] = u;
  }
}

// VARIANT as in tendency_fast_kernel: 0 = compiled without FMA contraction (bit-identical to the reference's CPU arithmetic),
// 1 = instantiated in tendency_fma.cu only (FMA contraction + regrouped WENO5 weights).
template <typename FT, int VARIANT>
__global__ void __launch_bounds__(TP_THREADS, TP_MINBLOCKS) tendency_plane_kernel(const __grid_constant__ PlaneJob<FT> P) {
  extern __shared__ unsigned char tp_raw[];
  typedef TPSizes<FT> Z;
  const FastTendJob<FT>& J = P.J;
  const GridDev<FT>& g = J.g;
  // carve-up: [stage][field] planes (128-byte aligned for TMA) | [buffer] x- and y-flux staging | mbarriers
  unsigned char* base = tp_raw + ((128u - (tp_smem_u32(tp_raw) & 127u)) & 127u);
  FT* const fluxes = reinterpret_cast<FT*>(base + Z::planes);
  const unsigned bars = tp_smem_u32(base + Z::planes + Z::flux_bytes), fbars = bars + 8 * TP_STAGES;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool regular = warp < TP_TY;
  const int i0 = blockIdx.x * TP_TX + 1, j0 = tf_tile_row(J) * TP_TY + 1, k0 = blockIdx.z * P.zseg + 1;
  const int nx = min(TP_TX, g.N[0] - i0 + 1), ny = min(TP_TY, g.N[1] - j0 + 1), nz = min(P.zseg, g.N[2] - k0 + 1);
  const int sz = (int)g.sz;
  const FT Ax = g.dy * g.dz, Ay = g.dx * g.dz, Az = g.dx * g.dy;
  const FT Vinv = FT(1) / ((g.dx * g.dy) * g.dz);
  const int mode = J.mode;

  // ---- roles.  Rows of the tile: thread (li, lj) = (lane, warp) computes x-flux slot li and y-flux slot lj of its row / column,
  // the z-flux, the divergence and the substep of its own column.  Extra warp: y-flux slot TY of column `lane` (all lanes) and
  // x-flux slot TX of row `lane` (lanes < TY).
  const int li = regular ? lane : TP_TX, lj = regular ? warp : min(lane, TP_TY - 1);       // x-flux slot (li, lj)
  const int yi = lane, yj = regular ? warp : TP_TY;                                        // y-flux slot (yi, yj)
  const int tb = (warp + 3) * TP_BW + lane + 4;                                            // own cell in a staged plane (rows only)
  const int tbx = (lj + 3) * TP_BW + li + 4, tby = (yj + 3) * TP_BW + yi + 4;
  const bool do_x = regular || lane < TP_TY;
  const int fxo = lj * (TP_TX + 1) + li, fyo = yj * TP_TX + yi;
  const int fr_x = warp * (TP_TX + 1) + lane, fr_y = warp * TP_TX + lane;                  // read positions of the row threads
  const bool store_ok = regular && lane < nx && warp < ny;
  // own columns (clamped into the allocated halo so that every load is in bounds; clamped values are never used):
  // the u window belongs to the x-flux slot's column, the v window to the y-flux slot's column, w to the own cell
  const int Imax = g.N[0] + g.H[0], Jmax = g.N[1] + g.H[1];
  const int nu = g.idx32(min(i0 + li, Imax), min(j0 + lj, Jmax), k0);
  const int nv = g.idx32(min(i0 + yi, Imax), min(j0 + yj, Jmax), k0);
  const int nw = g.idx32(min(i0 + lane, Imax), min(j0 + min(warp, TP_TY - 1), Jmax), k0);
  const int ncell = g.idx32(i0 + lane, j0 + warp, k0);                                     // used by store_ok threads only
  FT* const vp = (J.v_prev != nullptr && j0 + warp == 1) ? J.v_prev + g.N[1] * (int)g.sy : nullptr;

  const FT* __restrict__ U0 = J.U[0];
  const FT* __restrict__ U1 = J.U[1];
  const FT* __restrict__ U2 = J.U[2];

  // ---- mbarriers + the first two planes
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < TP_STAGES; ++s) tp_mbar_init(bars + 8 * s, 1);
#pragma unroll
    for (int s = 0; s < 4; ++s) tp_mbar_init(fbars + 8 * s, TP_THREADS / 32);      // one arrival per warp
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  const int c0 = (i0 - 4) - 1 + g.ox, c1 = (j0 - 3) - 1 + g.oy, c2 = k0 - 1 + g.oz;
  const unsigned planes_u32 = tp_smem_u32(base);
  auto issue = [&](int p, int stage) {
    const unsigned bar = bars + 8 * stage;
    tp_mbar_expect_tx(bar, 3u * Z::box_bytes);
#pragma unroll
    for (int f = 0; f < 3; ++f) tp_tma_load_3d(planes_u32 + (stage * 3 + f) * Z::slot, &P.tm[f], bar, c0, c1, c2 + p);
  };
#if TP_NO_TMA
  // A/B variant: the same boxes copied by the CTA's threads (LDG → STS) instead of the bulk-copy engine
  const FT* const Uf[3] = {J.U[0], J.U[1], J.U[2]};
  auto copy_planes = [&](int p, int stage) {
    for (int e = tid; e < 3 * TP_BW * TP_BH; e += TP_THREADS) {
      const int f = e / (TP_BW * TP_BH), r = e - f * (TP_BW * TP_BH), row = r / TP_BW, col = r - row * TP_BW;
      const int x = c0 + col, y = c1 + row;
      FT v = FT(0);
      if (x < g.px && y < g.py) v = Uf[f][x + g.px * (y + g.py * (c2 + p))];
      reinterpret_cast<FT*>(base + (stage * 3 + f) * Z::slot)[r] = v;
    }
  };
  copy_planes(0, 0); copy_planes(1, 1);
  __syncthreads();
#else
  if (tid == 0) { issue(0, 0); issue(1, 1); }
#endif

  // ---- z windows: planes k-3 … k+2 of the own columns, k = k0
  FT Su[6], Sv[6], Sw[6];
#pragma unroll
  for (int m = 0; m < 6; ++m) { Su[m] = U0[nu + (m - 3) * sz]; Sv[m] = U1[nv + (m - 3) * sz]; Sw[m] = U2[nw + (m - 3) * sz]; }
  FT fzu = FT(0), fzv = FT(0), fzw = FT(0);

  // Synchronisation between the warps of the CTA is split-phase: after writing its x / y fluxes of plane p a warp ARRIVES on the
  // flux barrier of staging buffer p mod 4 and only WAITS for it one phase later, just before it differences the fluxes of plane
  // p (in step p+1) — so a warp is never held at a CTA-wide barrier.  Hazards: (1) buffer (p+3) mod 4 = (p-1) mod 4 is rewritten
  // in step p+3, which a warp enters only after its wait of step p+2, i.e. after every warp has finished step p+1 and with it the
  // reads of buffer (p-1) mod 4 in step p (four buffers, not three); (2) the TMA slot of plane p-1 is refilled (plane p+2) by
  // thread 0 after its wait in step p, when every warp has arrived after reading plane p-1.
  int stage = 0;
  unsigned phase = 0;
  int koff = 0;                                  // (k - k0)·sz
  const bool gm_needed = store_ok && (mode == 1 || mode == 2);
  constexpr int FXC = TP_TY * (TP_TX + 1), FYC = (TP_TY + 1) * TP_TX, FBUF = Z::fx_elems + Z::fy_elems;
#pragma unroll 1
  for (int p = 0; p <= nz; ++p) {
#if TP_NO_TMA
    if (p + 2 <= nz) copy_planes(p + 2, stage == 0 ? 2 : stage - 1);
#else
    if (tid == 0 && p == 0 && 2 <= nz) issue(2, 2);
#endif
    // loads that are only needed at the end of the step or in the next one: next plane of the z windows, G⁻ of plane k-1
    FT nu5 = FT(0), nv5 = FT(0), nw5 = FT(0), gm0 = FT(0), gm1 = FT(0), gm2 = FT(0);
    const int n_prev = ncell + koff - sz;
    if (p < nz) { nu5 = U0[nu + koff + 3 * sz]; nv5 = U1[nv + koff + 3 * sz]; nw5 = U2[nw + koff + 3 * sz]; }
    if (gm_needed && p >= 1) { gm0 = J.Gm[0][n_prev]; gm1 = J.Gm[1][n_prev]; gm2 = J.Gm[2][n_prev]; }
#if !TP_NO_TMA
    tp_mbar_wait(bars + 8 * stage, phase);
#endif
    const FT* Pu = reinterpret_cast<const FT*>(base + (stage * 3 + 0) * Z::slot);
    const FT* Pv = reinterpret_cast<const FT*>(base + (stage * 3 + 1) * Z::slot);
    const FT* Pw = reinterpret_cast<const FT*>(base + (stage * 3 + 2) * Z::slot);
    FT* cur = fluxes + (p & 3) * FBUF;
    const FT* prev = fluxes + ((p + 3) & 3) * FBUF;
    const bool do_xy = p < nz;
    FT fu = FT(0), fv = FT(0), fw = FT(0);
    if (regular) {
      fu = tp_row_fluxes<FT, VARIANT, 0>(J.tab, Pu, Pv, Pw, tb, do_xy, Su, Sv, Sw, cur, fxo, fyo, Ax, Ay, Az);
      fv = tp_row_fluxes<FT, VARIANT, 1>(J.tab, Pu, Pv, Pw, tb, do_xy, Su, Sv, Sw, cur, fxo, fyo, Ax, Ay, Az);
      fw = tp_row_fluxes<FT, VARIANT, 2>(J.tab, Pu, Pv, Pw, tb, do_xy, Su, Sv, Sw, cur, fxo, fyo, Ax, Ay, Az);
    }
    else if (do_xy) tp_extra_fluxes<FT, VARIANT>(J.tab, base + stage * 3 * Z::slot, Z::slot, tbx, tby, do_x, Su, Sv, cur, fxo, fyo, Ax, Ay);
#if TP_NO_TMA
    __syncthreads();
#else
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(fbars + 8 * (p & 3)) : "memory");
    if (p >= 1) {
      tp_mbar_wait(fbars + 8 * ((p + 3) & 3), (unsigned)(((p - 1) >> 2) & 1));
      if (tid == 0 && p + 2 <= nz) issue(p + 2, stage == 0 ? 2 : stage - 1);
    }
#endif
    if (store_ok && p >= 1) {
      tp_tendency<FT, 0>(J, prev, prev + Z::fx_elems, fr_x, fr_y, fzu, fu, Su[2], gm0, n_prev, Vinv, mode, nullptr);
      tp_tendency<FT, 1>(J, prev + FXC, prev + Z::fx_elems + FYC, fr_x, fr_y, fzv, fv, Sv[2], gm1, n_prev, Vinv, mode, vp);
      tp_tendency<FT, 2>(J, prev + 2 * FXC, prev + Z::fx_elems + 2 * FYC, fr_x, fr_y, fzw, fw, Sw[2], gm2, n_prev, Vinv, mode, nullptr);
    }
#if TP_NO_TMA
    __syncthreads();
#endif
    fzu = fu; fzv = fv; fzw = fw;
#pragma unroll
    for (int m = 0; m < 5; ++m) { Su[m] = Su[m + 1]; Sv[m] = Sv[m + 1]; Sw[m] = Sw[m + 1]; }
    Su[5] = nu5; Sv[5] = nv5; Sw[5] = nw5;
    koff += sz;
    if (++stage == TP_STAGES) { stage = 0; phase ^= 1u; }
  }
}

// ---- host side ---------------------------------------------------------------------------------------------------------
// cuTensorMapEncodeTiled through the runtime's driver entry point (libnhstep.so links cudart statically and not libcuda)
template <typename FT>
inline void tp_encode_field_map(CUtensorMap* out, const FT* field, const GridDev<FT>& g) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  if (!encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    NH_REQUIRE(fn != nullptr && q == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available from this driver");
    encode = (EncodeFn)fn;
  }
  const cuuint64_t dims[3] = {(cuuint64_t)g.px, (cuuint64_t)g.py, (cuuint64_t)g.pz};
  const cuuint64_t strides[2] = {(cuuint64_t)g.px * sizeof(FT), (cuuint64_t)g.px * g.py * sizeof(FT)};
  const cuuint32_t box[3] = {TP_BW, TP_BH, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = encode(out, sizeof(FT) == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)field, dims,
                      strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  NH_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed for a velocity field (error " + std::to_string((int)r) + ")");
}

// planes per CTA: the longest segment (a power of two between 8 and 64 planes) that still gives every SM's TP_MINBLOCKS CTA
// slots at least six CTAs over the launch, so that the slots stay filled until the end (measured at 128³ … 512³: tools/plane_probe.cu)
inline int tp_choose_zseg(int Nx, int Ny, int Nz, int sms) {
  const long long cols = (long long)ceil_div(Nx, TP_TX) * ceil_div(Ny, TP_TY), want = 6LL * sms * TP_MINBLOCKS;
  int zseg = 8;
  for (int z = 16; z <= 64; z *= 2)
    if (z <= Nz && cols * ceil_div(Nz, z) >= want) zseg = z;
  return std::min(zseg, Nz);
}

template <typename FT> void launch_plane_tendency_fma(const PlaneJob<FT>& P, cudaStream_t stream);

template <typename FT>
inline void launch_plane_tendency(const PlaneJob<FT>& P, cudaStream_t stream, bool exact) {
  if (!exact) { launch_plane_tendency_fma<FT>(P, stream); return; }
  const GridDev<FT>& g = P.J.g;
  dim3 grd(ceil_div(g.N[0], TP_TX), P.J.ny_tiles > 0 ? P.J.ny_tiles : ceil_div(g.N[1], TP_TY), ceil_div(g.N[2], P.zseg));
  tendency_plane_kernel<FT, 0><<<grd, TP_THREADS, TPSizes<FT>::total, stream>>>(P);
}
