// dist.cuh — multi-GPU (y-slab) pieces of the fused fast path: one process per GPU, local grid Nx × Ny/R × Nz.
//   src/DistributedComputations/halo_communication.jl:114-366      (halo exchange: pack → send/recv → unpack)
//   src/DistributedComputations/distributed_transpose.jl:25-191, distributed_fft_based_poisson_solver.jl:141-178
//                                                                   (z-local → y-local transposes around the FFTs)
// The library only packs/unpacks and transforms; the messages themselves (NCCL send/recv and all-to-all over NVLink)
// are issued by the host driver on the buffers exposed through nh_dist_buffer().
//
// Spectral layout (shared with poisson_fast.cuh): element (kx, jl, kz) of the local half spectrum lives at
//   (kx % Cx) + Cx·(jl + NyL·(kz + Nz·(kx / Cx)))            — R contiguous blocks of Cx columns, block d goes to rank d
// so the x-forward / z kernels write the all-to-all send buffer directly (no pack kernel), and after the exchange rank r
// holds, for its Cx columns, block s = the rows of rank s:  (kxl, j = s·NyL + jl, kz) at kxl + Cx·(jl + NyL·(kz + Nz·s)),
// which the y-solve kernel indexes in place (no unpack kernel).
#pragma once
#include "poisson_fast.cuh"

// in-place c2c along z on the blocked layout.  grid (R·Cx/8, NyL), block N = Nz
template <typename FT, typename CT, int N, bool INV>
__global__ void __launch_bounds__(N, pf_min_ctas(N)) pd_z_kernel(const __grid_constant__ PFJob<FT, CT> J, int R) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  constexpr int pitch = FFPitch8<N>::value;
  const int tid = threadIdx.x, c = tid & (PF_L - 1), t = tid >> 3;
  const int kx = blockIdx.x * PF_L + c, jl = blockIdx.y;
  const bool active = kx <= J.Nx / 2;
  const long long plane = (long long)J.Cx * J.Ny;                 // stride between kz planes inside a block
  CT* base = J.S + (kx % J.Cx) + (long long)J.Cx * jl + plane * J.Nz * (kx / J.Cx) + plane * t;
  const long long step = plane * (N / 8);
  CT v[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) { if (active) v[m] = base[m * step]; else { v[m].x = 0; v[m].y = 0; } }
  ff_line_fft_regs<CT, N, INV>(v, X + c * pitch, t, J.twz);
  if (active) {
    const FT scale = FT(1) / FT(N);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      if (INV) { v[m].x *= scale; v[m].y *= scale; }
      base[m * step] = v[m];
    }
  }
}

// forward z pass with the transpose folded in: block d of the result is stored straight into rank d's y-local buffer
// (block `rank` of it) through its peer pointer, so no all-to-all follows — only a barrier.  grid (R·Cx/8, NyL), block Nz
template <typename FT, typename CT, int N>
__global__ void __launch_bounds__(N, pf_min_ctas(N)) pd_z_scatter_kernel(const __grid_constant__ PFJob<FT, CT> J, int rank) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  constexpr int pitch = FFPitch8<N>::value;
  const int tid = threadIdx.x, c = tid & (PF_L - 1), t = tid >> 3;
  const int kx = blockIdx.x * PF_L + c, jl = blockIdx.y;
  const bool active = kx <= J.Nx / 2;
  const long long plane = (long long)J.Cx * J.Ny;
  const long long inner = (kx % J.Cx) + (long long)J.Cx * jl + plane * t;
  const int d = kx / J.Cx;
  const CT* src = J.S + inner + plane * J.Nz * d;
  CT* dst = J.peer[d] + inner + plane * J.Nz * rank;
  const long long step = plane * (N / 8);
  CT v[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) { if (active) v[m] = src[m * step]; else { v[m].x = 0; v[m].y = 0; } }
  ff_line_fft_regs<CT, N, false>(v, X + c * pitch, t, J.twz);
  if (active) {
#pragma unroll
    for (int m = 0; m < 8; ++m) dst[m * step] = v[m];
  }
}

// The transpose alone (vertically Bounded models: z is not transformed before the y pass): block d of the z-local buffer is
// copied into block `rank` of rank d's y-local buffer.  One thread per element.
template <typename CT>
struct ScatterJob { const CT* src; CT* peer[PF_MAXR]; long long chunk; int R, rank; };
template <typename CT>
__global__ void __launch_bounds__(256) pd_scatter_kernel(const __grid_constant__ ScatterJob<CT> J) {
  const long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= J.chunk * J.R) return;
  const int d = (int)(e / J.chunk);
  J.peer[d][J.chunk * J.rank + (e - d * J.chunk)] = J.src[e];
}

// forward c2c along the GLOBAL y, spectral division, inverse c2c along y, on the transposed buffer T (this rank's Cx
// columns, all Ny rows).  grid (Cx/8, Nz), block N = Ny (global)
// MODE 0: all three; 1: forward transform only, in place (a Thomas sweep along z follows, pf_thomas_kernel<…, true>);
// 2: inverse transform only, with the 1/N scale and the scatter back.
template <typename FT, typename CT, int N, int MODE = 0>
__global__ void __launch_bounds__(N, pf_min_ctas(N)) pd_y_solve_kernel(const __grid_constant__ PFJob<FT, CT> J, CT* __restrict__ T, int rank) {
  extern __shared__ __align__(16) unsigned char pf_smem[];
  CT* X = reinterpret_cast<CT*>(pf_smem);
  constexpr int pitch = FFPitch8<N>::value;
  const int tid = threadIdx.x, c = tid & (PF_L - 1), t = tid >> 3;
  const int kxl = blockIdx.x * PF_L + c, kx = rank * J.Cx + kxl, kz = blockIdx.y;
  const bool active = kx <= J.Nx / 2;
  const int NyL = J.Ny;
  const long long plane = (long long)J.Cx * NyL;
  CT* base = T + kxl + plane * kz;
  CT v[8];
  long long off[8];
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    const int j = t + m * (N / 8);
    off[m] = (long long)J.Cx * (j % NyL) + plane * J.Nz * (j / NyL);
    if (active) v[m] = base[off[m]]; else { v[m].x = 0; v[m].y = 0; }
  }
  if (MODE != 2) ff_line_fft_regs<CT, N, false>(v, X + c * pitch, t, J.twy);
  if (MODE == 1) {
    if (active) {
#pragma unroll
      for (int m = 0; m < 8; ++m) base[off[m]] = v[m];
    }
    return;
  }
  if (MODE == 0) {
    const double lamx = active ? J.lx[kx] : 1.0, lamz = J.lz[kz];
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      const int j = t + m * (N / 8);
      // same association as the single-GPU kernel and the reference: ((λx + λy) + λz) - m
      double den = ((lamx + J.ly[j]) + lamz) - J.m;
      const bool den_ok = nh_den_ok(den);
      const double yr = nh_rcp_refined(den);
      CT r;
      r.x = (FT)nh_div_checked((double)(-v[m].x), den, yr, den_ok);
      r.y = (FT)nh_div_checked((double)(-v[m].y), den, yr, den_ok);
      if (J.m == 0.0 && kx == 0 && j == 0 && kz == 0) { r.x = 0; r.y = 0; }
      v[m] = r;
    }
  }
  ff_line_fft_regs<CT, N, true>(v, X + c * pitch, t, J.twy);
  if (active) {
    const FT scale = FT(1) / FT(N);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
      v[m].x *= scale; v[m].y *= scale;
      if (J.peer[0] == nullptr) base[off[m]] = v[m];
      else {
        // transpose back folded in: rows of rank s go to block `rank` of rank s's z-local buffer
        const int j = t + m * (N / 8), s = j / NyL;
        J.peer[s][kxl + (long long)J.Cx * (j - s * NyL) + plane * kz + plane * J.Nz * rank] = v[m];
      }
    }
  }
}

// y-halo pack / unpack: `w` rows next to each y face over the full parent extent in x and z, nf fields back to back.
//   lo buffer ↔ rows 1…w (send) / 1-w…0 (receive);  hi buffer ↔ rows Ny-w+1…Ny (send) / Ny+1…Ny+w (receive)
template <typename FT>
struct HaloPackJob { GridDev<FT> g; FT* f[NH_MAXTR + 4]; int nf; int w; FT* lo; FT* hi; };

template <typename FT>
__global__ void __launch_bounds__(256) y_halo_pack_kernel(const __grid_constant__ HaloPackJob<FT> J, int unpack) {
  const GridDev<FT>& g = J.g;
  const int Ex = g.N[0] + 2 * g.H[0], Ez = g.N[2] + 2 * g.H[2];
  const long long per_field = (long long)Ex * J.w * Ez;
  long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= per_field * J.nf) return;
  int f = (int)(e / per_field);
  long long r = e - f * per_field;
  int pi = (int)(r % Ex), row = (int)((r / Ex) % J.w), pk = (int)(r / ((long long)Ex * J.w));
  int i = pi + 1 - g.H[0], k = pk + 1 - g.H[2];
  if (!unpack) {
    J.lo[e] = J.f[f][g.idx(i, 1 + row, k)];
    J.hi[e] = J.f[f][g.idx(i, g.N[1] - J.w + 1 + row, k)];
  } else {
    J.f[f][g.idx(i, 1 - J.w + row, k)] = J.lo[e];
    J.f[f][g.idx(i, g.N[1] + 1 + row, k)] = J.hi[e];
  }
}


// ---- peer-to-peer barrier: one-sided flag writes over NVLink instead of a collective --------------------------------
// Thread r < R publishes this rank's epoch into slot `rank` of rank r's flag array (release, system scope: every store of
// the kernels that ran before this one on the stream — including their stores into peer memory — is ordered before it),
// then waits until slot r of its own array has reached the epoch (acquire).  All ranks issue the same sequence of
// barriers, so epochs only grow.  A wait that lasts longer than `timeout_ns` traps (a peer died) instead of hanging the GPU:
// 10 minutes by default, like a collective watchdog (ordinary host skew between ranks — output, a gather, a Python pause —
// must not kill the context); NHSTEP_BARRIER_TIMEOUT_S overrides it, 0 waits for ever.
struct PeerBarrierJob {
  unsigned int* peer_flags[PF_MAXR];     // flag arrays of all ranks (own array at index rank)
  int rank, R;
  unsigned int epoch;
  unsigned long long timeout_ns;         // 0: no timeout
};

static __global__ void peer_barrier_kernel(const __grid_constant__ PeerBarrierJob J) {
  const int r = threadIdx.x;
  if (r >= J.R) return;
  __threadfence_system();
  unsigned int* dst = J.peer_flags[r] + J.rank;
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(J.epoch) : "memory");
  const unsigned int* src = J.peer_flags[J.rank] + r;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(src) : "memory");
    if ((int)(v - J.epoch) >= 0) break;
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (J.timeout_ns != 0ull && t1 - t0 > J.timeout_ns) __trap();
  }
  __threadfence_system();
}
