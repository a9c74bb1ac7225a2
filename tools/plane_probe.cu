// plane_probe.cu — standalone A/B harness for the two fused tendency sweeps (development aid, not part of libnhstep.so):
// runs tendency_fast_kernel (tile) and tendency_plane_kernel (TMA plane-marching) on the same random fields, compares Gⁿ and U*
// bit for bit and times both.   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a [-fmad=false] -DVAR=0|1 \
//     tools/plane_probe.cu oceananigans.jl_b200/csrc/build/grid_host.o -o plane_probe ;  ./plane_probe [N] [zseg] [mode]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <random>
#include "../oceananigans.jl_b200/csrc/fft_generic.cuh"
#include "../oceananigans.jl_b200/csrc/tendency_plane.cuh"
#include "../oceananigans.jl_b200/csrc/grid_host.h"
#ifndef VAR
#define VAR 1
#endif
void nh_set_error(const std::string&) {}
template <typename FT> void launch_plane_tendency_fma(const PlaneJob<FT>&, cudaStream_t) {}
template <typename FT> void launch_fused_tendency_fma(const FastTendJob<FT>&, cudaStream_t) {}
typedef double FT;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

int main(int argc, char** argv) {
  const int N = argc > 1 ? atoi(argv[1]) : 64;
  const int Nx = N, Ny = argc > 4 ? atoi(argv[4]) : N, Nz = argc > 5 ? atoi(argv[5]) : N;
  int zseg = argc > 2 ? atoi(argv[2]) : 0;
  const int mode = argc > 3 ? atoi(argv[3]) : 1;
  GridDev<FT> g;
  memset(&g, 0, sizeof g);
  g.N[0] = Nx; g.N[1] = Ny; g.N[2] = Nz;
  for (int e = 0; e < 3; ++e) { g.H[e] = 3; g.topo[e] = NH_PERIODIC; }
  g.ox = 16; g.px = (g.ox + Nx + 1 + 3 + 15) / 16 * 16; g.oy = 3; g.py = Ny + 7; g.oz = 3; g.pz = Nz + 7;
  g.sy = g.px; g.sz = (long long)g.px * g.py;
  g.dx = 1.0 / Nx; g.dy = 1.0 / Ny; g.dz = 1.0 / Nz; g.z_regular = 1; g.zoff = 5;
  const size_t vol = (size_t)g.px * g.py * g.pz;
  std::vector<FT> h(vol);
  std::mt19937_64 rng(1);
  std::uniform_real_distribution<FT> d(-1, 1);
  FT *U[3], *Gm[3], *Gn[2][3], *Un[2][3];
  for (int a = 0; a < 3; ++a) {
    CK(cudaMalloc(&U[a], vol * sizeof(FT))); CK(cudaMalloc(&Gm[a], vol * sizeof(FT)));
    for (auto& v : h) v = d(rng);
    CK(cudaMemcpy(U[a], h.data(), vol * sizeof(FT), cudaMemcpyHostToDevice));
    for (auto& v : h) v = d(rng);
    CK(cudaMemcpy(Gm[a], h.data(), vol * sizeof(FT), cudaMemcpyHostToDevice));
    for (int w = 0; w < 2; ++w) {
      CK(cudaMalloc(&Gn[w][a], vol * sizeof(FT))); CK(cudaMalloc(&Un[w][a], vol * sizeof(FT)));
      CK(cudaMemset(Gn[w][a], 0, vol * sizeof(FT))); CK(cudaMemset(Un[w][a], 0, vol * sizeof(FT)));
    }
  }
  nhhost::Tables<FT> T;
  nhhost::make_scheme_tables<FT>(T);
  FastTendJob<FT> J;
  J.g = g;
  static_assert(sizeof(T) == sizeof(J.tab), "tables");
  memcpy(&J.tab, &T, sizeof T);
  J.dt = 1e-3; J.gamma = 5.0 / 12; J.zeta = -17.0 / 60; J.mode = mode;
  for (int a = 0; a < 3; ++a) { J.U[a] = U[a]; J.Gm[a] = Gm[a]; }
  CK(cudaFuncSetAttribute(tendency_fast_kernel<FT, VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(TF_SMEM_ELEMS * sizeof(FT))));
  CK(cudaFuncSetAttribute(tendency_plane_kernel<FT, VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, TPSizes<FT>::total));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  if (zseg <= 0) zseg = tp_choose_zseg(Nx, Ny, Nz, sms);
  PlaneJob<FT> P;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float ms[2] = {0, 0};
  for (int which = 0; which < 2; ++which) {
    for (int a = 0; a < 3; ++a) { J.Gn[a] = Gn[which][a]; J.Unew[a] = Un[which][a]; }
    P.J = J; P.zseg = zseg;
    for (int a = 0; a < 3; ++a) tp_encode_field_map<FT>(&P.tm[a], U[a], g);
    const int reps = 5;
    for (int r = 0; r < reps + 1; ++r) {
      if (r == 1) cudaEventRecord(e0);
      if (which == 0) {
        dim3 grd(ceil_div(Nx, TF_TX), ceil_div(Ny, TF_TY), ceil_div(Nz, TF_TZ));
        tendency_fast_kernel<FT, VAR><<<grd, TF_THREADS, TF_SMEM_ELEMS * sizeof(FT)>>>(J);
      } else {
        dim3 grd(ceil_div(Nx, TP_TX), ceil_div(Ny, TP_TY), ceil_div(Nz, zseg));
        tendency_plane_kernel<FT, VAR><<<grd, TP_THREADS, TPSizes<FT>::total>>>(P);
      }
      CK(cudaGetLastError());
    }
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("kernel %d (%s): %s\n", which, which ? "plane" : "tile", cudaGetErrorString(e)); return 2; }
    cudaEventElapsedTime(&ms[which], e0, e1);
    ms[which] /= reps;
  }
  long long bad = 0, cells = 0;
  std::vector<FT> a0(vol), a1(vol);
  for (int f = 0; f < 6; ++f) {
    FT* p0 = f < 3 ? Gn[0][f] : Un[0][f - 3];
    FT* p1 = f < 3 ? Gn[1][f] : Un[1][f - 3];
    cudaMemcpy(a0.data(), p0, vol * sizeof(FT), cudaMemcpyDeviceToHost);
    cudaMemcpy(a1.data(), p1, vol * sizeof(FT), cudaMemcpyDeviceToHost);
    long long fb = 0; int first[3] = {-1, -1, -1};
    for (int k = 1; k <= Nz; ++k) for (int j = 1; j <= Ny; ++j) for (int i = 1; i <= Nx; ++i) {
      const size_t n = g.idx(i, j, k);
      ++cells;
      if (memcmp(&a0[n], &a1[n], sizeof(FT)) != 0) { if (!fb) { first[0] = i; first[1] = j; first[2] = k; } ++fb; }
    }
    if (fb) printf("field %d: %lld cells differ, first at (%d,%d,%d): tile %.17g plane %.17g\n", f, fb, first[0], first[1], first[2],
                   a0[g.idx(first[0], first[1], first[2])], a1[g.idx(first[0], first[1], first[2])]);
    bad += fb;
  }
  printf("N=%dx%dx%d VAR=%d mode=%d zseg=%d: tile %.4f ms, plane %.4f ms, %lld of %lld values differ\n", Nx, Ny, Nz, VAR, mode, zseg, ms[0], ms[1],
         bad, cells);
  return bad ? 3 : 0;
}
