// tendency_fma.cu — the fused tendency + substep sweep compiled WITH floating-point contraction (the Makefile builds this
// one translation unit with -fmad=true).  Selected by nh_model_desc.exact_arithmetic = 0: the same source as the exact
// kernel, so every a·b + c of the reference's expressions becomes one FMA (one rounding instead of two).  Results are not
// bit-identical to the reference's CPU code any more but stay within the north_star tolerance (tests/test_gpu_parity.py
// test_contracted_arithmetic_*), with roughly a third fewer FP64 instructions per cell.
// Measured at 256³ Float64 on B200 (tools/ab_pipeline.sh): 1.148 ms per launch as the exact kernel's source compiles,
// 1.120 with the unconditional six-point x window, 1.106 with the software-pipelined marching loops on top.  The exact
// build gains nothing from either (FP64-pipe bound with its checked divisions), so the switches live here.
#ifndef TF_XWINDOW
#define TF_XWINDOW 1
#endif
#ifndef TF_PIPELINE
#define TF_PIPELINE 1
#endif
#include "fft_generic.cuh"      // cx_of (two-element vector types)
#include "tendency_fast.cuh"
#include "tendency_plane.cuh"
#include "tendency_generic.cuh"

template <typename FT>
void launch_fused_tendency_fma(const FastTendJob<FT>& J, cudaStream_t stream) {
  // the attribute is per device: set once for each device a model of this process lives on
  static bool attr_set[64] = {};
  int dev = 0;
  CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 64 || !attr_set[dev]) {
    CUDA_CHECK(cudaFuncSetAttribute(tendency_fast_kernel<FT, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)(TF_SMEM_ELEMS * sizeof(FT))));
    if (dev < 64) attr_set[dev] = true;
  }
  const GridDev<FT>& g = J.g;
  dim3 grd(ceil_div(g.N[0], TF_TX), J.ny_tiles > 0 ? J.ny_tiles : ceil_div(g.N[1], TF_TY), ceil_div(g.N[2], TF_TZ));
  tendency_fast_kernel<FT, 1><<<grd, TF_THREADS, TF_SMEM_ELEMS * sizeof(FT), stream>>>(J);
}

template void launch_fused_tendency_fma<float>(const FastTendJob<float>&, cudaStream_t);
template void launch_fused_tendency_fma<double>(const FastTendJob<double>&, cudaStream_t);

// the plane-marching TMA kernel (tendency_plane.cuh) in the contracted build
template <typename FT>
void launch_plane_tendency_fma(const PlaneJob<FT>& P, cudaStream_t stream) {
  static bool attr_set[64] = {};
  int dev = 0;
  CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 64 || !attr_set[dev]) {
    CUDA_CHECK(cudaFuncSetAttribute(tendency_plane_kernel<FT, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TPSizes<FT>::total));
    if (dev < 64) attr_set[dev] = true;
  }
  const GridDev<FT>& g = P.J.g;
  dim3 grd(ceil_div(g.N[0], TP_TX), P.J.ny_tiles > 0 ? P.J.ny_tiles : ceil_div(g.N[1], TP_TY), ceil_div(g.N[2], P.zseg));
  tendency_plane_kernel<FT, 1><<<grd, TP_THREADS, TPSizes<FT>::total, stream>>>(P);
}
template void launch_plane_tendency_fma<float>(const PlaneJob<float>&, cudaStream_t);
template void launch_plane_tendency_fma<double>(const PlaneJob<double>&, cudaStream_t);

// the complete path's tiled tendency kernel (all fields, one launch) in the contracted build: the fast flux stage uses the
// quotient-free WENO5 weights, everything else is the same source compiled with FMA contraction
template <typename FT>
void launch_tiled_tendency_fma(const TendParams<FT>& P, dim3 grid, int fast_ok, cudaStream_t stream) {
  static bool attr_set[64] = {};
  int dev = 0;
  CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 64 || !attr_set[dev]) {
    CUDA_CHECK(cudaFuncSetAttribute(tendency_tiled_kernel<FT, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(TG_SMEM_ELEMS * sizeof(FT))));
    if (dev < 64) attr_set[dev] = true;
  }
  tendency_tiled_kernel<FT, 1><<<grid, dim3(TG_TX, TG_TY, 1), TG_SMEM_ELEMS * sizeof(FT), stream>>>(P, -1, fast_ok);
}
template void launch_tiled_tendency_fma<float>(const TendParams<float>&, dim3, int, cudaStream_t);
template void launch_tiled_tendency_fma<double>(const TendParams<double>&, dim3, int, cudaStream_t);
