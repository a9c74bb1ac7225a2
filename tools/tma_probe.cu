// tma_probe.cu — minimal check of the TMA helpers of tendency_plane.cuh (development aid): one CTA loads a box of a 3-D Float64
// tensor into shared memory and writes it back out.   ./tma_probe [dtype: 0 = FLOAT64, 1 = UINT64] [box width] [thread that issues]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../oceananigans.jl_b200/csrc/fft_generic.cuh"
#include "../oceananigans.jl_b200/csrc/tendency_plane.cuh"
void nh_set_error(const std::string&) {}
template <typename FT> void launch_plane_tendency_fma(const PlaneJob<FT>&, cudaStream_t) {}
template <typename FT> void launch_fused_tendency_fma(const FastTendJob<FT>&, cudaStream_t) {}

struct alignas(64) Probe { CUtensorMap tm; double* out; const double* src; const CUtensorMap* gtm; int bw, bh, c0, c1, c2, mode; };

__global__ void probe_kernel(const __grid_constant__ Probe P) {
  extern __shared__ unsigned char raw[];
  unsigned char* base = raw + ((128u - (tp_smem_u32(raw) & 127u)) & 127u);
  const unsigned bar = tp_smem_u32(base + 8192);
  if (threadIdx.x == 0) {
    tp_mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (P.mode == 0) {                      // mbarrier only: a plain arrive completes the phase
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
    } else if (P.mode == 1) {               // non-tensor bulk copy (no descriptor)
      tp_mbar_expect_tx(bar, (unsigned)(P.bw * P.bh * 8));
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(tp_smem_u32(base)), "l"(P.src), "r"(P.bw * P.bh * 8), "r"(bar) : "memory");
    } else if (P.mode == 3) {               // descriptor in global memory
      tp_mbar_expect_tx(bar, (unsigned)(P.bw * P.bh * 8));
      tp_tma_load_3d(tp_smem_u32(base), P.gtm, bar, P.c0, P.c1, P.c2);
    } else {
      tp_mbar_expect_tx(bar, (unsigned)(P.bw * P.bh * 8));
      tp_tma_load_3d(tp_smem_u32(base), &P.tm, bar, P.c0, P.c1, P.c2);
    }
  }
  tp_mbar_wait(bar, 0);
  for (int e = threadIdx.x; e < P.bw * P.bh; e += blockDim.x) P.out[e] = reinterpret_cast<double*>(base)[e];
}

int main(int argc, char** argv) {
  const int dtype = argc > 1 ? atoi(argv[1]) : 0, bw = argc > 2 ? atoi(argv[2]) : 40, bh = 14;
  const int px = 96, py = 30, pz = 12;
  std::vector<double> h((size_t)px * py * pz);
  for (size_t q = 0; q < h.size(); ++q) h[q] = (double)q;
  double *d, *out;
  cudaMalloc(&d, h.size() * 8); cudaMalloc(&out, bw * bh * 8);
  cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  Probe P;
  const int wide = dtype >= 2 ? 2 : 1;      // 2: the doubles described as pairs of 4-byte elements
  const cuuint64_t dims[3] = {(cuuint64_t)px * wide, py, pz}, strides[2] = {px * 8, (cuuint64_t)px * py * 8};
  const cuuint32_t box[3] = {(cuuint32_t)(bw * wide), (cuuint32_t)bh, 1}, es[3] = {1, 1, 1};
  const CUtensorMapDataType dts[4] = {CU_TENSOR_MAP_DATA_TYPE_FLOAT64, CU_TENSOR_MAP_DATA_TYPE_UINT64, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, CU_TENSOR_MAP_DATA_TYPE_UINT32};
  CUresult r = ((EncodeFn)fn)(&P.tm, dts[dtype], 3, d, dims, strides, box, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)(argc > 4 ? atoi(argv[4]) : 3),
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode rc=%d\n", (int)r);
  CUtensorMap* gtm; cudaMalloc(&gtm, sizeof(CUtensorMap)); cudaMemcpy(gtm, &P.tm, sizeof(CUtensorMap), cudaMemcpyHostToDevice); P.gtm = gtm;
  { const unsigned long long* w = (const unsigned long long*)&P.tm; printf("desc:"); for (int q = 0; q < 16; ++q) printf(" %016llx", w[q]); printf("\n"); }
  P.mode = argc > 3 ? atoi(argv[3]) : 2; P.src = d; P.out = out; P.bw = bw; P.bh = bh; P.c0 = (argc > 5 ? atoi(argv[5]) : 13) * wide; P.c1 = 2; P.c2 = 5;
  probe_kernel<<<1, 128, 16384>>>(P);
  cudaError_t e = cudaDeviceSynchronize();
  printf("mode=%d dtype=%d bw=%d: %s\n", P.mode, dtype, bw, cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<double> o(bw * bh);
  cudaMemcpy(o.data(), out, o.size() * 8, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int row = 0; row < bh; ++row) for (int col = 0; col < bw; ++col)
    if (o[row * bw + col] != h[(P.c0 / wide + col) + px * ((2 + row) + (size_t)py * 5)]) ++bad;
  printf("mismatches: %d (first values %.0f %.0f, expect %.0f)\n", bad, o[0], o[1], h[13 + px * (2 + (size_t)py * 5)]);
  return bad != 0;
}
