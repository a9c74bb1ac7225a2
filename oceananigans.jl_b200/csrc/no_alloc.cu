// no_alloc.cu — nh_tridiag_solve_static: the batched tridiagonal solve for callers that may not allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  Same sweep as nh_tridiag_solve (thomas_generic_kernel,
// batched_tridiagonal_solver.jl:209-235), but every array is the caller's and the scratch `t` of the forward sweep lives
// in a file-scope __device__ array instead of a cudaMalloc'ed buffer.
//
// * The array costs NH_TRIDIAG_STATIC_MAX_ELEMENTS · 8 B = 8 MiB on every GPU that loads libnhstep.so, until exit.
// * Every call on a device shares it and it keeps the last call's contents.  Nothing in it has to start at zero: the sweep
//   writes t[k] (k ≥ 1 along the solve direction) before the back-substitution reads it and never reads t[0].
//   Calls that overlap on one GPU (two streams, two threads) race on it; serialise them.
// * This entry point calls no cudaMalloc / cudaFree or allocating library routine.  The first call on a device looks the
//   array's device address up (cudaGetSymbolAddress) — make that call outside a graph capture.
#include "model.cuh"
#include <mutex>

static __device__ __align__(256) unsigned char nh_static_scratch[(size_t)NH_TRIDIAG_STATIC_MAX_ELEMENTS * sizeof(double)];

// In host code `nh_static_scratch` is only a handle for the symbol, not a device address: the address a kernel may
// dereference is looked up once per device, with that device current.
static const int kMaxDevices = 64;
static void* g_scratch_addr[kMaxDevices];
static std::mutex g_scratch_mutex;

static void* static_scratch_on_current_device(int device) {
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_addr[device]) {
    void* ptr = nullptr;
    CUDA_CHECK(cudaGetSymbolAddress((void**)&ptr, nh_static_scratch));
    NH_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null address for the static scratch array");
    g_scratch_addr[device] = ptr;
  }
  return g_scratch_addr[device];
}

extern "C" int32_t nh_tridiag_solve_static(int32_t device, int32_t float_type, const int32_t N[3], int32_t direction,
                                           const void* a_dev, const void* b_dev, int32_t b_is_3d, const void* c_dev,
                                           const void* rhs_dev, void* phi_dev, void* cuda_stream) {
  try {
    NH_REQUIRE(N && a_dev && b_dev && c_dev && rhs_dev && phi_dev, "null argument");
    NH_REQUIRE(float_type == NH_F32 || float_type == NH_F64, "float_type must be NH_F32 or NH_F64");
    NH_REQUIRE(direction >= 0 && direction < 3, "bad tridiagonal direction");
    NH_REQUIRE(N[0] >= 1 && N[1] >= 1 && N[2] >= 1, "size must be >= 1");
    NH_REQUIRE(device >= 0 && device < kMaxDevices, "bad device index");
    long long tot = (long long)N[0] * N[1] * N[2];
    if (tot > (long long)NH_TRIDIAG_STATIC_MAX_ELEMENTS)
      throw NhError("nh_tridiag_solve_static: Nx*Ny*Nz = " + std::to_string(tot) + " exceeds the static scratch array (" +
                    std::to_string((long long)NH_TRIDIAG_STATIC_MAX_ELEMENTS) +
                    " elements, NH_TRIDIAG_STATIC_MAX_ELEMENTS); use nh_tridiag_solve");
    CUDA_CHECK(cudaSetDevice(device));
    void* scratch = static_scratch_on_current_device(device);
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    long long lines = tot / N[direction];
    int blocks = (int)((lines + 127) / 128);
    if (float_type == NH_F32)
      thomas_generic_kernel<float><<<blocks, 128, 0, stream>>>(N[0], N[1], N[2], direction, (const float*)a_dev, (const float*)b_dev,
                                                               b_is_3d, (const float*)c_dev, (const float*)rhs_dev, (float*)phi_dev,
                                                               (float*)scratch);
    else
      thomas_generic_kernel<double><<<blocks, 128, 0, stream>>>(N[0], N[1], N[2], direction, (const double*)a_dev, (const double*)b_dev,
                                                                b_is_3d, (const double*)c_dev, (const double*)rhs_dev, (double*)phi_dev,
                                                                (double*)scratch);
    CUDA_CHECK(cudaGetLastError());
  } catch (const std::exception& e) { nh_set_error(e.what()); return 1; } catch (...) { nh_set_error("unknown error"); return 1; }
  return 0;
}
