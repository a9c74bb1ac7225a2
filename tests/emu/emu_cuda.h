// emu_cuda.h — TEST INFRASTRUCTURE ONLY.  A CUDA-semantics stand-in that lets g++ compile the kernels of
// oceananigans.jl_b200/csrc/ and run them on host cores: every launch runs its blocks one after another, the threads of a block as
// cooperatively scheduled fibers (emu_runtime.cpp) so that __syncthreads() has its meaning.  It exists so that the index logic and
// the IEEE arithmetic of the shipped kernels can be checked against the oracle in a container without a GPU (tests/emu/README.md).
// It is never built or loaded by the product (the package loads oceananigans.jl_b200/libnhstep.so only and fails without a
// CUDA device); nothing measured with it is a performance number.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>

// ---- language keywords -------------------------------------------------------------------------------------------------
#define __host__
#define __device__
#define __global__
#define __forceinline__ inline
#define __noinline__
#define __launch_bounds__(...)
#define __grid_constant__
#define __align__(n) __attribute__((aligned(n)))
#define __shared__ static

struct uint3 { unsigned x, y, z; };
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct float2 { float x, y; };
struct double2 { double x, y; };
struct int2 { int x, y; };
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
static inline double2 make_double2(double x, double y) { return double2{x, y}; }
static inline int2 make_int2(int x, int y) { return int2{x, y}; }

namespace emu {
extern uint3 thread_idx, block_idx;
extern dim3 block_dim, grid_dim;
void launch(dim3 grid, dim3 block, size_t shared_bytes, const std::function<void()>& thread_body, const void* kernel = nullptr,
            const char* name = "");
void set_max_dynamic_shared(const void* kernel, int bytes);
void syncthreads();
void syncwarp();
unsigned char* dynamic_shared();
unsigned long long launches();
// mbarrier objects and bulk tensor copies of the TMA sweep (addresses are the 32-bit shared-window values of
// __cvta_generic_to_shared, i.e. truncated pointers into the launch's dynamic shared memory)
void mbar_init(unsigned bar, int count);
void mbar_arrive(unsigned bar);
void mbar_expect_tx(unsigned bar, unsigned bytes);
void mbar_wait(unsigned bar, unsigned parity);
void tma_load_3d(unsigned dst, const void* tensor_map, unsigned bar, int c0, int c1, int c2);
}  // namespace emu
#define threadIdx emu::thread_idx
#define blockIdx emu::block_idx
#define blockDim emu::block_dim
#define gridDim emu::grid_dim
static const int warpSize = 32;

static inline void __syncthreads() { emu::syncthreads(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { emu::syncwarp(); }
static inline void __threadfence_system() {}
static inline void __threadfence() {}
[[noreturn]] static inline void __trap() { std::fprintf(stderr, "emu: __trap()\n"); std::abort(); }

// ---- arithmetic intrinsics (IEEE operations; compile with -ffp-contract=off) ---------------------------------------------
static inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
static inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline float __frcp_rn(float a) { return 1.0f / a; }
static inline int __double2hiint(double d) { uint64_t u; std::memcpy(&u, &d, 8); return (int)(u >> 32); }
static inline int __double2loint(double d) { uint64_t u; std::memcpy(&u, &d, 8); return (int)(u & 0xffffffffu); }
static inline double __hiloint2double(int hi, int lo) {
  uint64_t u = ((uint64_t)(uint32_t)hi << 32) | (uint32_t)lo; double d; std::memcpy(&d, &u, 8); return d;
}
template <typename T> static inline T __ldg(const T* p) { return *p; }
static inline double rsqrt(double x) { return 1.0 / std::sqrt(x); }
static inline float rsqrtf(float x) { return 1.0f / std::sqrt(x); }
static inline unsigned __cvta_generic_to_shared(const void* p) { return (unsigned)(uintptr_t)p; }
using std::fabs; using std::sqrt; using std::cbrt; using std::fmin; using std::fmax; using std::floor; using std::ceil;
using std::sin; using std::cos; using std::exp; using std::log; using std::pow; using std::isnan; using std::isfinite;
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
static inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }
static inline long long min(long long a, long long b) { return a < b ? a : b; }
static inline long long max(long long a, long long b) { return a > b ? a : b; }
static inline long long min(long long a, int b) { return a < b ? a : b; }
static inline long long max(long long a, int b) { return a > b ? a : b; }
static inline long long min(int a, long long b) { return a < b ? a : b; }
static inline long long max(int a, long long b) { return a > b ? a : b; }
static inline size_t min(size_t a, size_t b) { return a < b ? a : b; }
static inline size_t max(size_t a, size_t b) { return a > b ? a : b; }
static inline float min(float a, float b) { return std::fmin(a, b); }
static inline float max(float a, float b) { return std::fmax(a, b); }
static inline double min(double a, double b) { return std::fmin(a, b); }
static inline double max(double a, double b) { return std::fmax(a, b); }
template <typename T> static inline T atomicAdd(T* p, T v) { T o = *p; *p = o + v; return o; }
template <typename T> static inline T atomicMax(T* p, T v) { T o = *p; if (v > o) *p = v; return o; }

// ---- runtime API ------------------------------------------------------------------------------------------------------
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorNotSupported = 801, cudaErrorInvalidValue = 1 };
typedef void* cudaStream_t;
typedef void* cudaEvent_t;
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount = 16 };
enum { cudaEnableDefault = 0, cudaIpcMemLazyEnablePeerAccess = 1 };
enum cudaDriverEntryPointQueryResult { cudaDriverEntryPointSuccess = 0, cudaDriverEntryPointSymbolNotFound = 1 };
struct cudaIpcMemHandle_t { char reserved[64]; };

static inline const char* cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "emulated CUDA runtime error"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
// sixteen "devices": a rank process of the multi-rank tests addresses the one with its rank's index
inline int emu_current_device = 0;
static inline cudaError_t cudaSetDevice(int d) { emu_current_device = d; return cudaSuccess; }
static inline cudaError_t cudaGetDevice(int* d) { *d = emu_current_device; return cudaSuccess; }
static inline cudaError_t cudaGetDeviceCount(int* n) { *n = 16; return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaDeviceGetAttribute(int* v, int attr, int) { *v = attr == cudaDevAttrMultiProcessorCount ? 148 : 0; return cudaSuccess; }
template <typename F> static inline cudaError_t cudaFuncSetAttribute(F f, int attr, int value) {
  if (attr == 8) emu::set_max_dynamic_shared((const void*)f, value);      // cudaFuncAttributeMaxDynamicSharedMemorySize
  return value > 227 * 1024 ? 1 : 0;
}
cudaError_t cudaMalloc(void** p, size_t bytes);
template <typename T> static inline cudaError_t cudaMalloc(T** p, size_t bytes) { return cudaMalloc((void**)p, bytes); }
cudaError_t cudaFree(void* p);
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { std::memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { std::memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemset(void* d, int v, size_t n) { std::memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = nullptr) { std::memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = nullptr; return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = nullptr) { return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.0f; return cudaSuccess; }
template <typename T> static inline cudaError_t cudaGetSymbolAddress(void** p, T& symbol) { *p = (void*)&symbol; return cudaSuccess; }
// IPC handles: the pointer itself within one process; with NHSTEP_EMU_SHM=1 every allocation is a file under /dev/shm and the
// handle names it, so that several rank processes open each other's buffers as they do over NVLink (emu_runtime.cpp)
cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t* h, void* p);
cudaError_t cudaIpcOpenMemHandle(void** p, cudaIpcMemHandle_t h, unsigned flags);
cudaError_t cudaIpcCloseMemHandle(void* p);

// ---- driver API pieces the TMA sweep names ---------------------------------------------------------------------------------
typedef int CUresult;
typedef uint32_t cuuint32_t;
typedef uint64_t cuuint64_t;
enum { CUDA_SUCCESS = 0 };
struct alignas(64) CUtensorMap { unsigned long long opaque[16]; };
enum CUtensorMapDataType { CU_TENSOR_MAP_DATA_TYPE_FLOAT32 = 7, CU_TENSOR_MAP_DATA_TYPE_FLOAT64 = 8 };
enum CUtensorMapInterleave { CU_TENSOR_MAP_INTERLEAVE_NONE = 0 };
enum CUtensorMapSwizzle { CU_TENSOR_MAP_SWIZZLE_NONE = 0 };
enum CUtensorMapL2promotion { CU_TENSOR_MAP_L2_PROMOTION_L2_256B = 3 };
enum CUtensorMapFloatOOBfill { CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE = 0 };
// the tensor map of the stand-in: base address, element size, the three extents, the two byte strides, the box
CUresult emu_cuTensorMapEncodeTiled(CUtensorMap* out, CUtensorMapDataType type, cuuint32_t rank, void* address, const cuuint64_t* dims,
                                    const cuuint64_t* strides, const cuuint32_t* box, const cuuint32_t* element_strides,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static inline cudaError_t cudaGetDriverEntryPoint(const char* name, void** fn, unsigned long long, cudaDriverEntryPointQueryResult* q = nullptr) {
  const bool known = std::strcmp(name, "cuTensorMapEncodeTiled") == 0;
  *fn = known ? (void*)&emu_cuTensorMapEncodeTiled : nullptr;
  if (q) *q = known ? cudaDriverEntryPointSuccess : cudaDriverEntryPointSymbolNotFound;
  return known ? cudaSuccess : cudaErrorNotSupported;
}
