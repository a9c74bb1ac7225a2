// common.cuh — internal types shared by the kernels of libnhstep.so (sm_100a only; no other backend).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include <stdexcept>
#include "../../include/nhstep.h"

#define NH_MAXTR NH_MAX_TRACERS

struct NhError : std::runtime_error { using std::runtime_error::runtime_error; };

void nh_set_error(const std::string& msg);

#define CUDA_CHECK(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { \
  throw NhError(std::string(#expr) + ": " + cudaGetErrorString(_e)); } } while (0)

#define NH_REQUIRE(cond, msg) do { if (!(cond)) throw NhError(msg); } while (0)

// ---- HBM layout -------------------------------------------------------------------------------------------
// Every field of a model shares one geometry: element (i,j,k) (the reference's 1-based indices, halo ≤ 0 and > N)
// lives at  (i-1+ox) + px*((j-1+oy) + py*(k-1+oz)).  ox is padded so that i = 1 starts a 128-byte line and px is a
// multiple of 16 elements; the reference's dense `parent` layout is produced only at nh_get_field/nh_set_field.
template <typename FT>
struct GridDev {
  int N[3], H[3], topo[3];
  int ox, oy, oz;
  int px, py, pz;
  long long sy, sz;           // strides in elements: sy = px, sz = px*py
  FT dx, dy;                  // Δx, Δy (1 when Flat)
  const FT* dzc;              // Δzᵃᵃᶜ[k]  at dzc[k + zoff]
  const FT* dzf;              // Δzᵃᵃᶠ[k]  at dzf[k + zoff]
  int zoff;
  int z_regular;              // 1 ⇒ dzc, dzf hold one constant
  FT dz;                      // that constant (regular z)
  __host__ __device__ __forceinline__ long long idx(int i, int j, int k) const {
    return (long long)(i - 1 + ox) + sy * (j - 1 + oy) + sz * (k - 1 + oz);
  }
  // 32-bit variant for kernels that are only used while the parent volume is below 2^31 elements
  __host__ __device__ __forceinline__ int idx32(int i, int j, int k) const {
    return (i - 1 + ox) + px * ((j - 1 + oy) + py * (k - 1 + oz));
  }
  __device__ __forceinline__ FT Dzc(int k) const { return dzc[k + zoff]; }
  __device__ __forceinline__ FT Dzf(int k) const { return dzf[k + zoff]; }
};

// One reconstruction scheme per direction.  kind: NH_ADV_*; B = buffer (1..3).
// Coefficient tables are generated on the host exactly as the reference does
// (src/Advection/reconstruction_coefficients.jl:100-115: exact rational → FT, last = 1 - Σ others).
template <typename FT>
struct SchemeTables {
  FT sym[3][4];        // sym[Bc][·]: Centered(2·Bc) coefficients, Bc = 1,2   (index 0 unused)
  FT upl[4][5];        // upl[B][·]: UpwindBiased(2B-1) left-biased coefficients, B = 1..3
  FT upr[4][5];        // right-biased
  FT wc[4][3][3];      // wc[B][r][·]: WENO sub-stencil coefficients, B = 2,3
  FT cstar[4][3];      // optimal weights
  FT eps;              // FT(1f-8)
};

struct SchemeDev { int kind, B; };

template <typename FT>
struct TendParams {
  GridDev<FT> g;
  const FT* U[3];
  const FT* c[NH_MAXTR];
  const FT* pHY;
  FT* GU[3];
  FT* Gc[NH_MAXTR];
  int ntr;
  SchemeDev adv[3];
  SchemeTables<FT> tab;
  int buoy_kind, iT, iS;
  FT grav, alpha, beta;
  int w_buoyancy;             // buoyancy acts on w directly (pHY′ is nothing)
  int has_coriolis; FT fcor;
  int has_closure; FT nu; FT kappa[NH_MAXTR];
  int vitd;                   // VerticallyImplicitTimeDiscretization: z-fluxes explicit on the walls k = 1, Nz+1 only
  const FT* nue;              // eddy viscosity νₑ at ccc (Smagorinsky closures); nullptr ⇒ the constants above
  const FT* kape[NH_MAXTR];   // eddy diffusivity field of each tracer at ccc: νₑ itself (Smagorinsky) or κₑ[n] (AMD)
  FT Pr[NH_MAXTR];            // κ on a face = ℑ(kape[n]) / Pr[n]  (Pr = 1 for AMD)
  // WENO(grid) on a stretched z: per-index reconstruction coefficients wz[loc][B-2] (loc 0 = Face, 1 = Center target; B = 2, 3),
  // each (Nz + 2) × (B + 1) × B values [i = 0 … Nz+1][r = −1 … B−1][t]; nullptr ⇒ the uniform coefficients of `tab`
  const FT* wz[2][2];
};

struct FieldExtent { int n[3]; };   // interior size of a field (N, or N+1 on a Bounded Face dim)

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
