/* abi_smoke.c — the C ABI of libnhstep.so driven from plain C: fills nh_grid_desc / nh_model_desc by hand, checks the struct
 * layout the Julia shim (julia/OceananigansB200Ext.jl) and the ctypes binding assume, and — when a CUDA device is present —
 * sets and steps a 16^3 triply periodic WENO5 model.  Built and run by tests/test_abi_c.py.
 *   usage: abi_smoke layout   → prints sizes / offsets, needs no device
 *          abi_smoke step     → creates, sets, steps; prints max |div U| and the clock */
#include <math.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "nhstep.h"

#define CHECK(call) do { if ((call) != 0) { fprintf(stderr, "%s failed: %s\n", #call, nh_last_error()); return 2; } } while (0)

static int layout(void) {
  printf("sizeof(nh_grid_desc)=%zu\n", sizeof(nh_grid_desc));
  printf("offsetof(nh_grid_desc,x)=%zu\n", offsetof(nh_grid_desc, x));
  printf("offsetof(nh_grid_desc,z_faces)=%zu\n", offsetof(nh_grid_desc, z_faces));
  printf("sizeof(nh_model_desc)=%zu\n", sizeof(nh_model_desc));
  printf("offsetof(nh_model_desc,ab2_chi)=%zu\n", offsetof(nh_model_desc, ab2_chi));
  printf("offsetof(nh_model_desc,n_tracers)=%zu\n", offsetof(nh_model_desc, n_tracers));
  printf("offsetof(nh_model_desc,gravitational_acceleration)=%zu\n", offsetof(nh_model_desc, gravitational_acceleration));
  printf("offsetof(nh_model_desc,coriolis_f)=%zu\n", offsetof(nh_model_desc, coriolis_f));
  printf("offsetof(nh_model_desc,nu)=%zu\n", offsetof(nh_model_desc, nu));
  printf("offsetof(nh_model_desc,kappa)=%zu\n", offsetof(nh_model_desc, kappa));
  printf("offsetof(nh_model_desc,vertically_implicit)=%zu\n", offsetof(nh_model_desc, vertically_implicit));
  printf("offsetof(nh_model_desc,smagorinsky_C)=%zu\n", offsetof(nh_model_desc, smagorinsky_C));
  printf("offsetof(nh_model_desc,amd_Cnu)=%zu\n", offsetof(nh_model_desc, amd_Cnu));
  printf("offsetof(nh_model_desc,hydrostatic_pressure_anomaly)=%zu\n", offsetof(nh_model_desc, hydrostatic_pressure_anomaly));
  printf("offsetof(nh_model_desc,exact_arithmetic)=%zu\n", offsetof(nh_model_desc, exact_arithmetic));
  printf("sizeof(nh_clock)=%zu\n", sizeof(nh_clock));
  /* the host-only entry point works without a device: Centered(4) coefficients in Float32 (reconstruction_coefficients.jl:154-171) */
  double c[8]; int32_t n = 0;
  CHECK(nh_reconstruction_coefficients(NH_F32, 0, 2, 0, c, &n));
  printf("centered4_f32=%.9g,%.9g,%.9g,%.9g\n", c[0], c[1], c[2], c[3]);
  return 0;
}

static int step(void) {
  enum { N = 16 };
  nh_grid_desc gd;
  memset(&gd, 0, sizeof gd);
  gd.float_type = NH_F64;
  for (int e = 0; e < 3; ++e) { gd.N[e] = N; gd.H[e] = 3; gd.topology[e] = NH_PERIODIC; }
  gd.x[0] = gd.y[0] = gd.z[0] = 0.0;
  gd.x[1] = gd.y[1] = gd.z[1] = 1.0;
  nh_grid* grid = NULL;
  if (nh_grid_create(&gd, 0, &grid) != 0) {
    /* no CPU fallback: without a device the create call must fail, loudly */
    printf("no_device: %s\n", nh_last_error());
    return 3;
  }
  nh_model_desc md;
  memset(&md, 0, sizeof md);
  for (int e = 0; e < 3; ++e) { md.advection_kind[e] = NH_ADV_WENO; md.advection_order[e] = 5; }
  md.timestepper = NH_RK3;
  md.ab2_chi = 0.1;
  md.hydrostatic_pressure_anomaly = -1;
  md.exact_arithmetic = 1;
  nh_model* m = NULL;
  CHECK(nh_model_create(grid, &md, &m));
  double* a = (double*)malloc(sizeof(double) * N * N * N);
  const double pi = 3.14159265358979323846;
  for (int f = 0; f < 3; ++f) {
    for (int k = 0; k < N; ++k) for (int j = 0; j < N; ++j) for (int i = 0; i < N; ++i) {
      const double x = (i + 0.5) / N, y = (j + 0.5) / N, z = (k + 0.5) / N;
      a[i + N * (j + N * k)] = f == 0 ? sin(2 * pi * x) * cos(2 * pi * y) : f == 1 ? -cos(2 * pi * x) * sin(2 * pi * y) + 0.1 * sin(2 * pi * z)
                                                                                  : 0.1 * cos(2 * pi * x);
    }
    CHECK(nh_set_field(m, f, NH_INTERIOR, a, (int64_t)N * N * N, NH_HOST));
    CHECK(nh_fill_halo_regions(m, 1u << f, 1));
  }
  CHECK(nh_set_state_finalize(m, 1));
  CHECK(nh_time_step(m, 1e-3, 0, 3));
  double div = -1;
  CHECK(nh_max_abs_divergence(m, &div));
  nh_clock c;
  CHECK(nh_clock_get(m, &c));
  int64_t sz[3];
  CHECK(nh_field_size(m, NH_FIELD_U, NH_PARENT, sz));
  CHECK(nh_get_field(m, NH_FIELD_U, NH_INTERIOR, a, (int64_t)N * N * N, NH_HOST));
  double s = 0;
  for (int q = 0; q < N * N * N; ++q) s += a[q] * a[q];
  printf("stepped: iteration=%lld time=%.17g max_div=%.3e parent=%lldx%lldx%lld u_rms=%.6f\n", (long long)c.iteration, c.time, div,
         (long long)sz[0], (long long)sz[1], (long long)sz[2], sqrt(s / (N * N * N)));
  free(a);
  nh_model_destroy(m);
  nh_grid_destroy(grid);
  return (c.iteration == 3 && div < 1e-12 && s > 0) ? 0 : 4;
}

int main(int argc, char** argv) {
  if (argc > 1 && strcmp(argv[1], "step") == 0) return step();
  return layout();
}
