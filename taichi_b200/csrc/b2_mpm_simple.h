// b2_mpm_simple.h -- launch table of the plain-atomics MLS-MPM kernels (see b2_mpm_simple.cu)
#pragma once
#include <stdint.h>

enum { B2_MODEL_MPM88 = 0, B2_MODEL_MPM3D = 1, B2_MODEL_MPM128 = 2 };
enum { B2_K_RESET = 0, B2_K_P2G = 1, B2_K_GRIDOP = 2, B2_K_G2P = 3, B2_K_INIT = 4 };

struct B2SimpleParams {
  int n_grid;
  float dt, E, p_rho, gravity;
  int bound;
  // mpm128 only
  float nu, gx, gy, ax, ay, astrength;
};
struct B2SimpleArgs {
  long long n_particles = 0;
  float *x = nullptr, *v = nullptr, *C = nullptr, *J = nullptr, *F = nullptr, *Jp = nullptr, *pos3 = nullptr;
  int *material = nullptr;
  float *grid_v = nullptr, *grid_m = nullptr;
};
struct TiRuntime_t;
void b2_simple_launch(TiRuntime_t *rt, int model, int kernel, const B2SimpleParams &P, const B2SimpleArgs &a);
