// b2_mpm_simple.cu -- the "plain atomics" MLS-MPM kernels on ndarrays in the reference's own
// field layouts (x[N][D], v[N][D], C[N][D][D], J[N], grid_v[n^D][D], grid_m[n^D]).
// They back the prebuilt AOT-style modules "builtin:mpm88|mpm3d|mpm128" of the C-API
// (kernel names and argument order of tests/cpp/aot/python_scripts/mpm88_graph_aot.py:29-108,
// client c_api/tests/mpm88_test.cpp:26-88) and are the un-binned baseline the block-local
// kernels of b2_mpm.cu are measured against (reference K15: 27x4 / 9x3 global atomics per
// particle, codegen_llvm.cpp:1495-1569).
// Arithmetic follows mpm3d.py:36-81, test_mpm88.py:34-80 and mpm128.py:28-102 statement by
// statement (fp32, int() = truncation); nvcc contracts a*b+c to FMA as the reference's NVPTX
// backend does (jit_cuda.cpp:107-121).
#include "b2_host.h"
#include "b2_mpm_simple.h"

template <int D>
struct Stencil {
  int base[D];
  float fx[D];
  float w[3][D];
};
template <int D>
__device__ __forceinline__ void make_stencil(const float *xp, float inv_dx_or_dx, bool divide, Stencil<D> &s) {
#pragma unroll
  for (int d = 0; d < D; d++) {
    float X = divide ? xp[d] / inv_dx_or_dx : xp[d] * inv_dx_or_dx;
    s.base[d] = (int)(X - 0.5f);
    s.fx[d] = X - (float)s.base[d];
    s.w[0][d] = 0.5f * (1.5f - s.fx[d]) * (1.5f - s.fx[d]);
    s.w[1][d] = 0.75f - (s.fx[d] - 1.0f) * (s.fx[d] - 1.0f);
    s.w[2][d] = 0.5f * (s.fx[d] - 0.5f) * (s.fx[d] - 0.5f);
  }
}

template <int D>
__global__ void k_simple_reset(float *gv, float *gm, long long cells) {
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < cells; g += (long long)gridDim.x * blockDim.x) {
#pragma unroll
    for (int d = 0; d < D; d++) gv[g * D + d] = 0.f;
    gm[g] = 0.f;
  }
}

// J model (mpm3d.py / test_mpm88.py / mpm88_graph_aot.py)
template <int D>
__global__ void __launch_bounds__(128) k_simple_p2g(B2SimpleParams P, const float *__restrict__ x, const float *__restrict__ v,
                                                     const float *__restrict__ C, const float *__restrict__ J, float *gv, float *gm, long long N) {
  const int n = P.n_grid;
  const float dx = 1.0f / (float)n;
  const float p_vol = (dx * 0.5f) * (dx * 0.5f), p_mass = p_vol * P.p_rho;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < N; p += (long long)gridDim.x * blockDim.x) {
    Stencil<D> s;
    make_stencil<D>(x + p * D, dx, true, s);
    const float stress = -P.dt * 4.0f * P.E * p_vol * (J[p] - 1.0f) / (dx * dx);
    float A[D][D], vp[D];
#pragma unroll
    for (int i = 0; i < D; i++) {
      vp[i] = v[p * D + i];
#pragma unroll
      for (int j = 0; j < D; j++) A[i][j] = (i == j ? stress : 0.0f) + p_mass * C[p * D * D + i * D + j];
    }
    constexpr int NN = D == 2 ? 9 : 27;
#pragma unroll
    for (int q = 0; q < NN; q++) {
      int o[3] = {D == 2 ? q / 3 : q / 9, D == 2 ? q % 3 : (q / 3) % 3, q % 3};
      float dpos[D], weight = 1.0f;
      long long g = 0;
#pragma unroll
      for (int d = 0; d < D; d++) {
        dpos[d] = ((float)o[d] - s.fx[d]) * dx;
        weight *= s.w[o[d]][d];
        g = g * n + (s.base[d] + o[d]);
      }
#pragma unroll
      for (int a = 0; a < D; a++) {
        float Ad = 0.f;
#pragma unroll
        for (int b = 0; b < D; b++) Ad += A[a][b] * dpos[b];
        atomicAdd(&gv[g * D + a], weight * (p_mass * vp[a] + Ad));
      }
      atomicAdd(&gm[g], weight * p_mass);
    }
  }
}

// all_cells: mpm3d.py applies gravity + walls to every cell; mpm88 only where m > 0
template <int D>
__global__ void k_simple_gridop(B2SimpleParams P, float *gv, const float *gm, long long cells, int all_cells) {
  const int n = P.n_grid;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < cells; g += (long long)gridDim.x * blockDim.x) {
    int I[D];
    long long r = g;
#pragma unroll
    for (int d = D - 1; d >= 0; d--) {
      I[d] = (int)(r % n);
      r /= n;
    }
    const float m = gm[g];
    float vv[D];
#pragma unroll
    for (int d = 0; d < D; d++) vv[d] = gv[g * D + d];
    if (all_cells) {
      if (m > 0.f) {
#pragma unroll
        for (int d = 0; d < D; d++) vv[d] /= m;
      }
    } else {
      if (!(m > 0.f)) continue;
      const float inv_m = 1.0f / m;
#pragma unroll
      for (int d = 0; d < D; d++) vv[d] = inv_m * vv[d];
    }
    vv[1] -= P.dt * P.gravity;
#pragma unroll
    for (int d = 0; d < D; d++) {
      if (I[d] < P.bound && vv[d] < 0.f) vv[d] = 0.f;
      if (I[d] > n - P.bound && vv[d] > 0.f) vv[d] = 0.f;
    }
#pragma unroll
    for (int d = 0; d < D; d++) gv[g * D + d] = vv[d];
  }
}

template <int D>
__global__ void __launch_bounds__(128) k_simple_g2p(B2SimpleParams P, float *x, float *v, float *C, float *J, const float *__restrict__ gv,
                                                     float *pos3, long long N) {
  const int n = P.n_grid;
  const float dx = 1.0f / (float)n;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < N; p += (long long)gridDim.x * blockDim.x) {
    Stencil<D> s;
    make_stencil<D>(x + p * D, dx, true, s);
    float nv[D], nC[D][D];
#pragma unroll
    for (int a = 0; a < D; a++) {
      nv[a] = 0.f;
#pragma unroll
      for (int b = 0; b < D; b++) nC[a][b] = 0.f;
    }
    constexpr int NN = D == 2 ? 9 : 27;
#pragma unroll
    for (int q = 0; q < NN; q++) {
      int o[3] = {D == 2 ? q / 3 : q / 9, D == 2 ? q % 3 : (q / 3) % 3, q % 3};
      float dpos[D], weight = 1.0f;
      long long g = 0;
#pragma unroll
      for (int d = 0; d < D; d++) {
        dpos[d] = ((float)o[d] - s.fx[d]) * dx;
        weight *= s.w[o[d]][d];
        g = g * n + (s.base[d] + o[d]);
      }
#pragma unroll
      for (int a = 0; a < D; a++) {
        const float ga = gv[g * D + a];
        nv[a] += weight * ga;
#pragma unroll
        for (int b = 0; b < D; b++) nC[a][b] += 4.0f * weight * (ga * dpos[b]) / (dx * dx);
      }
    }
    float tr = 0.f;
#pragma unroll
    for (int a = 0; a < D; a++) {
      v[p * D + a] = nv[a];
      const float xn = x[p * D + a] + P.dt * nv[a];
      x[p * D + a] = xn;
      if (pos3) pos3[p * 3 + a] = xn;
      tr += nC[a][a];
#pragma unroll
      for (int b = 0; b < D; b++) C[p * D * D + a * D + b] = nC[a][b];
    }
    if (pos3 && D == 2) pos3[p * 3 + 2] = 0.f;
    J[p] *= 1.0f + P.dt * tr;
  }
}

// ------------------------- mpm128 (2-D, F + SVD model) -------------------------------------
// ti.polar_decompose / ti.svd for 2x2: python/taichi/_funcs.py:53-157
__device__ __forceinline__ void polar2d(const float A[4], float U[4], float Pm[4]) {
  U[0] = 1.f, U[1] = 0.f, U[2] = 0.f, U[3] = 1.f;
  Pm[0] = A[0], Pm[1] = A[1], Pm[2] = A[2], Pm[3] = A[3];
  if (A[0] == 0.f && A[1] == 0.f && A[2] == 0.f && A[3] == 0.f) return;
  const float detA = A[0] * A[3] - A[2] * A[1];
  const float adetA = fabsf(detA);
  float B[4] = {A[0] + A[3], A[1] - A[2], A[2] - A[1], A[3] + A[0]};
  if (detA < 0.f) {
    B[0] = A[0] - A[3], B[1] = A[1] + A[2], B[2] = A[2] + A[1], B[3] = A[3] - A[0];
  }
  const float adetB = fabsf(B[0] * B[3] - B[2] * B[1]);
  const float k = 1.0f / sqrtf(adetB);
#pragma unroll
  for (int i = 0; i < 4; i++) U[i] = B[i] * k;
  Pm[0] = (A[0] * A[0] + A[2] * A[2] + adetA) * k;
  Pm[1] = (A[0] * A[1] + A[2] * A[3]) * k;
  Pm[2] = (A[1] * A[0] + A[3] * A[2]) * k;
  Pm[3] = (A[1] * A[1] + A[3] * A[3] + adetA) * k;
}
__device__ __forceinline__ void svd2d(const float A[4], float U[4], float sig[2], float V[4]) {
  float R[4], S[4];
  polar2d(A, R, S);
  float c, s, s1, s2;
  if (fabsf(S[1]) < 1e-5f) {
    c = 1.f, s = 0.f, s1 = S[0], s2 = S[3];
  } else {
    const float tao = 0.5f * (S[0] - S[3]);
    const float w = sqrtf(tao * tao + S[1] * S[1]);
    const float t = tao > 0.f ? S[1] / (tao + w) : S[1] / (tao - w);
    c = 1.0f / sqrtf(t * t + 1.0f);
    s = -t * c;
    s1 = c * c * S[0] - 2.f * c * s * S[1] + s * s * S[3];
    s2 = s * s * S[0] + 2.f * c * s * S[1] + c * c * S[3];
  }
  if (s1 < s2) {
    const float tmp = s1;
    s1 = s2, s2 = tmp;
    V[0] = -s, V[1] = c, V[2] = -c, V[3] = -s;
  } else {
    V[0] = c, V[1] = s, V[2] = -s, V[3] = c;
  }
  U[0] = R[0] * V[0] + R[1] * V[2], U[1] = R[0] * V[1] + R[1] * V[3];
  U[2] = R[2] * V[0] + R[3] * V[2], U[3] = R[2] * V[1] + R[3] * V[3];
  sig[0] = s1, sig[1] = s2;
}
__device__ __forceinline__ void mm2(const float a[4], const float b[4], float o[4]) {
  o[0] = a[0] * b[0] + a[1] * b[2], o[1] = a[0] * b[1] + a[1] * b[3];
  o[2] = a[2] * b[0] + a[3] * b[2], o[3] = a[2] * b[1] + a[3] * b[3];
}

__global__ void __launch_bounds__(128) k_mpm128_p2g(B2SimpleParams P, const float *__restrict__ x, const float *__restrict__ v, const float *__restrict__ C,
                                                     float *F, const int *__restrict__ material, float *Jp, float *gv, float *gm, long long N) {
  const int n = P.n_grid;
  const float dx = 1.0f / (float)n, inv_dx = (float)n;
  const float p_vol = (dx * 0.5f) * (dx * 0.5f), p_mass = p_vol * P.p_rho;
  const float mu_0 = P.E / (2.f * (1.f + P.nu)), lambda_0 = P.E * P.nu / ((1.f + P.nu) * (1.f - 2.f * P.nu));
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < N; p += (long long)gridDim.x * blockDim.x) {
    Stencil<2> s;
    make_stencil<2>(x + p * 2, inv_dx, false, s);
    float Fp[4], Cp[4], Fn[4];
#pragma unroll
    for (int i = 0; i < 4; i++) Fp[i] = F[p * 4 + i], Cp[i] = C[p * 4 + i];
    const float M[4] = {1.0f + P.dt * Cp[0], P.dt * Cp[1], P.dt * Cp[2], 1.0f + P.dt * Cp[3]};
    mm2(M, Fp, Fn);
#pragma unroll
    for (int i = 0; i < 4; i++) Fp[i] = Fn[i];
    const int mat = material[p];
    float jp = Jp[p];
    float h = fmaxf(0.1f, fminf(5.0f, expf(10.0f * (1.0f - jp))));
    if (mat == 1) h = 0.3f;
    float mu = mu_0 * h;
    const float la = lambda_0 * h;
    if (mat == 0) mu = 0.0f;
    float U[4], sig[2], V[4];
    svd2d(Fp, U, sig, V);
    float Jd = 1.0f;
#pragma unroll
    for (int d = 0; d < 2; d++) {
      float new_sig = sig[d];
      if (mat == 2) new_sig = fminf(fmaxf(sig[d], 1.f - 2.5e-2f), 1.f + 4.5e-3f);
      jp *= sig[d] / new_sig;
      sig[d] = new_sig;
      Jd *= new_sig;
    }
    const float Vt[4] = {V[0], V[2], V[1], V[3]};
    if (mat == 0) {
      const float sq = sqrtf(Jd);
      Fp[0] = sq, Fp[1] = 0.f, Fp[2] = 0.f, Fp[3] = sq;
    } else if (mat == 2) {
      const float Us[4] = {U[0] * sig[0], U[1] * sig[1], U[2] * sig[0], U[3] * sig[1]};
      mm2(Us, Vt, Fp);
    }
    float UVt[4], Dm[4], DFt[4];
    const float Ft[4] = {Fp[0], Fp[2], Fp[1], Fp[3]};
    mm2(U, Vt, UVt);
#pragma unroll
    for (int i = 0; i < 4; i++) Dm[i] = Fp[i] - UVt[i];
    mm2(Dm, Ft, DFt);
    const float k = -P.dt * p_vol * 4.f * inv_dx * inv_dx;
    float A[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const float st = 2.f * mu * DFt[i] + ((i == 0 || i == 3) ? la * Jd * (Jd - 1.f) : 0.0f);
      A[i] = k * st + p_mass * Cp[i];
    }
#pragma unroll
    for (int i = 0; i < 4; i++) F[p * 4 + i] = Fp[i];
    Jp[p] = jp;
    const float v0 = v[p * 2], v1 = v[p * 2 + 1];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) {
        const float d0 = ((float)i - s.fx[0]) * dx, d1 = ((float)j - s.fx[1]) * dx;
        const float weight = s.w[i][0] * s.w[j][1];
        const long long g = (long long)(s.base[0] + i) * n + (s.base[1] + j);
        atomicAdd(&gv[g * 2 + 0], weight * (p_mass * v0 + (A[0] * d0 + A[1] * d1)));
        atomicAdd(&gv[g * 2 + 1], weight * (p_mass * v1 + (A[2] * d0 + A[3] * d1)));
        atomicAdd(&gm[g], weight * p_mass);
      }
  }
}
__global__ void k_mpm128_gridop(B2SimpleParams P, float *gv, const float *gm) {
  const int n = P.n_grid;
  const float dx = 1.0f / (float)n;
  const long long cells = (long long)n * n;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < cells; g += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(g / n), j = (int)(g % n);
    const float m = gm[g];
    if (!(m > 0.f)) continue;
    const float im = 1.f / m;
    float g0 = im * gv[g * 2], g1 = im * gv[g * 2 + 1];
    g0 += P.dt * P.gx * 30.f, g1 += P.dt * P.gy * 30.f;
    const float d0 = P.ax - dx * (float)i, d1 = P.ay - dx * (float)j;
    const float nrm = sqrtf(d0 * d0 + d1 * d1);
    g0 += d0 / (0.01f + nrm) * P.astrength * P.dt * 100.f;
    g1 += d1 / (0.01f + nrm) * P.astrength * P.dt * 100.f;
    if (i < 3 && g0 < 0.f) g0 = 0.f;
    if (i > n - 3 && g0 > 0.f) g0 = 0.f;
    if (j < 3 && g1 < 0.f) g1 = 0.f;
    if (j > n - 3 && g1 > 0.f) g1 = 0.f;
    gv[g * 2] = g0, gv[g * 2 + 1] = g1;
  }
}
__global__ void __launch_bounds__(128) k_mpm128_g2p(B2SimpleParams P, float *x, float *v, float *C, const float *__restrict__ gv, long long N) {
  const int n = P.n_grid;
  const float inv_dx = (float)n;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < N; p += (long long)gridDim.x * blockDim.x) {
    Stencil<2> s;
    make_stencil<2>(x + p * 2, inv_dx, false, s);
    float nv[2] = {0.f, 0.f}, nC[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) {
        const float dpos[2] = {(float)i - s.fx[0], (float)j - s.fx[1]};
        const long long g = (long long)(s.base[0] + i) * n + (s.base[1] + j);
        const float weight = s.w[i][0] * s.w[j][1];
#pragma unroll
        for (int a = 0; a < 2; a++) {
          const float ga = gv[g * 2 + a];
          nv[a] += weight * ga;
#pragma unroll
          for (int b = 0; b < 2; b++) nC[a * 2 + b] += 4.f * inv_dx * weight * (ga * dpos[b]);
        }
      }
    v[p * 2] = nv[0], v[p * 2 + 1] = nv[1];
#pragma unroll
    for (int i = 0; i < 4; i++) C[p * 4 + i] = nC[i];
    x[p * 2] += P.dt * nv[0], x[p * 2 + 1] += P.dt * nv[1];
  }
}

// test_mpm88.py:81-84 / mpm88_graph_aot.py init_particles use ti.random (thread-mapping dependent,
// runtime.cpp:1790-1807); here a counter-based hash so that results do not depend on the launch shape
__device__ __forceinline__ float hash01(unsigned a, unsigned b) {
  unsigned h = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u) * 0x85EBCA77u;
  h ^= h >> 15;
  h *= 0x2C1B3C6Du;
  h ^= h >> 12;
  h *= 0x297A2D39u;
  h ^= h >> 15;
  return (float)(h >> 8) * (1.0f / 16777216.0f);
}
__global__ void k_init_particles(float *x, float *v, float *J, long long N, int D, float lo, float ext, float vy) {
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < N; p += (long long)gridDim.x * blockDim.x) {
    for (int d = 0; d < D; d++) {
      x[p * D + d] = hash01((unsigned)p, (unsigned)d) * ext + lo;
      if (v) v[p * D + d] = d == 1 ? vy : 0.f;
    }
    if (J) J[p] = 1.f;
  }
}

// ------------------------------------------------------------------------------------------
static int grid1d(TiRuntime_t *rt, long long n, int block) { return rt->grid_for(n, block, 16); }

void b2_simple_launch(TiRuntime_t *rt, int model, int kernel, const B2SimpleParams &P, const B2SimpleArgs &a) {
  cudaStream_t st = rt->stream();
  const int D = model == B2_MODEL_MPM3D ? 3 : 2;
  long long cells = 1;
  for (int d = 0; d < D; d++) cells *= P.n_grid;
  const long long N = a.n_particles;
  switch (kernel) {
    case B2_K_RESET:
      if (D == 2)
        k_simple_reset<2><<<grid1d(rt, cells, 256), 256, 0, st>>>(a.grid_v, a.grid_m, cells);
      else
        k_simple_reset<3><<<grid1d(rt, cells, 256), 256, 0, st>>>(a.grid_v, a.grid_m, cells);
      break;
    case B2_K_P2G:
      if (N == 0) return;
      if (model == B2_MODEL_MPM128)
        k_mpm128_p2g<<<grid1d(rt, N, 128), 128, 0, st>>>(P, a.x, a.v, a.C, a.F, a.material, a.Jp, a.grid_v, a.grid_m, N);
      else if (D == 2)
        k_simple_p2g<2><<<grid1d(rt, N, 128), 128, 0, st>>>(P, a.x, a.v, a.C, a.J, a.grid_v, a.grid_m, N);
      else
        k_simple_p2g<3><<<grid1d(rt, N, 128), 128, 0, st>>>(P, a.x, a.v, a.C, a.J, a.grid_v, a.grid_m, N);
      break;
    case B2_K_GRIDOP:
      if (model == B2_MODEL_MPM128)
        k_mpm128_gridop<<<grid1d(rt, cells, 256), 256, 0, st>>>(P, a.grid_v, a.grid_m);
      else if (D == 2)
        k_simple_gridop<2><<<grid1d(rt, cells, 256), 256, 0, st>>>(P, a.grid_v, a.grid_m, cells, 0);
      else
        k_simple_gridop<3><<<grid1d(rt, cells, 256), 256, 0, st>>>(P, a.grid_v, a.grid_m, cells, 1);
      break;
    case B2_K_G2P:
      if (N == 0) return;
      if (model == B2_MODEL_MPM128)
        k_mpm128_g2p<<<grid1d(rt, N, 128), 128, 0, st>>>(P, a.x, a.v, a.C, a.grid_v, N);
      else if (D == 2)
        k_simple_g2p<2><<<grid1d(rt, N, 128), 128, 0, st>>>(P, a.x, a.v, a.C, a.J, a.grid_v, a.pos3, N);
      else
        k_simple_g2p<3><<<grid1d(rt, N, 128), 128, 0, st>>>(P, a.x, a.v, a.C, a.J, a.grid_v, a.pos3, N);
      break;
    case B2_K_INIT:
      if (N == 0) return;
      if (model == B2_MODEL_MPM3D)
        k_init_particles<<<grid1d(rt, N, 256), 256, 0, st>>>(a.x, a.v, a.J, N, 3, 0.15f, 0.4f, 0.f);
      else
        k_init_particles<<<grid1d(rt, N, 256), 256, 0, st>>>(a.x, a.v, a.J, N, 2, 0.2f, 0.4f, -1.f);
      break;
    default: throw B2Error(TI_ERROR_NAME_NOT_FOUND, "unknown builtin kernel");
  }
  rt->count_launch();
  B2_CUDA(cudaGetLastError());
}
