// b2_svd.cuh -- 2x2 / 3x3 singular value and polar decompositions for struct-for bodies.
//
// Stands in for ti.svd / ti.polar_decompose (reference python/taichi/_funcs.py:45-203: 2-D closed form through the
// polar factor, 3-D through the `sifakis_svd` builtin) with the same result convention: A = U diag(sig) V^T,
// det U = det V = +1, |sig| sorted descending, the sign of det A carried by the LAST singular value.
// The algorithm is not the reference's approximate-Givens one: V comes from cyclic Jacobi sweeps with exact
// rotations on A^T A, U and sig from a Gram-Schmidt QR of A V, which keeps sig accurate to eps * |A|.
// Matrices are row-major T[9] / T[4].  Plain C++ (no CUDA types) so the same text compiles under NVRTC inside a
// run-time body, under nvcc, and under g++ for the host-side check in tests/test_svd_host.py.
#pragma once
#if defined(__CUDACC__) || defined(__CUDACC_RTC__)
#define B2_MATH __host__ __device__ __forceinline__
#else
#include <math.h>
#define B2_MATH static inline
#endif

B2_MATH float b2_sqrt(float x) { return sqrtf(x); }
B2_MATH double b2_sqrt(double x) { return sqrt(x); }
template <typename T> B2_MATH T b2_abs(T x) { return x < (T)0 ? -x : x; }

// One Jacobi rotation in the (P, Q) plane: S <- J^T S J (S symmetric, full 3x3 storage), V <- V J.
template <int P, int Q, typename T> B2_MATH void b2_jacobi_rot(T *S, T *V) {
  const T spq = S[P * 3 + Q];
  const T spp = S[P * 3 + P], sqq = S[Q * 3 + Q];
  if (b2_abs(spq) <= (T)1e-30 + (T)1e-12 * (b2_abs(spp) + b2_abs(sqq))) return;
  const T theta = (sqq - spp) / ((T)2 * spq);
  const T t = (theta >= (T)0 ? (T)1 : (T)-1) / (b2_abs(theta) + b2_sqrt(theta * theta + (T)1));
  const T c = (T)1 / b2_sqrt(t * t + (T)1), s = t * c;
  constexpr int R = 3 - P - Q;  // the untouched axis
  const T srp = S[R * 3 + P], srq = S[R * 3 + Q];
  S[P * 3 + P] = spp - t * spq;
  S[Q * 3 + Q] = sqq + t * spq;
  S[P * 3 + Q] = S[Q * 3 + P] = (T)0;
  S[R * 3 + P] = S[P * 3 + R] = c * srp - s * srq;
  S[R * 3 + Q] = S[Q * 3 + R] = s * srp + c * srq;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const T vp = V[k * 3 + P], vq = V[k * 3 + Q];
    V[k * 3 + P] = c * vp - s * vq;
    V[k * 3 + Q] = s * vp + c * vq;
  }
}

// Swap columns I and J of B and V, negating one of them so that det V stays +1.
template <int I, int J, typename T> B2_MATH void b2_swap_cols(T *B, T *V, T *n2) {
  if (n2[I] >= n2[J]) return;
  T t = n2[I]; n2[I] = n2[J]; n2[J] = t;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    t = B[k * 3 + I]; B[k * 3 + I] = B[k * 3 + J]; B[k * 3 + J] = -t;
    t = V[k * 3 + I]; V[k * 3 + I] = V[k * 3 + J]; V[k * 3 + J] = -t;
  }
}

// A = U diag(sig) V^T.  `sweeps` Jacobi sweeps (5 for fp32 and 8 for fp64 reach working precision; the reference's
// default iteration counts are the same numbers, _funcs.py:170-174).
template <typename T> B2_MATH void b2_svd3(const T *A, T *U, T *sig, T *V, int sweeps = sizeof(T) == 4 ? 5 : 8) {
  T S[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) S[i * 3 + j] = A[0 + i] * A[0 + j] + A[3 + i] * A[3 + j] + A[6 + i] * A[6 + j];
#pragma unroll
  for (int i = 0; i < 9; ++i) V[i] = (i % 4 == 0) ? (T)1 : (T)0;
  for (int s = 0; s < sweeps; ++s) {
    b2_jacobi_rot<0, 1>(S, V);
    b2_jacobi_rot<0, 2>(S, V);
    b2_jacobi_rot<1, 2>(S, V);
  }
  T B[9], n2[3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) B[i * 3 + j] = A[i * 3 + 0] * V[0 + j] + A[i * 3 + 1] * V[3 + j] + A[i * 3 + 2] * V[6 + j];
#pragma unroll
  for (int j = 0; j < 3; ++j) n2[j] = B[j] * B[j] + B[3 + j] * B[3 + j] + B[6 + j] * B[6 + j];
  b2_swap_cols<0, 1>(B, V, n2);
  b2_swap_cols<0, 2>(B, V, n2);
  b2_swap_cols<1, 2>(B, V, n2);
  // Gram-Schmidt QR of B (its columns are orthogonal up to rounding): u0 | b0, u1 | b1 - (u0.b1) u0, u2 = u0 x u1.
  T u0[3] = {B[0], B[3], B[6]}, u1[3] = {B[1], B[4], B[7]};
  const T tiny = sizeof(T) == 4 ? (T)1e-18 : (T)1e-150;
  T l0 = b2_sqrt(n2[0]);
  if (l0 > tiny) {
    const T r = (T)1 / l0;
    u0[0] *= r; u0[1] *= r; u0[2] *= r;
  } else {
    u0[0] = (T)1; u0[1] = (T)0; u0[2] = (T)0;
  }
  T d = u0[0] * u1[0] + u0[1] * u1[1] + u0[2] * u1[2];
  u1[0] -= d * u0[0]; u1[1] -= d * u0[1]; u1[2] -= d * u0[2];
  T l1 = b2_sqrt(u1[0] * u1[0] + u1[1] * u1[1] + u1[2] * u1[2]);
  if (l1 <= (T)(sizeof(T) == 4 ? 1e-6 : 1e-14) * l0 || l1 <= tiny) {
    // rank <= 1: any unit vector orthogonal to u0 -- cross it with the axis it is least aligned with
    const T ax = b2_abs(u0[0]), ay = b2_abs(u0[1]), az = b2_abs(u0[2]);
    T e[3] = {(T)0, (T)0, (T)0};
    e[(ax <= ay && ax <= az) ? 0 : (ay <= az ? 1 : 2)] = (T)1;
    u1[0] = u0[1] * e[2] - u0[2] * e[1];
    u1[1] = u0[2] * e[0] - u0[0] * e[2];
    u1[2] = u0[0] * e[1] - u0[1] * e[0];
    l1 = b2_sqrt(u1[0] * u1[0] + u1[1] * u1[1] + u1[2] * u1[2]);
  }
  {
    const T r = (T)1 / l1;
    u1[0] *= r; u1[1] *= r; u1[2] *= r;
  }
  const T u2[3] = {u0[1] * u1[2] - u0[2] * u1[1], u0[2] * u1[0] - u0[0] * u1[2], u0[0] * u1[1] - u0[1] * u1[0]};
  sig[0] = u0[0] * B[0] + u0[1] * B[3] + u0[2] * B[6];
  sig[1] = u1[0] * B[1] + u1[1] * B[4] + u1[2] * B[7];
  sig[2] = u2[0] * B[2] + u2[1] * B[5] + u2[2] * B[8];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    U[k * 3 + 0] = u0[k];
    U[k * 3 + 1] = u1[k];
    U[k * 3 + 2] = u2[k];
  }
}

// A = R P with R a rotation and P = V diag(sig) V^T symmetric (reference _polar_decompose3d, _funcs.py:98-112).
template <typename T> B2_MATH void b2_polar3(const T *A, T *R, T *P) {
  T U[9], s[3], V[9];
  b2_svd3(A, U, s, V);
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      R[i * 3 + j] = U[i * 3 + 0] * V[j * 3 + 0] + U[i * 3 + 1] * V[j * 3 + 1] + U[i * 3 + 2] * V[j * 3 + 2];
      P[i * 3 + j] = V[i * 3 + 0] * s[0] * V[j * 3 + 0] + V[i * 3 + 1] * s[1] * V[j * 3 + 1] + V[i * 3 + 2] * s[2] * V[j * 3 + 2];
    }
}

// 2x2: A = R P (R rotation, P symmetric; reference _polar_decompose2d, _funcs.py:45-95).
template <typename T> B2_MATH void b2_polar2(const T *A, T *R, T *P) {
  const T x = A[0] + A[3], y = A[2] - A[1];
  const T scale = b2_sqrt(x * x + y * y);
  T c = (T)1, s = (T)0;
  if (scale > (T)0) {
    c = x / scale;
    s = y / scale;
  }
  R[0] = c; R[1] = -s; R[2] = s; R[3] = c;
  P[0] = c * A[0] + s * A[2];
  P[1] = c * A[1] + s * A[3];
  P[2] = -s * A[0] + c * A[2];
  P[3] = -s * A[1] + c * A[3];
}

// 2x2: A = U diag(sig) V^T, sig[0] >= sig[1] (sig[1] < 0 when det A < 0); reference _svd2d, _funcs.py:115-156.
template <typename T> B2_MATH void b2_svd2(const T *A, T *U, T *sig, T *V) {
  T R[4], P[4];
  b2_polar2(A, R, P);
  T c = (T)1, s = (T)0, s1 = P[0], s2 = P[3];
  const T off = (T)0.5 * (P[1] + P[2]);
  if (b2_abs(off) > (T)1e-30 + (T)(sizeof(T) == 4 ? 1e-9 : 1e-18) * (b2_abs(s1) + b2_abs(s2))) {
    const T tau = (T)0.5 * (P[0] - P[3]);
    const T w = b2_sqrt(tau * tau + off * off);
    const T t = tau > (T)0 ? off / (tau + w) : off / (tau - w);
    c = (T)1 / b2_sqrt(t * t + (T)1);
    s = -t * c;
    s1 = c * c * P[0] - (T)2 * c * s * off + s * s * P[3];
    s2 = s * s * P[0] + (T)2 * c * s * off + c * c * P[3];
  }
  if (s1 < s2) {
    const T t = s1; s1 = s2; s2 = t;
    V[0] = -s; V[1] = c; V[2] = -c; V[3] = -s;
  } else {
    V[0] = c; V[1] = s; V[2] = -s; V[3] = c;
  }
  sig[0] = s1;
  sig[1] = s2;
  U[0] = R[0] * V[0] + R[1] * V[2];
  U[1] = R[0] * V[1] + R[1] * V[3];
  U[2] = R[2] * V[0] + R[3] * V[2];
  U[3] = R[2] * V[1] + R[3] * V[3];
}
