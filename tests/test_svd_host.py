"""ti.svd / ti.polar_decompose stand-ins (taichi_b200/csrc/b2_svd.cuh) checked on the host: the header is plain
C++, so the text a run-time struct-for body includes is compiled here with g++ and compared with numpy.linalg,
following the reference's own checks (tests/python/test_svd.py:27-63: U^T U = V^T V = I, U sig V^T = A, tol 1e-5
for f32 / 1e-12 for f64) plus the result convention of the `sifakis_svd` builtin the reference calls for 3x3
(python/taichi/_funcs.py:159-203): det U = det V = +1, |sig| descending, sign of det A in the last value."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "taichi_b200", "csrc")

HARNESS = r"""
#include "b2_svd.cuh"
extern "C" {
void svd3f(const float *A, float *U, float *s, float *V) { b2_svd3<float>(A, U, s, V); }
void svd3d(const double *A, double *U, double *s, double *V) { b2_svd3<double>(A, U, s, V); }
void svd2f(const float *A, float *U, float *s, float *V) { b2_svd2<float>(A, U, s, V); }
void svd2d(const double *A, double *U, double *s, double *V) { b2_svd2<double>(A, U, s, V); }
void polar3f(const float *A, float *R, float *P) { b2_polar3<float>(A, R, P); }
void polar2f(const float *A, float *R, float *P) { b2_polar2<float>(A, R, P); }
}
"""


@pytest.fixture(scope="module")
def H(tmp_path_factory):
    d = tmp_path_factory.mktemp("svd_host")
    src = d / "h.cpp"
    src.write_text(HARNESS)
    so = d / "h.so"
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I", CSRC, str(src), "-o", str(so)], check=True)
    return ctypes.CDLL(str(so))


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _svd(H, fn, A, n, dt):
    A = np.ascontiguousarray(A, dt)
    U, V, s = np.zeros((n, n), dt), np.zeros((n, n), dt), np.zeros(n, dt)
    getattr(H, fn)(_p(A), _p(U), _p(s), _p(V))
    return U, s, V


def _cases(n, rng):
    out = [rng.standard_normal((n, n)) for _ in range(1000)]
    out += [np.eye(n), np.zeros((n, n)), np.ones((n, n)), -np.eye(n), np.diag([3.0, 1e-4, 1e-7][:n]),
            np.diag([1.0, 2.0, 3.0][:n]), np.diag([2.0, 2.0, 2.0][:n])]
    for _ in range(300):                                     # deformation gradients: rotation x mild stretch
        q, _r = np.linalg.qr(rng.standard_normal((n, n)))
        out.append(q @ np.diag(1 + 0.1 * rng.standard_normal(n)))
    for _ in range(50):                                      # rank one
        out.append(np.outer(rng.standard_normal(n), rng.standard_normal(n)))
    return out


@pytest.mark.parametrize("fn,n,dt,tol", [("svd3f", 3, np.float32, 1e-5), ("svd3d", 3, np.float64, 1e-11),
                                         ("svd2f", 2, np.float32, 1e-5), ("svd2d", 2, np.float64, 1e-12)])
def test_svd_matches_numpy(H, fn, n, dt, tol):
    rng = np.random.default_rng(11)
    for A in _cases(n, rng):
        A = A.astype(dt)
        U, s, V = _svd(H, fn, A, n, dt)
        sc = max(1.0, float(np.abs(A).max()))
        assert np.abs(U @ np.diag(s) @ V.T - A).max() <= tol * sc
        assert np.abs(U.T @ U - np.eye(n)).max() <= tol and np.abs(V.T @ V - np.eye(n)).max() <= tol
        assert abs(np.linalg.det(U) - 1) < 1e-4 and abs(np.linalg.det(V) - 1) < 1e-4
        want = np.linalg.svd(A.astype(np.float64), compute_uv=False)
        assert np.abs(np.sort(np.abs(s))[::-1] - want).max() <= tol * sc
        assert np.all(np.abs(s[:-1]) >= np.abs(s[1:]) - tol * sc) and np.all(s[:-1] >= -tol * sc)
        if abs(np.linalg.det(A.astype(np.float64))) > 1e-3:
            assert np.sign(s[-1]) == np.sign(np.linalg.det(A.astype(np.float64)))


@pytest.mark.parametrize("fn,n,dt,tol", [("svd3f", 3, np.float32, 1e-5), ("svd3d", 3, np.float64, 1e-12),
                                         ("svd2f", 2, np.float32, 1e-5), ("svd2d", 2, np.float64, 1e-12)])
def test_reference_svd_case(H, fn, n, dt, tol):
    """The matrices and tolerances of the reference test (test_svd.py:47-63)."""
    A = np.array([[1, 1, 3], [9, -3, 2], [-3, 4, 2]] if n == 3 else [[1, 1], [2, 3]], dt)
    U, s, V = _svd(H, fn, A, n, dt)
    assert np.abs(U.T @ U - np.eye(n)).max() < tol and np.abs(V.T @ V - np.eye(n)).max() < tol
    assert np.abs(U @ np.diag(s) @ V.T - A).max() < tol


def test_polar_decompositions(H):
    rng = np.random.default_rng(5)
    for n, fn in ((3, "polar3f"), (2, "polar2f")):
        for _ in range(300):
            A = rng.standard_normal((n, n)).astype(np.float32)
            if np.linalg.det(A) < 0.05:                      # R is a rotation only for det A > 0 (test_polar_decomp)
                continue
            R, P = np.zeros((n, n), np.float32), np.zeros((n, n), np.float32)
            getattr(H, fn)(_p(A), _p(R), _p(P))
            assert np.abs(R @ P - A).max() < 2e-5 and np.abs(R.T @ R - np.eye(n)).max() < 1e-5
            assert np.abs(P - P.T).max() < 2e-5 and abs(np.linalg.det(R) - 1) < 1e-4
            assert np.all(np.linalg.eigvalsh(0.5 * (P + P.T).astype(np.float64)) > -1e-5)


def test_svd_body_compiles_without_a_device():
    """A run-time body sees b2_svd.cuh: NVRTC compiles a range-for that calls b2_svd3 to an sm_100a cubin."""
    from taichi_b200 import _lib as L
    ok = L.lib.b2_jit_check(b"""B2_RANGE_BODY(c) {
      const float *F = c.arg<const float *>(0) + 9 * c.i; float *o = c.arg<float *>(1) + 21 * c.i;
      float U[9], s[3], V[9]; b2_svd3(F, U, s, V);
      for (int k = 0; k < 9; k++) { o[k] = U[k]; o[12 + k] = V[k]; } for (int k = 0; k < 3; k++) o[9 + k] = s[k]; }""", b"svd_body")
    if ok < 0 and L.last_error()[0] == -1:
        L.lib.ti_set_last_error(0, None)
        pytest.skip("libnvrtc not installed on this host")
    assert ok > 1000, L.last_error()
