"""Pins oracle/mpm_oracle.c against the reference's own regression vector for this path:
tests/python/test_mpm88.py:81-101 (4096 particles, 128^2 grid, 500 substeps; moments of the
positions, rel 1e-2).  CPU only."""
import ctypes as C
import numpy as np

from _desc import load_oracle

ORC = load_oracle()
f32p = C.POINTER(C.c_float)


def fp(a):
    return a.ctypes.data_as(f32p)


def mpm88_init():
    N = 64
    n = N * N
    i = np.arange(n)
    x = np.stack([(i % N) / N * 0.4 + 0.2, i / N / N * 0.4 + 0.05], axis=1).astype(np.float32)
    v = np.tile(np.array([0, -3], np.float32), (n, 1))
    Cm = np.zeros((n, 2, 2), np.float32)
    J = np.ones(n, np.float32)
    return x, v, Cm, J


def test_mpm88_regression_vector():
    x, v, Cm, J = mpm88_init()
    n_grid = 128
    gv = np.zeros((n_grid, n_grid, 2), np.float32)
    gm = np.zeros((n_grid, n_grid), np.float32)
    ORC.orc_mpm88_substeps(n_grid, C.c_int64(len(x)), fp(x), fp(v), fp(Cm), fp(J), C.c_float(2.0e-4),
                           C.c_float(400.0), 500, fp(gv), fp(gm))
    pos = x.copy()
    pos[:, 1] *= 2
    regression = [0.31722742, 0.15826741, 0.10224003, 0.07810827]
    for k in range(4):
        assert abs((pos.astype(np.float64) ** (k + 1)).mean() - regression[k]) <= 1e-2 * regression[k]


def test_svd2d_reconstructs():
    rng = np.random.default_rng(0)
    for _ in range(200):
        A = rng.normal(size=4).astype(np.float32)
        U, s, V = np.zeros(4, np.float32), np.zeros(2, np.float32), np.zeros(4, np.float32)
        ORC.orc_svd2d(fp(A), fp(U), fp(s), fp(V))
        R = U.reshape(2, 2) @ np.diag(s) @ V.reshape(2, 2).T
        assert np.allclose(R, A.reshape(2, 2), atol=2e-5)
        assert s[0] >= s[1]
        assert np.allclose(np.sort(np.abs(s)), np.sort(np.linalg.svd(A.reshape(2, 2).astype(np.float64))[1]), atol=2e-5)


def test_mpm3d_conservation_and_modes():
    rng = np.random.default_rng(2)
    n_grid, N = 32, 4096
    x = (rng.random((N, 3)) * 0.4 + 0.15).astype(np.float32)
    v = np.zeros((N, 3), np.float32)
    Cm = np.zeros((N, 3, 3), np.float32)
    J = np.ones(N, np.float32)
    outs = []
    for mode in (0, 1):
        xs, vs, Cs, Js = x.copy(), v.copy(), Cm.copy(), J.copy()
        gv = np.zeros((n_grid,) * 3 + (3,), np.float32)
        gm = np.zeros((n_grid,) * 3, np.float32)
        pv, pm = np.zeros_like(gv), np.zeros_like(gm)
        ORC.orc_mpm3d_substeps(n_grid, C.c_int64(N), fp(xs), fp(vs), fp(Cs), fp(Js), C.c_float(4e-4),
                               C.c_float(400.0), C.c_float(1.0), C.c_float(9.8), 3, 5, mode, 4,
                               fp(gv), fp(gm), fp(pv), fp(pm))
        outs.append((xs, vs, pm))
        dx = 1.0 / n_grid
        p_mass = (dx * 0.5) ** 2
        assert abs(pm.astype(np.float64).sum() - N * p_mass) < 1e-5 * N * p_mass  # partition of unity
    assert np.allclose(outs[0][0], outs[1][0], rtol=1e-4, atol=1e-6)
    assert np.allclose(outs[0][2], outs[1][2], rtol=1e-4, atol=1e-9)
