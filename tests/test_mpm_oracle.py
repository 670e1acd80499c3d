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


def run_sparse(n_grid, x, v, Cm, J, dt, nsteps, mode, nthreads, deactivate_every, cap=4096):
    xs, vs, Cs, Js = x.copy(), v.copy(), Cm.copy(), J.copy()
    nblk = C.c_int32(0)
    keys = np.zeros(cap, np.int32)
    p2g = np.zeros((cap, 64, 4), np.float32)
    grid = np.zeros((cap, 64, 4), np.float32)
    rc = ORC.orc_mpm3d_sparse_substeps(n_grid, C.c_int64(len(x)), fp(xs), fp(vs), fp(Cs), fp(Js), C.c_float(dt), C.c_float(400.0),
                                       C.c_float(1.0), C.c_float(9.8), 3, nsteps, mode, nthreads, deactivate_every, cap,
                                       C.byref(nblk), keys.ctypes.data_as(C.POINTER(C.c_int32)), fp(p2g), fp(grid))
    assert rc == 0
    n = nblk.value
    order = np.argsort(keys[:n])
    return xs, vs, Cs, Js, keys[:n][order], p2g[:n][order], grid[:n][order]


def dense_blocks(field, keys, nb):
    """[n][n][n][k] dense image -> [len(keys)][64][k] blocks in key order"""
    bx, by, bz = keys // (nb * nb), (keys // nb) % nb, keys % nb
    out = np.zeros((len(keys), 64, field.shape[-1]), field.dtype)
    for t in range(64):
        out[:, t] = field[bx * 4 + (t >> 4), by * 4 + ((t >> 2) & 3), bz * 4 + (t & 3)]
    return out


def test_sparse_port_equals_dense_port():
    """The blocked active-set variant (the honest ti.cpu cost model, also the full-size checker) runs the same
    arithmetic as the dense restatement that the mpm88 vector pins indirectly: single-threaded fp64 accumulation
    is bit-identical; the active set is exactly the set of blocks holding a touched node."""
    rng = np.random.default_rng(7)
    n_grid, N, dt, steps = 32, 5000, 4e-4, 6
    nb = n_grid // 4
    x = (rng.random((N, 3)) * 0.3 + 0.2).astype(np.float32)
    v = ((rng.random((N, 3)) - 0.5) * 2).astype(np.float32)
    Cm = ((rng.random((N, 3, 3)) - 0.5) * 2).astype(np.float32)
    J = (1 + (rng.random(N) - 0.5) * 0.05).astype(np.float32)
    xd, vd, Cd, Jd = x.copy(), v.copy(), Cm.copy(), J.copy()
    gv = np.zeros((n_grid,) * 3 + (3,), np.float32)
    gm = np.zeros((n_grid,) * 3, np.float32)
    pv, pm = np.zeros_like(gv), np.zeros_like(gm)
    ORC.orc_mpm3d_substeps(n_grid, C.c_int64(N), fp(xd), fp(vd), fp(Cd), fp(Jd), C.c_float(dt), C.c_float(400.0), C.c_float(1.0),
                           C.c_float(9.8), 3, steps, 0, 1, fp(gv), fp(gm), fp(pv), fp(pm))
    xs, vs, Cs, Js, keys, p2g, grid = run_sparse(n_grid, x, v, Cm, J, dt, steps, 0, 1, 1)
    assert np.array_equal(xs, xd) and np.array_equal(vs, vd) and np.array_equal(Cs, Cd) and np.array_equal(Js, Jd)
    dense_p2g = dense_blocks(np.concatenate([pv, pm[..., None]], -1), keys, nb)
    assert np.array_equal(p2g, dense_p2g)
    # with deactivate_every = 1 the active set is exactly the blocks the LAST substep's particles touch
    # (positions BEFORE the last substep: rerun steps - 1 substeps)
    xb = run_sparse(n_grid, x, v, Cm, J, dt, steps - 1, 0, 1, 1)[0]
    base = (xb * np.float32(n_grid) - np.float32(0.5)).astype(np.int32)
    want = set()
    for d in np.ndindex(2, 2, 2):
        b = (base + 2 * np.array(d, np.int32)) >> 2
        want.update(((b[:, 0] * nb + b[:, 1]) * nb + b[:, 2]).tolist())
    assert sorted(want) == keys.tolist()
    # never deactivating only grows the set; threaded fp32 CAS mode agrees within the reorder tolerance
    xk, vk, Ck, Jk, keys_keep, _, _ = run_sparse(n_grid, x, v, Cm, J, dt, steps, 1, 4, 0)
    assert set(keys.tolist()) <= set(keys_keep.tolist())
    assert np.allclose(xk, xs, rtol=0, atol=2e-6) and np.allclose(vk, vs, rtol=1e-4, atol=1e-4) and np.allclose(Jk, Js, rtol=1e-5)
    # block capacity overflow is reported, not written out of bounds
    nblk = C.c_int32(0)
    rc = ORC.orc_mpm3d_sparse_substeps(n_grid, C.c_int64(N), fp(x.copy()), fp(v.copy()), fp(Cm.copy()), fp(J.copy()), C.c_float(dt),
                                       C.c_float(400.0), C.c_float(1.0), C.c_float(9.8), 3, 1, 1, 2, 0, 4, C.byref(nblk), None, None, None)
    assert rc == -1
