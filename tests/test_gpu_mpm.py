"""Parity of the CUDA MLS-MPM kernels (through the C ABI) against oracle/mpm_oracle.c on the
same seeded inputs.  Tolerances are stated per test: fp32 atomics reorder sums, and the
block-local kernels re-associate the affine term (b2_mpm.cu), so grid values are compared
per node against the oracle's fp64-accumulated P2G, particles after N substeps relative to
the state's scale.  Active block sets are compared bit-exactly.
Reference: mpm3d.py:36-81, test_mpm88.py:34-101, mpm128.py:28-102, bls_test_template.py:85-96,258-272."""
import ctypes as C
import numpy as np
import pytest

from _desc import load_oracle
from _b2 import have_gpu, runtime

pytestmark = pytest.mark.gpu
ORC = load_oracle()
f32p = C.POINTER(C.c_float)


def fp(a):
    return a.ctypes.data_as(f32p)


@pytest.fixture(scope="module")
def rt():
    if not have_gpu():
        pytest.skip("no CUDA device")
    import taichi_b200 as tb
    r = tb.Runtime.__new__(tb.Runtime)
    r.handle = runtime()
    r.memories = []
    return r


def oracle_mpm3d(n_grid, x, v, Cm, J, dt, nsteps, E=400.0, mode=0):
    xs, vs, Cs, Js = x.copy(), v.copy(), Cm.copy(), J.copy()
    gv = np.zeros((n_grid,) * 3 + (3,), np.float32)
    gm = np.zeros((n_grid,) * 3, np.float32)
    pv, pm = np.zeros_like(gv), np.zeros_like(gm)
    ORC.orc_mpm3d_substeps(n_grid, C.c_int64(len(x)), fp(xs), fp(vs), fp(Cs), fp(Js), C.c_float(dt), C.c_float(E),
                           C.c_float(1.0), C.c_float(9.8), 3, nsteps, mode, 8, fp(gv), fp(gm), fp(pv), fp(pm))
    return xs, vs, Cs, Js, gv, gm, pv, pm


def cloud(seed, N, lo=0.15, ext=0.4, vel=None):
    rng = np.random.default_rng(seed)
    x = (rng.random((N, 3)) * ext + lo).astype(np.float32)
    v = np.zeros((N, 3), np.float32) if vel is None else np.tile(np.array(vel, np.float32), (N, 1))
    v += (rng.random((N, 3)).astype(np.float32) - 0.5) * 0.1
    Cm = ((rng.random((N, 3, 3)) - 0.5) * 2).astype(np.float32)
    J = (1 + (rng.random(N) - 0.5) * 0.05).astype(np.float32)
    return x, v, Cm, J


def expected_blocks(x, n_grid, halo):
    """x64 semantics: blocks that contain a touched node base..base+2;
    halo: the CUDA-BLS epilogue activates all 8 blocks {0,+1}^3 of every occupied block."""
    base = (x.astype(np.float32) * np.float32(n_grid) - np.float32(0.5)).astype(np.int32)
    out = set()
    if halo:
        for b in np.unique(base >> 2, axis=0):
            for d in np.ndindex(2, 2, 2):
                out.add(tuple(b + np.array(d)))
    else:
        for d in np.ndindex(2, 2, 2):
            nb = np.unique((base + 2 * np.array(d, np.int32)) >> 2, axis=0)
            out.update(map(tuple, nb))
    nbk = n_grid // 4
    arr = np.array(sorted(t for t in out if max(t) < nbk), np.int32).reshape(-1, 3)
    return arr[np.lexsort(arr.T[::-1])]


# --------------------------------------------------------------------------- plain kernels / C-API
def test_capi_mpm88_regression_vector(rt):
    """c_api/tests/mpm88_test.cpp flow with test_mpm88.py's deterministic init and regression
    moments (test_mpm88.py:81-101, rel 1e-2)."""
    N, n_grid = 64 * 64, 128
    i = np.arange(N)
    x0 = np.stack([(i % 64) / 64 * 0.4 + 0.2, i / 64 / 64 * 0.4 + 0.05], axis=1).astype(np.float32)
    mod = rt.load_aot_module("builtin:mpm88")
    x, v = rt.allocate_ndarray((N,), (2,)), rt.allocate_ndarray((N,), (2,))
    Cm, J = rt.allocate_ndarray((N,), (2, 2)), rt.allocate_ndarray((N,))
    pos = rt.allocate_ndarray((N,), (3,))
    gv, gm = rt.allocate_ndarray((n_grid, n_grid), (2,)), rt.allocate_ndarray((n_grid, n_grid))
    x.write(x0)
    v.write(np.tile(np.array([0, -3], np.float32), (N, 1)))
    J.write(np.ones(N, np.float32))
    k_reset, k_p2g = mod.get_kernel("substep_reset_grid"), mod.get_kernel("substep_p2g")
    k_grid, k_g2p = mod.get_kernel("substep_update_grid_v"), mod.get_kernel("substep_g2p")
    for _ in range(500):
        k_reset.launch(gv, gm)
        k_p2g.launch(x, v, Cm, J, gv, gm)
        k_grid.launch(gv, gm)
        k_g2p.launch(x, v, Cm, J, gv, pos)
    rt.wait()
    p = x.read().astype(np.float64)
    p[:, 1] *= 2
    regression = [0.31722742, 0.15826741, 0.10224003, 0.07810827]
    for k in range(4):
        assert abs((p ** (k + 1)).mean() - regression[k]) <= 1e-2 * regression[k]
    assert np.allclose(pos.read()[:, :2], x.read())
    # unknown kernel / module names use the reference's error codes
    from taichi_b200 import _lib as L
    assert L.lib.ti_get_aot_module_kernel(mod.handle, b"nope") is None and L.last_error()[0] == -3
    L.lib.ti_set_last_error(0, None)
    assert L.lib.ti_load_aot_module(rt.handle, b"/tmp/not_a_module") is None and L.last_error()[0] == -2
    L.lib.ti_set_last_error(0, None)


def test_capi_mpm3d_plain_vs_oracle(rt):
    n_grid, N, dt, steps = 32, 4096, 4e-4, 10
    x0, v0, C0, J0 = cloud(2, N)
    mod = rt.load_aot_module(f"builtin:mpm3d?dt={dt}")
    x, v = rt.allocate_ndarray((N,), (3,)), rt.allocate_ndarray((N,), (3,))
    Cm, J = rt.allocate_ndarray((N,), (3, 3)), rt.allocate_ndarray((N,))
    gv, gm = rt.allocate_ndarray((n_grid,) * 3, (3,)), rt.allocate_ndarray((n_grid,) * 3)
    for nd, a in ((x, x0), (v, v0), (Cm, C0), (J, J0)):
        nd.write(a)
    ks = [mod.get_kernel(k) for k in ("substep_reset_grid", "substep_p2g", "substep_update_grid_v", "substep_g2p")]
    for _ in range(steps):
        ks[0].launch(gv, gm)
        ks[1].launch(x, v, Cm, J, gv, gm)
        ks[2].launch(gv, gm)
        ks[3].launch(x, v, Cm, J, gv)
    rt.wait()
    xs, vs, Cs, Js, ogv, ogm, _, _ = oracle_mpm3d(n_grid, x0, v0, C0, J0, dt, steps)
    assert np.allclose(x.read(), xs, rtol=0, atol=2e-6)
    assert np.allclose(v.read(), vs, rtol=1e-4, atol=1e-4)
    assert np.allclose(J.read(), Js, rtol=1e-5)
    assert np.allclose(gm.read(), ogm, rtol=1e-5, atol=1e-10)


def test_capi_mpm128_vs_oracle(rt):
    """config[0]: mpm128 2-D, 128^2, 8192 particles (SURVEY 8d C1 recipe). The SVD/snow branch has
    no regression vector in the reference ("parity unpinned" beyond the C restatement)."""
    N, n_grid, dt, steps = 8192, 128, 1e-4, 50
    rng = np.random.default_rng(1)
    g = np.arange(N) // (N // 3 + 1)
    x0 = (rng.random((N, 2)) * 0.2 + np.stack([0.3 + 0.1 * g, 0.05 + 0.32 * g], 1)).astype(np.float32)
    mat0 = g.astype(np.int32)
    F0 = np.tile(np.eye(2, dtype=np.float32), (N, 1, 1))
    mod = rt.load_aot_module("builtin:mpm128")
    x, v = rt.allocate_ndarray((N,), (2,)), rt.allocate_ndarray((N,), (2,))
    Cm, F = rt.allocate_ndarray((N,), (2, 2)), rt.allocate_ndarray((N,), (2, 2))
    mat, Jp = rt.allocate_ndarray((N,), (), np.int32), rt.allocate_ndarray((N,))
    gv, gm = rt.allocate_ndarray((n_grid, n_grid), (2,)), rt.allocate_ndarray((n_grid, n_grid))
    x.write(x0), F.write(F0), mat.write(mat0), Jp.write(np.ones(N, np.float32))
    ks = [mod.get_kernel(k) for k in ("substep_reset_grid", "substep_p2g", "substep_update_grid_v", "substep_g2p")]
    for _ in range(steps):
        ks[0].launch(gv, gm)
        ks[1].launch(x, v, Cm, F, mat, Jp, gv, gm)
        ks[2].launch(gv, gm)
        ks[3].launch(x, v, Cm, gv)
    rt.wait()
    xs, vs = x0.copy(), np.zeros((N, 2), np.float32)
    Cs, Fs, ms, Jps = np.zeros((N, 2, 2), np.float32), F0.copy(), mat0.copy(), np.ones(N, np.float32)
    ogv, ogm = np.zeros((n_grid, n_grid, 2), np.float32), np.zeros((n_grid, n_grid), np.float32)
    ORC.orc_mpm128_substeps(n_grid, C.c_int64(N), fp(xs), fp(vs), fp(Cs), fp(Fs), ms.ctypes.data_as(C.POINTER(C.c_int32)),
                            fp(Jps), C.c_float(dt), C.c_float(5e3), C.c_float(0.2), C.c_float(0), C.c_float(-1),
                            C.c_float(0), C.c_float(0), C.c_float(0), steps, fp(ogv), fp(ogm))
    assert np.allclose(x.read(), xs, rtol=0, atol=5e-6)
    assert np.allclose(v.read(), vs, rtol=0, atol=2e-3 * max(1.0, np.abs(vs).max()))
    assert np.allclose(F.read(), Fs, rtol=0, atol=1e-3)
    assert np.allclose(Jp.read(), Jps, rtol=1e-3)


# --------------------------------------------------------------------------- block-binned solver
LAYOUT_CASES = [("dense", True), ("pointer_bitmasked", True), ("pointer_pointer", True),
                ("pointer_bitmasked", False), ("dense", False)]


@pytest.mark.parametrize("layout,bls", LAYOUT_CASES)
def test_binned_one_substep_vs_oracle(rt, layout, bls):
    import taichi_b200 as tb
    n_grid, N, dt = 64, 30000, 2e-4
    x0, v0, C0, J0 = cloud(11, N, vel=(0.5, -1.0, 0.25))
    s = tb.Mpm3D(rt, n_grid, N, dt, layout=layout, block_local=bls)
    s.set_particles(x0, v0, C0, J0)
    assert s.stats()["particles"] == N
    # identity round trip through the bins (ids restore the caller's order)
    xr, vr, Cr, Jr = s.get_particles()
    assert np.array_equal(xr, x0) and np.array_equal(vr, v0) and np.array_equal(Cr, C0) and np.array_equal(Jr, J0)
    xs, vs, Cs, Js, ogv, ogm, pv, pm = oracle_mpm3d(n_grid, x0, v0, C0, J0, dt, 1)
    s.clear_grid()
    s.p2g()
    g = s.grid_to_numpy()
    # grid mass / momentum after P2G: per node, relative to the node's own scale plus a floor of
    # 1e-6 of the largest node (weights span 8 orders of magnitude)
    assert np.allclose(g[..., 3], pm, rtol=2e-5, atol=1e-6 * pm.max())
    assert np.allclose(g[..., :3], pv, rtol=2e-5, atol=2e-6 * np.abs(pv).max())
    if layout != "dense":
        assert np.array_equal(s.active_blocks(), expected_blocks(x0, n_grid, halo=False))
    s.grid_op()
    g2 = s.grid_to_numpy()
    act = g2[..., 3] > 0
    assert np.allclose(g2[..., :3][act], ogv[act], rtol=1e-4, atol=1e-5)
    s.g2p()
    xr, vr, Cr, Jr = s.get_particles()
    assert np.allclose(xr, xs, rtol=0, atol=1e-6)
    assert np.allclose(vr, vs, rtol=1e-4, atol=1e-4)
    # C = 4/dx^2 * sum w g (x) dpos: a difference of O(|v|/dx) terms, so its error is the velocity tolerance (1e-4 of the
    # velocity scale) divided by dx / 4: BASELINE.md section 6 "rel 1e-4", stated on the quantity that carries it
    assert np.allclose(Cr, Cs, rtol=1e-4, atol=1e-4 * 4.0 * n_grid * max(1.0, float(np.abs(vs).max())))
    assert np.allclose(Jr, Js, rtol=1e-5)
    assert s.stats()["particles"] == N
    s.destroy()


@pytest.mark.parametrize("layout", ["dense", "pointer_bitmasked"])
def test_binned_ten_substeps_vs_oracle(rt, layout):
    import taichi_b200 as tb
    n_grid, N, dt, steps = 64, 20000, 2e-4, 10
    x0, v0, C0, J0 = cloud(5, N, vel=(0.0, -2.0, 0.0))
    C0[:] = 0
    J0[:] = 1
    s = tb.Mpm3D(rt, n_grid, N, dt, layout=layout)
    s.set_particles(x0, v0, C0, J0)
    s.substep(steps)
    xr, vr, Cr, Jr = s.get_particles()
    xs, vs, Cs, Js, *_ = oracle_mpm3d(n_grid, x0, v0, C0, J0, dt, steps)
    assert np.allclose(xr, xs, rtol=0, atol=2e-6)          # particle state rel 1e-4 after 10 substeps
    assert np.allclose(vr, vs, rtol=1e-4, atol=2e-4)
    assert np.allclose(Jr, Js, rtol=1e-4)
    assert s.stats()["particles"] == N
    s.destroy()


@pytest.mark.parametrize("case", ["fat_cells", "multi_pass", "one_per_block", "mixed"])
def test_p2g_ragged_bins_vs_oracle(rt, case):
    """Edge cases of the unit-balanced block-local P2G (b2_mpm_p2g.cuh): bins of more than one 576-entry
    pass, thousands of particles in one cell (unit length L > 31), blocks holding a single particle,
    and a mix; grid after P2G per node against the oracle's fp64-accumulated P2G, active sets bit-exact."""
    import taichi_b200 as tb
    n_grid, dt = 32, 2e-4
    rng = np.random.default_rng({"fat_cells": 21, "multi_pass": 22, "one_per_block": 23, "mixed": 24}[case])
    dx = 1.0 / n_grid
    if case == "fat_cells":      # 3000 particles inside ONE cell + 500 in a second cell of the same block
        x = np.concatenate([(np.array([9.6, 9.7, 9.8]) + rng.random((3000, 3)) * 0.8) * dx,
                            (np.array([10.6, 8.6, 11.6]) + rng.random((500, 3)) * 0.8) * dx])
    elif case == "multi_pass":   # 40 particles per cell over two adjacent blocks: 2560 entries per bin
        x = (np.array([8.5, 8.5, 8.5]) + rng.random((5120, 3)) * np.array([8.0, 4.0, 4.0])) * dx
    elif case == "one_per_block":
        bl = np.stack(np.meshgrid(np.arange(1, 7), np.arange(1, 7), np.arange(1, 7), indexing="ij"), -1).reshape(-1, 3)
        x = (bl * 4 + 0.5 + rng.random((len(bl), 3)) * 3.9) * dx
    else:
        x = np.concatenate([(np.array([4.5, 4.5, 4.5]) + rng.random((7000, 3)) * 20.0) * dx,
                            (np.array([13.2, 13.2, 13.2]) + rng.random((1500, 3)) * 0.5) * dx])
    x = x.astype(np.float32)
    N = len(x)
    v = ((rng.random((N, 3)) - 0.5) * 2).astype(np.float32)
    Cm = ((rng.random((N, 3, 3)) - 0.5) * 4).astype(np.float32)
    J = (1 + (rng.random(N) - 0.5) * 0.1).astype(np.float32)
    for layout in ("pointer_bitmasked", "dense"):
        s = tb.Mpm3D(rt, n_grid, N, dt, layout=layout)
        s.set_particles(x, v, Cm, J)
        *_, pv, pm = oracle_mpm3d(n_grid, x, v, Cm, J, dt, 1)
        s.clear_grid()
        s.p2g()
        g = s.grid_to_numpy()
        assert np.allclose(g[..., 3], pm, rtol=2e-5, atol=1e-6 * pm.max())
        assert np.allclose(g[..., :3], pv, rtol=2e-5, atol=2e-6 * np.abs(pv).max())
        if layout != "dense":
            assert np.array_equal(s.active_blocks(), expected_blocks(x, n_grid, halo=False))
        # the whole substep keeps every particle (re-binning of the ragged bins) and stays finite
        s.grid_op()
        s.g2p()
        s.substep(2)
        xr, vr, Cr, Jr = s.get_particles()
        assert s.stats()["particles"] == N and np.isfinite(xr).all() and np.isfinite(vr).all()
        s.destroy()


def test_bls_activate_halo_mode(rt):
    """the reference's CUDA-BLS epilogue activates all 8 blocks {0,+1}^3."""
    import taichi_b200 as tb
    n_grid, N = 64, 5000
    x0, v0, C0, J0 = cloud(3, N, lo=0.3, ext=0.2)
    out = {}
    for halo in (False, True):
        s = tb.Mpm3D(rt, n_grid, N, 2e-4, layout="pointer_bitmasked", bls_activate_halo=halo)
        s.set_particles(x0, v0, C0, J0)
        s.clear_grid()
        s.p2g()
        assert np.array_equal(s.active_blocks(), expected_blocks(x0, n_grid, halo=halo))
        out[halo] = s.grid_to_numpy()
        s.destroy()
    # values identical up to the order of the fp32 reductions (only zeros are added by the halo mode)
    assert np.allclose(out[False], out[True], rtol=1e-5, atol=1e-6 * np.abs(out[False]).max())


def test_bls_equals_plain(rt):
    """bls_test_template.py:85-96,258-272: block_local on/off agree (f32: abs 1e-4 on O(1) sums)."""
    import taichi_b200 as tb
    n_grid, N = 64, 40000
    x0, v0, C0, J0 = cloud(9, N, vel=(1.0, 0.0, -1.0))
    gs = []
    for bls in (True, False):
        s = tb.Mpm3D(rt, n_grid, N, 2e-4, layout="pointer_bitmasked", block_local=bls)
        s.set_particles(x0, v0, C0, J0)
        s.clear_grid()
        s.p2g()
        gs.append(s.grid_to_numpy())
        blocks = s.active_blocks()
        s.destroy()
    scale = np.abs(gs[1]).max()
    assert np.allclose(gs[0], gs[1], rtol=1e-4, atol=1e-4 * scale)


def test_conservation_full_size(rt):
    """config[1] size (128^3, 1M particles): size-independent properties -- partition of unity
    (total grid mass = N p_mass), momentum conservation P2G, particle count, and that particles
    stay binned in the block of their base cell."""
    import taichi_b200 as tb
    n_grid, N, dt = 128, 1 << 20, 8e-5
    x0, v0, C0, J0 = cloud(2, N, vel=(0.3, -1.0, 0.2))
    C0[:] = 0
    s = tb.Mpm3D(rt, n_grid, N, dt, layout="pointer_bitmasked")
    s.set_particles(x0, v0, C0, J0)
    s.clear_grid()
    s.p2g()
    g = s.grid_to_numpy().astype(np.float64)
    p_mass = (1.0 / n_grid * 0.5) ** 2
    assert abs(g[..., 3].sum() - N * p_mass) < 1e-6 * N * p_mass
    # total momentum = sum m v + stress terms that cancel (sum_i w_i dpos_i = 0 for quadratic B-splines with C = 0)
    mom = g[..., :3].reshape(-1, 3).sum(0)
    expect = p_mass * v0.astype(np.float64).sum(0)
    assert np.allclose(mom, expect, rtol=1e-4, atol=1e-6 * N * p_mass)
    s.grid_op()
    s.g2p()
    s.substep(20)
    st = s.stats()
    assert st["particles"] == N
    xr, vr, Cr, Jr = s.get_particles()
    assert np.isfinite(xr).all() and np.isfinite(vr).all() and (Jr > 0).all()
    assert xr.min() > 0.0 and xr.max() < 1.0
    # grid mass is conserved after 21 substeps too
    s.clear_grid()
    s.p2g()
    g = s.grid_to_numpy().astype(np.float64)
    assert abs(g[..., 3].sum() - N * p_mass) < 1e-6 * N * p_mass
    assert np.array_equal(s.active_blocks(), expected_blocks(xr, n_grid, halo=False)) or \
        set(map(tuple, expected_blocks(xr, n_grid, halo=False))) <= set(map(tuple, s.active_blocks()))
    s.destroy()


def test_pipelined_host_io_matches_blocking(rt):
    """Host arrays go through the solver's copy streams (double-buffered staging): three frames issued
    back to back with page-locked inputs and asynchronous downloads must return exactly what the same
    frames return when every call blocks (deterministic kernels: bit-exact)."""
    import torch
    import taichi_b200 as tb
    n_grid, N, dt, steps = 64, 50000, 2e-4, 5
    frames = []
    for k in range(3):
        x0, v0, C0, J0 = cloud(20 + k, N, vel=(0.0, -1.0 - k, 0.0))
        frames.append(tuple(torch.from_numpy(a).pin_memory() for a in (x0, v0, C0, J0)))
    s = tb.Mpm3D(rt, n_grid, N, dt, layout="pointer_bitmasked")
    blocking = []
    for fr in frames:
        s.set_particles(*fr)
        s.substep(steps)
        blocking.append([a.copy() for a in s.get_particles()])
    outs = [(torch.empty(N, 3).pin_memory(), torch.empty(N, 3).pin_memory(), torch.empty(N, 3, 3).pin_memory(), torch.empty(N).pin_memory())
            for _ in frames]
    for fr, o in zip(frames, outs):
        s.set_particles(*fr)
        s.substep(steps)
        s.get_particles(*o, wait=False)
    s.wait_io()
    for ref, o in zip(blocking, outs):
        xr, vr, Cr, Jr = ref
        # positions and J are order-independent sums of few terms; v and C come from fp32 atomics on the grid,
        # whose order may differ between runs: same tolerance as the BLS/plain comparison (1e-4)
        assert np.allclose(o[0].numpy(), xr, rtol=0, atol=1e-6)
        assert np.allclose(o[1].numpy(), vr, rtol=1e-4, atol=1e-4)
        assert np.allclose(o[3].numpy(), Jr, rtol=1e-5)
    # pageable numpy arrays take the same path (the driver stages them before the call returns)
    x0, v0, C0, J0 = (a.numpy().copy() for a in frames[0])
    s.set_particles(x0, v0, C0, J0)
    x0[:] = 0.5  # overwritten right after the call: the upload must already own its copy
    s.substep(steps)
    xr = s.get_particles()[0]
    assert np.allclose(xr, blocking[0][0], rtol=0, atol=1e-6)
    s.destroy()


def test_full_size_c3_properties(rt):
    """config[2] at full size (512^3 pointer -> bitmasked -> dense(4^3), 8M particles): the size-independent
    properties -- partition of unity (grid mass = N p_mass), momentum conservation of P2G, exact active-block set,
    particle count and finiteness after 10 substeps.  The grid sums run as a run-time compiled struct-for over the
    solver's own SNode tree (b2_jit_*), in fp64."""
    import sys
    import os
    import torch
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    import taichi_b200 as tb
    from taichi_b200 import _lib as L
    lib = L.lib
    n_grid, N, dt, layout, _ = bench.WORKLOADS["c3"]
    x0, v0, C0, J0 = bench.make_cloud("c3", N, seed=3)
    s = tb.Mpm3D(rt, n_grid, N, dt, layout=layout)
    s.set_particles(x0, v0, C0, J0)
    s.clear_grid()
    s.p2g()
    k = lib.b2_jit_compile(rt.handle, b"""B2_BODY(c) {
        double *acc = c.arg<double *>(0);
        const float m = c.read<float>(3, c.I);
        if (m != 0.0f) {
            atomicAdd(acc + 0, (double)m);
            atomicAdd(acc + 1, (double)c.read<float>(0, c.I));
            atomicAdd(acc + 2, (double)c.read<float>(1, c.I));
            atomicAdd(acc + 3, (double)c.read<float>(2, c.I));
            atomicAdd(acc + 4, 1.0);
        }
    }""", b"grid_sums")
    L.check()
    tree, leaf = lib.b2_mpm_tree(s.handle), lib.b2_mpm_leaf_snode(s.handle)
    places = (C.c_int32 * 4)(leaf + 1, leaf + 2, leaf + 3, leaf + 4)

    def grid_sums():
        acc = torch.zeros(5, dtype=torch.float64, device="cuda")
        args = (C.c_uint64 * 1)(acc.data_ptr())
        assert lib.b2_jit_launch(k, tree, leaf, places, 4, args, 1) == 0, L.last_error()
        torch.cuda.synchronize()
        return acc.cpu().numpy()

    def touched_blocks(x):  # expected_blocks(halo=False) on integer keys (8M particles: seconds instead of a minute)
        base = (x.astype(np.float32) * np.float32(n_grid) - np.float32(0.5)).astype(np.int32)
        nb = n_grid // 4
        keys = []
        for d in np.ndindex(2, 2, 2):
            b = (base + 2 * np.array(d, np.int32)) >> 2
            ok = (b < nb).all(1)
            keys.append(np.unique((b[ok, 0].astype(np.int64) * nb + b[ok, 1]) * nb + b[ok, 2]))
        return np.unique(np.concatenate(keys))

    def keys_of(blocks):
        nb = n_grid // 4
        return np.sort((blocks[:, 0].astype(np.int64) * nb + blocks[:, 1]) * nb + blocks[:, 2])

    p_mass = float(np.float32((1.0 / n_grid * 0.5) ** 2))
    g = grid_sums()
    assert abs(g[0] - N * p_mass) < 2e-6 * N * p_mass
    want_p = p_mass * v0.astype(np.float64).sum(0)            # C = 0, J = 1: the stress / affine terms cancel
    assert np.allclose(g[1:4], want_p, rtol=1e-5, atol=1e-7 * N * p_mass)
    blocks = s.active_blocks()
    assert np.array_equal(keys_of(blocks), touched_blocks(x0))
    assert g[4] <= len(blocks) * 64
    d0 = s.diagnostics()
    assert d0["particles"] == N and abs(d0["py"] - want_p[1]) < 1e-6 * abs(want_p[1]) and abs(d0["sum_J"] - N) < 1e-3
    s.grid_op()
    s.g2p()
    s.substep(9)
    st = s.stats()
    assert st["particles"] == N
    d1 = s.diagnostics()
    assert d1["particles"] == N and np.isfinite(list(d1.values())).all() and d1["max_speed"] < 10.0 and d1["max_dJ"] < 0.5
    s.clear_grid()
    s.p2g()
    g = grid_sums()
    assert abs(g[0] - N * p_mass) < 2e-6 * N * p_mass          # mass is still all on the grid after 10 substeps
    xr = s.get_particles()[0]
    assert np.isfinite(xr).all() and xr.min() > 0.0 and xr.max() < 1.0
    assert np.isin(touched_blocks(xr), keys_of(s.active_blocks())).all()
    lib.b2_jit_destroy(k)
    s.destroy()


def test_out_of_domain_particles_raise_the_device_error(rt):
    """The reference asserts in debug mode when an index leaves the field (`taichi_assert_format`, runtime.cpp:727-741);
    here a particle whose stencil leaves the grid, or a NaN position, sets the device-side error flag (and is clamped
    into the grid, never written out of bounds).  Reading the flag clears it."""
    import taichi_b200 as tb
    lib = tb._lib.lib
    buf = C.create_string_buffer(256)
    assert lib.b2_check_runtime_error(rt.handle, buf, 256) == 0, buf.value
    n_grid, N = 64, 1000
    x0, v0, C0, J0 = cloud(31, N)
    for bad in (np.array([1.5, 0.5, 0.5], np.float32), np.array([0.5, -0.2, 0.5], np.float32),
                np.array([0.5, 0.5, np.nan], np.float32)):
        s = tb.Mpm3D(rt, n_grid, N, 2e-4, layout="pointer_bitmasked")
        x = x0.copy()
        x[17] = bad
        s.set_particles(x, v0, C0, J0)
        with pytest.raises(tb._lib.B2RuntimeError, match=r"assertion failed \[arg 17\]"):
            rt.wait()                                                    # ti_wait surfaces it (llvm_runtime_executor.cpp:272-312)
        assert lib.b2_check_runtime_error(rt.handle, buf, 256) == 0      # cleared by the read
        assert s.stats()["particles"] == N                               # clamped into a border block, not dropped
        s.destroy()
    s = tb.Mpm3D(rt, n_grid, N, 2e-4, layout="pointer_bitmasked")
    s.set_particles(x0, v0, C0, J0)
    s.substep(2)
    rt.wait()
    assert lib.b2_check_runtime_error(rt.handle, buf, 256) == 0, buf.value  # a valid cloud raises nothing
    s.destroy()


# --------------------------------------------------------------------------- full-size oracle parity (VERDICT r1 #1a, #1b)
def oracle_sparse(n_grid, x, v, Cm, J, dt, nsteps, deactivate_every=0, cap=1 << 16, threads=None):
    """oracle/mpm_oracle.c orc_mpm3d_sparse_substeps, mode 0 (fp64 accumulators): particles after nsteps, and the
    active blocks (sorted keys) with the raw P2G cells of the LAST substep."""
    import os
    xs, vs, Cs, Js = x.copy(), v.copy(), Cm.copy(), J.copy()
    nblk = C.c_int32(0)
    keys = np.zeros(cap, np.int32)
    p2g = np.zeros((cap, 64, 4), np.float32)
    rc = ORC.orc_mpm3d_sparse_substeps(n_grid, C.c_int64(len(x)), fp(xs), fp(vs), fp(Cs), fp(Js), C.c_float(dt), C.c_float(400.0),
                                       C.c_float(1.0), C.c_float(9.8), 3, nsteps, 0, threads or min(os.cpu_count() or 1, 64),
                                       deactivate_every, cap, C.byref(nblk), keys.ctypes.data_as(C.POINTER(C.c_int32)), fp(p2g), None)
    assert rc == 0
    n = nblk.value
    order = np.argsort(keys[:n])
    return xs, vs, Cs, Js, keys[:n][order], p2g[:n][order]


def check_p2g_blocks(s, keys_o, p2g_o):
    """grid after P2G, per node over the whole active set: mass rtol 2e-5 (+1e-6 of the largest node), momentum
    rtol 2e-5 (+2e-6 of the largest component) -- the tolerances of the small cases above; block set bit-exact."""
    keys, cells = s.grid_blocks()
    assert np.array_equal(keys, keys_o)
    assert np.allclose(cells[..., 3], p2g_o[..., 3], rtol=2e-5, atol=1e-6 * p2g_o[..., 3].max())
    assert np.allclose(cells[..., :3], p2g_o[..., :3], rtol=2e-5, atol=2e-6 * np.abs(p2g_o[..., :3]).max())


def check_particles(s, xs, vs, Cs, Js, dx, vmax):
    """particle state rel 1e-4 after 10 substeps.  x: 2e-6 absolute (domain is [0,1)); v: rel 1e-4 of
    the velocity scale; J: rel 1e-4; C = 4/dx^2 * sum w g (x) dpos is a difference of O(|v|/dx) terms: its error is
    bounded by the velocity error divided by dx, i.e. 1e-4 * 4 |v|max / dx absolute."""
    xr, vr, Cr, Jr = s.get_particles()
    assert np.allclose(xr, xs, rtol=0, atol=2e-6)
    assert np.allclose(vr, vs, rtol=1e-4, atol=1e-4 * vmax)
    assert np.allclose(Jr, Js, rtol=1e-4)
    assert np.allclose(Cr, Cs, rtol=1e-4, atol=1e-4 * 4.0 * vmax / dx)


def test_c2_full_size_vs_oracle(rt):
    """config[1]: mpm3d dense 128^3, 1M particles (SURVEY 8d C2 recipe): P2G grid per node after one substep and the
    particle state after 10 substeps against the fp64-accumulating oracle."""
    import taichi_b200 as tb
    n_grid, N, dt = 128, 1 << 20, 8e-5
    rng = np.random.default_rng(2)
    x0 = (rng.random((N, 3)) * 0.4 + 0.15).astype(np.float32)
    v0 = np.zeros((N, 3), np.float32)
    C0 = np.zeros((N, 3, 3), np.float32)
    J0 = np.ones(N, np.float32)
    s = tb.Mpm3D(rt, n_grid, N, dt, layout="dense")
    s.set_particles(x0, v0, C0, J0)
    s.clear_grid()
    s.p2g()
    g = s.grid_to_numpy()
    _, _, _, _, keys_o, p2g_o = oracle_sparse(n_grid, x0, v0, C0, J0, dt, 1)
    nb = n_grid // 4
    bx, by, bz = keys_o // (nb * nb), (keys_o // nb) % nb, keys_o % nb
    for t in range(64):
        cell = g[bx * 4 + (t >> 4), by * 4 + ((t >> 2) & 3), bz * 4 + (t & 3)]
        assert np.allclose(cell[:, 3], p2g_o[:, t, 3], rtol=2e-5, atol=1e-6 * p2g_o[..., 3].max())
        assert np.allclose(cell[:, :3], p2g_o[:, t, :3], rtol=2e-5, atol=2e-6 * np.abs(p2g_o[..., :3]).max())
    assert abs(float(g[..., 3].astype(np.float64).sum()) - float(p2g_o[..., 3].astype(np.float64).sum())) < 1e-6 * N * (0.5 / n_grid) ** 2
    s.grid_op()
    s.g2p()
    s.substep(9)
    xs, vs, Cs, Js, _, _ = oracle_sparse(n_grid, x0, v0, C0, J0, dt, 10)
    check_particles(s, xs, vs, Cs, Js, 1.0 / n_grid, max(1.0, float(np.abs(vs).max())))
    s.destroy()


def test_c3_full_size_vs_oracle(rt):
    """config[2]: sparse 512^3 pointer -> bitmasked -> dense(4^3), 8M particles (bench.py's cloud): P2G grid per node
    over the whole active set and the exact active-block set after one substep, particle state after 10 substeps,
    against the sparse fp64-accumulating oracle (oracle/mpm_oracle.c, the same arithmetic as the dense restatement)."""
    import sys
    import os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    import taichi_b200 as tb
    n_grid, N, dt, layout, _ = bench.WORKLOADS["c3"]
    x0, v0, C0, J0 = bench.make_cloud("c3", N, seed=3)
    s = tb.Mpm3D(rt, n_grid, N, dt, layout=layout)
    s.set_particles(x0, v0, C0, J0)
    s.clear_grid()
    s.p2g()
    _, _, _, _, keys_o, p2g_o = oracle_sparse(n_grid, x0, v0, C0, J0, dt, 1)
    check_p2g_blocks(s, keys_o, p2g_o)
    s.grid_op()
    s.g2p()
    s.substep(9)
    xs, vs, Cs, Js, _, _ = oracle_sparse(n_grid, x0, v0, C0, J0, dt, 10)
    check_particles(s, xs, vs, Cs, Js, 1.0 / n_grid, max(1.0, float(np.abs(vs).max())))
    assert s.stats()["particles"] == N
    s.destroy()


# --------------------------------------------------------------------------- sparse life cycle
def test_deactivate_every_substep_life_cycle(rt):
    """The reference's sparse-MPM idiom (snode.py:286-304 deactivate_all, offload.cpp:671-695 GC after the deactivating
    task, test_gc.py:40-52, test_mpm_particle_list.py:9-53) inside the solver's own loop: with deactivate_every = 1 the
    active-block set after EVERY substep is exactly the set of blocks the particles touched in that substep (bit-exact
    against the oracle run with the same life cycle), the physics is unchanged, and the pointer level's allocator stays
    bounded however long the run is (nodes are recycled through the GC, `_num_dynamically_allocated`)."""
    import taichi_b200 as tb
    lib = tb._lib.lib
    n_grid, N, dt = 64, 20000, 2e-4
    x0, v0, C0, J0 = cloud(41, N, lo=0.25, ext=0.3, vel=(4.0, -6.0, 2.0))     # fast: the cloud crosses many blocks
    s = tb.Mpm3D(rt, n_grid, N, dt, layout="pointer_bitmasked", deactivate_every=1)
    keep = tb.Mpm3D(rt, n_grid, N, dt, layout="pointer_bitmasked")            # never deactivates: the set only grows
    for m in (s, keep):
        m.set_particles(x0, v0, C0, J0)
    xs, vs, Cs, Js = x0, v0, C0, J0
    for step in range(12):
        xb = xs
        xs, vs, Cs, Js, keys_o, _ = oracle_sparse(n_grid, xs, vs, Cs, Js, dt, 1, deactivate_every=1, cap=4096)
        s.substep(1)
        keep.substep(1)
        want = expected_blocks(xb, n_grid, halo=False)
        got = s.active_blocks()
        assert np.array_equal(got, want)
        nb = n_grid // 4
        assert np.array_equal(np.sort((got[:, 0] * nb + got[:, 1]) * nb + got[:, 2]), keys_o)
        assert set(map(tuple, got)) <= set(map(tuple, keep.active_blocks()))
    assert len(keep.active_blocks()) > len(s.active_blocks())                 # the cloud has left blocks behind
    xr, vr, Cr, Jr = s.get_particles()
    xk, vk, Ck, Jk = keep.get_particles()
    assert np.allclose(xr, xs, rtol=0, atol=2e-6) and np.allclose(vr, vs, rtol=1e-4, atol=1e-3) and np.allclose(Jr, Js, rtol=1e-4)
    assert np.allclose(xr, xk, rtol=0, atol=2e-6) and np.allclose(vr, vk, rtol=1e-4, atol=1e-3)
    # allocator: bounded over 100 more frames of 4 substeps (every substep recycles and re-allocates the containers)
    tree = lib.b2_mpm_tree(s.handle)
    seen = []
    for _ in range(100):
        s.substep(4)
        seen.append(int(lib.b2_num_dynamically_allocated(tree, 1)))
    st = np.zeros(4, np.int32)
    lib.b2_alloc_stats(tree, 1, st.ctypes.data_as(tb._lib.i32p))
    assert max(seen) <= (n_grid // 32) ** 3 and st[0] <= 2 * (n_grid // 32) ** 3   # never more nodes than pointer cells (x2: ping-pong)
    assert s.stats()["particles"] == N
    buf = C.create_string_buffer(256)
    assert lib.b2_check_runtime_error(rt.handle, buf, 256) == 0, buf.value
    s.destroy()
    keep.destroy()


def test_deactivate_every_k(rt):
    """deactivate_every = K > 1: the set is rebuilt on substeps 0, K, 2K, ... and grows in between."""
    import taichi_b200 as tb
    n_grid, N, dt, K = 64, 20000, 2e-4, 5
    x0, v0, C0, J0 = cloud(42, N, lo=0.25, ext=0.3, vel=(4.0, -6.0, 2.0))
    s = tb.Mpm3D(rt, n_grid, N, dt, layout="pointer_pointer", deactivate_every=K)
    s.set_particles(x0, v0, C0, J0)
    touched = set()
    for step in range(2 * K + 1):
        xb = s.get_particles()[0]
        if step % K == 0:
            touched = set()
        touched |= set(map(tuple, expected_blocks(xb, n_grid, halo=False)))
        s.substep(1)
        assert set(map(tuple, s.active_blocks())) == touched
    s.destroy()


def test_bin_slots_are_recycled(rt):
    """a block gets a bin slot the first time a particle lands in it; without recycling a cloud that keeps moving
    exhausts the table.  Here the table holds 24 bins, the cloud occupies <= 8 blocks at a time and crosses ~40 distinct
    blocks: the periodic compaction returns vacated slots, nothing is dropped, no error is raised.  With compaction off
    the same run overflows: the particles beyond the table are DROPPED and flagged (never aliased onto a foreign bin)."""
    import taichi_b200 as tb
    lib = tb._lib.lib
    n_grid, N, dt = 64, 4000, 2e-4
    rng = np.random.default_rng(77)
    x0 = ((np.array([6.0, 31.0, 31.0]) + rng.random((N, 3)) * 2.0) / n_grid).astype(np.float32)   # straddles 2 x 2 blocks in y, z
    v0 = np.tile(np.array([5.0, 0.0, 0.0], np.float32), (N, 1))
    C0, J0 = np.zeros((N, 3, 3), np.float32), np.ones(N, np.float32)
    buf = C.create_string_buffer(256)
    for compact in (16, 0):
        s = tb.Mpm3D(rt, n_grid, N, dt, layout="pointer_bitmasked", gravity=0.0, max_bins=24, deactivate_every=1)
        assert lib.b2_mpm_set_compact_every(s.handle, compact) == 0
        s.set_particles(x0, v0, C0, J0)
        s.substep(700)                       # 5 * 2e-4 * 64 = 0.064 cells per substep: 45 cells = 11 block layers along x
        st = s.stats()
        err = lib.b2_check_runtime_error(rt.handle, buf, 256)
        xr = s.get_particles()[0] if compact else None
        if compact:
            assert err == 0, buf.value
            assert st["particles"] == N and st["slots"] <= 24
            assert xr[:, 0].min() > (6.0 + 40.0) / n_grid and np.isfinite(xr).all()
        else:
            assert err != 0 and b"max_bins" in buf.value
            assert st["particles"] < N         # dropped, and flagged
        s.destroy()


def test_materials_model_in_the_binned_kernels(rt):
    """SURVEY 8 row f3 / VERDICT r1 #9: the mpm128 constitutive model (liquid / jelly / snow, deformation gradient, ti.svd)
    carried to 3-D (ggui_examples/mpm3d_ggui.py:52-88) INSIDE the block-binned solver: F, Jp and the material id travel with
    the particle through the bins, the material update runs before the block-local P2G.  Same cloud, parameters and float64
    NumPy restatement as the user-written version of test_gpu_jit.py; the sparse life cycle runs every substep."""
    import taichi_b200 as tb
    from test_gpu_jit import _np_mpm3d_material_substep
    n_grid, N, steps = 64, 20000, 6
    E, nu = 1e3, 0.2
    dx = 1.0 / n_grid
    par = dict(dt=1e-4, n_grid=n_grid, p_vol=(dx * 0.5) ** 3, p_mass=(dx * 0.5) ** 3 * 1.0, gravity=9.8, bound=3,
               mu_0=E / (2 * (1 + nu)), lambda_0=E * nu / ((1 + nu) * (1 - 2 * nu)))
    rng = np.random.default_rng(17)
    x0 = (rng.random((N, 3)) * 0.3 + 0.3).astype(np.float32)
    v0 = (rng.standard_normal((N, 3)) * 0.3).astype(np.float32)
    C0 = (rng.standard_normal((N, 3, 3)) * 2.0).astype(np.float32)
    q = np.linalg.qr(rng.standard_normal((N, 3, 3)))[0]
    q *= np.sign(np.linalg.det(q))[:, None, None]
    F0 = (q @ (np.eye(3) + 0.02 * rng.standard_normal((N, 3, 3)))).astype(np.float32)
    Jp0 = (1 + 0.02 * rng.standard_normal(N)).astype(np.float32)
    mat = rng.integers(0, 3, N).astype(np.int32)
    s = tb.Mpm3D(rt, n_grid, N, par["dt"], layout="pointer_bitmasked", E=E, nu=nu, p_rho=1.0, gravity=par["gravity"], bound=3,
                 p_vol=par["p_vol"], model="materials", deactivate_every=1)
    s.set_particles_model(x0, v0, C0, F0, Jp0, mat)
    S = dict(x=x0.astype(np.float64), v=v0.astype(np.float64), C=C0.astype(np.float64), F=F0.astype(np.float64),
             Jp=Jp0.astype(np.float64), mat=mat)
    for k in range(steps):
        x_before = S["x"].astype(np.float32)
        s.substep(1)
        want_m = _np_mpm3d_material_substep(S, par)
        if k in (0, steps - 1):  # grid mass of the substep (after G2P the grid still holds it) and the touched blocks
            got_m = s.grid_to_numpy()[..., 3]
            assert np.allclose(got_m, want_m, rtol=2e-4, atol=par["p_mass"] * 1e-4), k
            assert np.array_equal(s.active_blocks(), expected_blocks(x_before, n_grid, halo=False))
    x, v, Cm, F, Jp, m2 = s.get_particles_model()
    assert s.stats()["particles"] == N and np.array_equal(m2, mat)
    assert np.abs(x - S["x"]).max() < 2e-6
    assert np.abs(v - S["v"]).max() < 2e-3 * max(1.0, np.abs(S["v"]).max())
    assert np.abs(Cm - S["C"]).max() < 2e-3 * np.abs(S["C"]).max()
    assert np.abs(F - S["F"]).max() < 2e-5
    assert np.allclose(Jp, S["Jp"], rtol=1e-4)
    assert np.abs(S["v"] - v0).max() > 0.05                                  # the substeps did something
    assert (np.abs(Jp - Jp0)[mat == 2] > 1e-4).any()                         # the snow clamp was exercised
    # the fluid model refuses the model calls, and the other way round the plain getters still work
    x2, v2, C2, J2 = s.get_particles()
    assert np.array_equal(x2, x) and np.array_equal(C2, Cm)
    s.destroy()
    f = tb.Mpm3D(rt, n_grid, 1000, 1e-4)
    with pytest.raises(Exception):
        f.set_particles_model(x0[:1000])
    f.destroy()
