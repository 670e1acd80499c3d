"""Struct-for with run-time compiled bodies: the reference's sparse kernels written as
B2_BODY sources and run through b2_jit_compile / b2_jit_launch (tests/python/test_bitmasked.py:5-31,
test_sparse_basics.py pointer/activation idioms, test_struct_for_non_pot.py, docs/lang/articles/basic/sparse.md:
writes activate, reads do not, struct-for visits active cells only)."""
import numpy as np
import pytest

from _b2 import have_gpu

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ti():
    if not have_gpu():
        pytest.skip("no CUDA device")
    import taichi_b200 as ti
    ti.init(arch=ti.cuda, device_memory_GB=0.5)
    yield ti
    ti.reset()


def test_sum_over_bitmasked(ti):  # test_bitmasked.py:5-31: for i, j in x: s[None] += x[i, j]
    import torch
    x = ti.field(ti.i32)
    bm = ti.root.bitmasked(ti.ij, (3, 6)).bitmasked(ti.i, 8)
    bm.place(x)
    x[5, 1] = 2
    x[9, 4] = 20
    x[0, 3] = 20
    k = ti.struct_for("""B2_BODY(c) {
        atomicAdd(c.arg<int *>(0), c.read<int>(0, c.I));
        atomicAdd(c.arg<int *>(0) + 1, 1);
    }""", "sum_x")
    s = torch.zeros(2, dtype=torch.int32, device="cuda")
    k(bm, fields=[x], args=[s])
    assert s.tolist() == [42, 3]
    k(bm, fields=[x], args=[s])          # a second launch re-generates the list and accumulates
    assert s.tolist() == [84, 6]


def test_write_activates_read_does_not(ti):  # sparse.md: y[i + 1] = 2 x[i] activates; reading x[i - 1] does not
    import torch
    x, y = ti.field(ti.f32), ti.field(ti.f32)
    bx = ti.root.pointer(ti.i, 32).dense(ti.i, 8)
    by = ti.root.pointer(ti.i, 32).dense(ti.i, 8)
    bx.place(x)
    by.place(y)
    src = [3, 7, 8, 100, 255 - 1]
    for q in src:
        x[q] = float(q + 1)
    k = ti.struct_for("""B2_BODY(c) {
        int J[B2_MAX_AXES], K[B2_MAX_AXES];
        c.offset(J, 0, +1);
        c.offset(K, 0, -1);
        const float v = c.read<float>(0, c.I);
        const float left = c.I[0] > 0 ? c.read<float>(0, K) : 0.0f;   // inactive neighbours read as 0
        if (v != 0.0f) c.write<float>(1, J, c.arg<float>(0) * v + left);
    }""", "shift")
    before = ti.active_indices(bx.parent()).copy()
    k(bx, fields=[x, y], args=[2.0])
    img_x, img_y = x.to_numpy(), y.to_numpy()
    want = np.zeros(256, np.float32)
    for q in src:
        want[q + 1] = 2.0 * (q + 1) + img_x[q - 1]
    assert np.array_equal(img_y, want)
    assert np.array_equal(ti.active_indices(bx.parent()), before)     # reads did not activate x's blocks
    got_blocks = sorted(int(b) for b in ti.active_indices(by.parent())[:, 0])
    assert got_blocks == sorted({(q + 1) // 8 for q in src})          # writes activated exactly the touched blocks


def test_loop_over_pointer_level_and_non_pot(ti):  # struct-for over a non-leaf SNode; non power-of-two shapes
    import torch
    x = ti.field(ti.i32)
    ptr = ti.root.pointer(ti.ij, (3, 5))
    blk = ptr.dense(ti.ij, (3, 2))
    blk.place(x)
    for (a, b) in [(0, 0), (4, 7), (8, 9), (7, 1)]:
        x[a, b] = 1
    k = ti.struct_for("""B2_BODY(c) {
        const unsigned long long n = atomicAdd(c.arg<unsigned long long *>(0), 1ull);
        int *out = c.arg<int *>(1);
        out[2 * n] = c.I[0], out[2 * n + 1] = c.I[1];
    }""", "dump")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    out = torch.full((64, 2), -1, dtype=torch.int32, device="cuda")
    k(ptr, args=[cnt, out])
    n = int(cnt.item())
    cells = sorted(map(tuple, out[:n].tolist()))
    assert cells == sorted({(a // 3, b // 2) for (a, b) in [(0, 0), (4, 7), (8, 9), (7, 1)]})
    cnt.zero_()
    k(blk, args=[cnt, out])                                            # leaf level: every cell of the active blocks
    n = int(cnt.item())
    assert n == 4 * 6
    got = sorted(map(tuple, out[:n].tolist()))
    assert got == sorted(map(tuple, ti.active_indices(blk).tolist()))


def test_sparse_laplacian_3d(ti):  # pointer -> bitmasked -> dense(4^3): 7-point stencil into a second sparse field
    x, y = ti.field(ti.f32), ti.field(ti.f32)
    bx = ti.root.pointer(ti.ijk, 4).bitmasked(ti.ijk, 4).dense(ti.ijk, 4)
    by = ti.root.pointer(ti.ijk, 4).bitmasked(ti.ijk, 4).dense(ti.ijk, 4)
    bx.place(x)
    by.place(y)
    rng = np.random.default_rng(0)
    pts = rng.integers(8, 40, size=(300, 3))
    vals = rng.random(300).astype(np.float32)
    x[pts] = vals
    k = ti.struct_for("""B2_BODY(c) {
        float acc = -6.0f * c.read<float>(0, c.I);
        for (int a = 0; a < 3; a++) {
            int J[B2_MAX_AXES];
            c.offset(J, a, +1); acc += c.read<float>(0, J);
            c.offset(J, a, -1); acc += c.read<float>(0, J);
        }
        c.write<float>(1, c.I, acc);
    }""", "laplace")
    k(bx, fields=[x, y])
    X = x.to_numpy().astype(np.float32)
    act = np.zeros_like(X, bool)
    idx = ti.active_indices(bx)
    act[idx[:, 0], idx[:, 1], idx[:, 2]] = True
    P = np.pad(X, 1)
    lap = (-6.0 * X + P[2:, 1:-1, 1:-1] + P[:-2, 1:-1, 1:-1] + P[1:-1, 2:, 1:-1] + P[1:-1, :-2, 1:-1]
           + P[1:-1, 1:-1, 2:] + P[1:-1, 1:-1, :-2]).astype(np.float32)
    Y = y.to_numpy()
    assert np.allclose(Y[act], lap[act], rtol=1e-6, atol=1e-6)
    assert np.array_equal(ti.active_indices(by), idx)                 # y is active exactly where the loop ran
    assert np.all(Y[~act] == 0)


def test_compile_error_surfaces_log(ti):
    from taichi_b200._lib import B2RuntimeError
    with pytest.raises(B2RuntimeError) as e:
        ti.struct_for("B2_BODY(c) { undeclared_symbol = 1; }", "broken")
    assert "undeclared_symbol" in str(e.value) and "broken" in str(e.value)


def test_user_p2g_range_for_then_grid_struct_for(ti):
    """The reference's plain (non-BLS) MPM idiom written by the user: a particle range-for that scatters with
    quadratic B-spline weights into a SPARSE grid through activating atomic adds (mpm3d.py:40-55), followed by a
    struct-for over the active grid cells (mpm3d.py:56-61).  Grid mass / momentum against NumPy, active blocks
    against the touched-block set (x64 semantics)."""
    import torch
    from test_gpu_mpm import expected_blocks
    n_grid, N = 64, 20000
    gm, gv = ti.field(ti.f32), ti.field(ti.f32)
    blk = ti.root.pointer(ti.ijk, 4).bitmasked(ti.ijk, 4).dense(ti.ijk, 4)
    blk.place(gm, gv)
    rng = np.random.default_rng(7)
    x = (rng.random((N, 3)) * 0.3 + 0.3).astype(np.float32)
    vy = (rng.random(N) - 0.5).astype(np.float32)
    k = ti.struct_for("""
    B2_RANGE_BODY(c) {
      const float *x = c.arg<const float *>(0) + 3 * c.i;
      const float vy = c.arg<const float *>(1)[c.i];
      const float inv_dx = c.arg<float>(2);
      int base[3]; float w[3][3];
      for (int d = 0; d < 3; d++) {
        const float X = x[d] * inv_dx;
        base[d] = (int)(X - 0.5f);
        const float fx = X - (float)base[d];
        w[0][d] = 0.5f * (1.5f - fx) * (1.5f - fx);
        w[1][d] = 0.75f - (fx - 1.0f) * (fx - 1.0f);
        w[2][d] = 0.5f * (fx - 0.5f) * (fx - 0.5f);
      }
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) for (int kk = 0; kk < 3; kk++) {
        int J[B2_MAX_AXES] = {base[0] + i, base[1] + j, base[2] + kk};
        const float wt = w[i][0] * w[j][1] * w[kk][2];
        atomicAdd(c.ref<float>(0, J), wt);          // grid_m[base + offset] += weight * p_mass (p_mass = 1)
        atomicAdd(c.ref<float>(1, J), wt * vy);     // grid_v.y
      }
    }
    B2_BODY(c) {                                    // for I in ti.grouped(grid_m): if m > 0: v /= m
      const float m = c.read<float>(0, c.I);
      if (m > 0.0f) *c.ref<float>(1, c.I) = c.read<float>(1, c.I) / m;
    }""", "user_mpm")
    dx, dvy = torch.from_numpy(x).cuda(), torch.from_numpy(vy).cuda()
    k.range(0, N, fields=[gm, gv], args=[dx, dvy, float(n_grid)])
    M = gm.to_numpy()
    Xg = x * np.float32(n_grid)
    base = (Xg - np.float32(0.5)).astype(np.int32)
    fx = Xg - base.astype(np.float32)
    w = np.stack([np.float32(0.5) * (np.float32(1.5) - fx) ** 2, np.float32(0.75) - (fx - np.float32(1.0)) ** 2,
                  np.float32(0.5) * (fx - np.float32(0.5)) ** 2])           # (3, N, 3)
    want_m = np.zeros((n_grid,) * 3, np.float64)
    want_p = np.zeros((n_grid,) * 3, np.float64)
    for i in range(3):
        for j in range(3):
            for kk in range(3):
                wt = (w[i, :, 0] * w[j, :, 1] * w[kk, :, 2]).astype(np.float64)
                np.add.at(want_m, (base[:, 0] + i, base[:, 1] + j, base[:, 2] + kk), wt)
                np.add.at(want_p, (base[:, 0] + i, base[:, 1] + j, base[:, 2] + kk), wt * vy)
    assert np.allclose(M, want_m, rtol=2e-5, atol=1e-6)
    assert abs(M.sum() - N) < 1e-3 * N * 1e-2                             # partition of unity
    got_blocks = ti.active_indices(blk.parent())                          # bitmasked cells = 4^3 blocks
    assert np.array_equal(got_blocks, expected_blocks(x, n_grid, halo=False))
    k(blk, fields=[gm, gv])                                                # struct-for over active cells
    V = gv.to_numpy()
    nz = want_m > 0
    assert np.allclose(V[nz], (want_p[nz] / want_m[nz]), rtol=1e-3, atol=1e-5)
    assert np.array_equal(ti.active_indices(blk.parent()), got_blocks)     # the struct-for activated nothing new


def test_taichi_sparse_capi_example(ti):
    """c_api/tests/taichi_sparse_test.cpp + tests/cpp/aot/python_scripts/taichi_sparse_test_.py: three pointer levels
    over a dense(4^2) leaf, fill_img / block1_deactivate_all / activate(t) / paint / check_img_value, the kernels
    written as JIT bodies.  Known answer of the reference's own check: sum_{i<20} img[i, i] == 15.25; on top of it the
    whole image against a NumPy restatement of the same pipeline (taichi_logo: examples/patterns.py:5-36)."""
    n = 512
    res = n + n // 4 + n // 16 + n // 64
    x = ti.field(ti.i32)
    img = ti.field(ti.f32, shape=(res, res))
    block1 = ti.root.pointer(ti.ij, n // 64)
    block2 = block1.pointer(ti.ij, 4)
    block3 = block2.pointer(ti.ij, 4)
    leaf = block3.dense(ti.ij, 4)
    leaf.place(x)
    k = ti.struct_for("""
    __device__ int taichi_logo(float px, float py) {          // examples/patterns.py:5-36, scale = 1 / 1.11
      const float scale = 1.0f / 1.11f;
      const float x = (px - 0.5f) / scale + 0.5f, y = (py - 0.5f) / scale + 0.5f;
      auto d2 = [&](float cx, float cy) { return (x - cx) * (x - cx) + (y - cy) * (y - cy); };
      int ret = -1;
      if (!(d2(0.5f, 0.5f) <= 0.52f * 0.52f)) { if (ret == -1) ret = 0; }
      if (!(d2(0.5f, 0.5f) <= 0.495f * 0.495f)) { if (ret == -1) ret = 1; }
      if (d2(0.5f, 0.25f) <= 0.08f * 0.08f) { if (ret == -1) ret = 1; }
      if (d2(0.5f, 0.75f) <= 0.08f * 0.08f) { if (ret == -1) ret = 0; }
      if (d2(0.5f, 0.25f) <= 0.25f * 0.25f) { if (ret == -1) ret = 0; }
      if (d2(0.5f, 0.75f) <= 0.25f * 0.25f) { if (ret == -1) ret = 1; }
      if (x < 0.5f) { if (ret == -1) ret = 1; } else { if (ret == -1) ret = 0; }
      return 1 - ret;
    }
    __device__ int scatter(int i) { return i + i / 4 + i / 16 + i / 64 + 2; }
    B2_BODY(c) { c.write<float>(0, c.I, c.arg<float>(0)); }                 // fill_img: img.fill(0.05)
    B2_RANGE_BODY(c) {                                                     // for i, j in ti.ndrange(n, n)
      const int n = c.arg<int>(1), mode = c.arg<int>(2);
      const int i = (int)(c.i / n), j = (int)(c.i % n);
      int I[B2_MAX_AXES] = {i, j};
      if (mode == 0) {                                                     // activate(t)
        const float t = c.arg<float>(0);
        float px = (float)i / (float)n - 0.5f, py = (float)j / (float)n - 0.5f;
        const float cs = cosf(sinf(t)), sn = sinf(sinf(t));               // rotation2d(sin(t)) @ (p - 0.5) + 0.5
        const float qx = cs * px - sn * py + 0.5f, qy = sn * px + cs * py + 0.5f;
        if (taichi_logo(qx, qy) == 0) c.write<int>(1, I, 1);              // x[i, j] = 1 (activates 3 pointer levels)
      } else {                                                             // paint()
        int t = c.read<int>(1, I);
        t += c.is_active_up(1, I, 3) + c.is_active_up(1, I, 2) + c.is_active_up(1, I, 1);   // block1, block2, block3
        int P[B2_MAX_AXES] = {scatter(i), scatter(j)};
        c.write<float>(0, P, 1.0f - (float)t / 4.0f);
      }
    }""", "taichi_sparse")
    k(img.parent_snode, fields=[img], args=[np.float32(0.05)])
    x[100, 100] = 1                                   # something to deactivate
    assert len(ti.active_indices(block1)) == 1
    block1.deactivate_all()                           # block1_deactivate_all (children first, GC)
    assert len(ti.active_indices(block1)) == 0
    val = np.float32(0.05)
    k.range(0, n * n, fields=[img, x], args=[val, np.int32(n), np.int32(0)])
    k.range(0, n * n, fields=[img, x], args=[val, np.int32(n), np.int32(1)])
    im = img.to_numpy()
    acc = np.float32(0)
    for i in range(20):                                                    # check_img_value: s: ti.f32, s += img[i, i]
        acc = np.float32(acc + im[i, i])
    assert acc == np.float32(15.25)
    # NumPy restatement of the whole pipeline (fp32 like the kernels)
    f = np.float32
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    px, py = (ii.astype(f) / f(n) - f(0.5)), (jj.astype(f) / f(n) - f(0.5))
    cs, sn = np.cos(np.sin(val, dtype=f), dtype=f), np.sin(np.sin(val, dtype=f), dtype=f)
    qx, qy = cs * px - sn * py + f(0.5), sn * px + cs * py + f(0.5)
    sc = f(1.0) / f(1.11)
    X, Y = (qx - f(0.5)) / sc + f(0.5), (qy - f(0.5)) / sc + f(0.5)
    d2 = lambda cx, cy: (X - f(cx)) ** 2 + (Y - f(cy)) ** 2
    ret = np.full((n, n), -1)
    for cond, v in [(~(d2(0.5, 0.5) <= f(0.52) * f(0.52)), 0), (~(d2(0.5, 0.5) <= f(0.495) * f(0.495)), 1),
                    (d2(0.5, 0.25) <= f(0.08) * f(0.08), 1), (d2(0.5, 0.75) <= f(0.08) * f(0.08), 0),
                    (d2(0.5, 0.25) <= f(0.25) * f(0.25), 0), (d2(0.5, 0.75) <= f(0.25) * f(0.25), 1),
                    (X < f(0.5), 1), (np.ones((n, n), bool), 0)]:
        ret = np.where((ret == -1) & cond, v, ret)
    on = (1 - ret) == 0
    xs = x.to_numpy()
    mism = (xs == 1) != on
    assert mism.mean() < 2e-4, mism.sum()            # fp32 rounding on a handful of boundary pixels (libdevice vs NumPy sin/cos)
    act = lambda blk: blk.repeat(n // blk.shape[0], 0).repeat(n // blk.shape[1], 1)
    def level(cells):
        m = np.zeros((cells, cells), bool)
        s_ = n // cells
        np.logical_or.at(m, (ii[xs == 1] // s_, jj[xs == 1] // s_), True)
        return act(m)
    t = xs + level(n // 64).astype(int) + level(n // 16).astype(int) + level(n // 4).astype(int)
    want = np.full((res, res), f(0.05))
    s_idx = np.arange(n) + np.arange(n) // 4 + np.arange(n) // 16 + np.arange(n) // 64 + 2
    want[np.ix_(s_idx, s_idx)] = f(1.0) - t.astype(f) / f(4.0)
    assert np.array_equal(im, want)                  # given the same x, paint is exact
    assert np.array_equal(ti.active_indices(block3), np.argwhere(level(n // 4)[::4, ::4]))


COUNT_CELLS = "B2_BODY(c) { atomicAdd(c.arg<int *>(0), 1); }"


def _count(ti, snode, fields):
    import torch
    s = torch.zeros(1, dtype=torch.int32, device="cuda")
    ti.struct_for(COUNT_CELLS, "count")(snode, fields=fields, args=[s])
    return int(s.item())


def test_sparse_basics_known_answers(ti):
    """tests/python/test_sparse_basics.py:8-113 with the kernels as JIT bodies: the reference's own expected counts."""
    import torch
    n = 128
    # test_pointer (:8-28): for i in x: s += 1  -> 256
    x = ti.field(ti.f32)
    ptr = ti.root.pointer(ti.i, n)
    leaf = ptr.dense(ti.i, n)
    leaf.place(x)
    # test_pointer2 (:74-100): pointer -> pointer -> dense, 7 writes into 5 distinct leaf containers -> 5 n
    y = ti.field(ti.f32)
    leaf2 = ti.root.pointer(ti.i, n).pointer(ti.i, n).dense(ti.i, n)
    leaf2.place(y)
    # test_pointer_is_active_2 (:52-71): root.dense(i).pointer(j).place -> 3 active cells
    z = ti.field(ti.f32)
    pz = ti.root.dense(ti.i, n).pointer(ti.j, n)
    pz.place(z)
    for q in (0, 127, 256):
        x[q] = 1
    assert _count(ti, leaf, [x]) == 256
    # test_pointer_is_active (:31-49): for i in range(n * n): s += is_active(ptr, rescale_index(x, ptr, [i])) -> 256
    k = ti.struct_for("""B2_RANGE_BODY(c) {
        int I[B2_MAX_AXES] = {(int)c.i};
        atomicAdd(c.arg<int *>(0), (int)c.is_active_up(0, I, 1));
    }""", "is_active_ptr")
    s = torch.zeros(1, dtype=torch.int32, device="cuda")
    k.range(0, n * n, fields=[x], args=[s])
    assert int(s.item()) == 256
    for q in (0, 127, 254, 256 + n * n, 257 + n * n, 257 + n * n * 2, 257 + n * n * 5):
        y[q] = 1
    assert _count(ti, leaf2, [y]) == 5 * n
    assert y[257 + n * n * 7] == 0                      # the read of an inactive cell (:99) ...
    assert _count(ti, leaf2, [y]) == 5 * n              # ... activates nothing (:100)
    for (a, b) in ((0, 0), (0, 127), (127, 127)):
        z[a, b] = 1
    k2 = ti.struct_for("""B2_RANGE_BODY(c) {
        const int n = c.arg<int>(1);
        int I[B2_MAX_AXES] = {(int)(c.i / n), (int)(c.i % n)};
        atomicAdd(c.arg<int *>(0), (int)c.is_active_up(0, I, 0));
    }""", "is_active_2")
    s.zero_()
    k2.range(0, n * n, fields=[z], args=[s, np.int32(n)])
    assert int(s.item()) == 3


def test_sparse_parallel_activation_and_gc(ti):
    """tests/python/test_sparse_parallel.py:5-78: activation from inside kernels by many threads at once (the lock-free
    pointer activation under contention), the reference's expected sums, and its fill / clear cycle x10 on a
    512 x 512 pointer grid of dense 8 x 8 blocks with in-kernel ti.deactivate + GC after the task."""
    import torch
    n = 128
    x = ti.field(ti.f32)
    ptr = ti.root.pointer(ti.i, n)
    leaf = ptr.dense(ti.i, n)
    leaf.place(x)
    a = ti.field(ti.f32)
    N = 512
    pa = ti.root.pointer(ti.ij, [N, N])
    la = pa.dense(ti.ij, [8, 8])
    la.place(a)
    k = ti.struct_for("""
    B2_RANGE_BODY(c) {
      const int mode = c.arg<int>(1), n = c.arg<int>(2);
      if (mode == 0) { int I[B2_MAX_AXES] = {(int)c.i * n}; c.write<float>(0, I, 0.0f); }           // x[i * n] = 0
      else if (mode == 1) { int I[B2_MAX_AXES] = {(int)c.i}; c.write<float>(0, I, (float)c.i); }    // x[i] = i
      else { int I[B2_MAX_AXES] = {(int)(c.i / n), (int)(c.i % n)}; c.write<float>(0, I, 2.0f); }   // a[i, j] = 2.0
    }
    B2_BODY(c) {
      const int mode = c.arg<int>(1);
      if (mode == 0) atomicAdd(c.arg<int *>(0), 1);                                                  // s += 1
      else if (mode == 1) atomicAdd(c.arg<int *>(0), c.I[0]);                                        // s += i
      else c.deactivate_up(0, c.I, 1);                        // ti.deactivate(ptr, rescale_index(a, ptr, [i, j]))
    }""", "sparse_parallel")
    s = torch.zeros(1, dtype=torch.int32, device="cuda")
    k.range(0, n, fields=[x], args=[s, np.int32(0), np.int32(n)])
    k(leaf, fields=[x], args=[s, np.int32(0)])
    assert int(s.item()) == n * n                                                                    # :27
    s.zero_()
    k.range(0, n * n, fields=[x], args=[s, np.int32(1), np.int32(n)])
    k(leaf, fields=[x], args=[s, np.int32(1)])
    NN = n * n
    assert int(s.item()) == NN * (NN - 1) // 2                                                       # :52
    assert np.array_equal(x.to_numpy(), np.arange(NN, dtype=np.float32))
    for it in range(10):                                                                             # :55-78
        k.range(0, (N * 8) * (N * 8), fields=[a], args=[s, np.int32(2), np.int32(N * 8)])
        if it == 0:
            assert len(ti.active_indices(pa)) == N * N and pa._num_dynamically_allocated == N * N
        k(la, fields=[a], args=[s, np.int32(2)], gc=[pa])
        ti.sync()
        assert len(ti.active_indices(pa)) == 0
    assert pa._num_dynamically_allocated == N * N      # ten cycles re-used the same N*N nodes through the GC (test_gc.py:40-52) ...
    k.range(0, 64, fields=[a], args=[s, np.int32(2), np.int32(N * 8)])
    assert a[0, 5] == 2.0 and a[0, 70] == 0.0           # ... zero-filled: a recycled block reads 0 outside the new writes
    assert a.to_numpy().sum() == 2.0 * 64


def test_particle_id_lists_per_block(ti):
    """The reference's sparse-MPM idiom (tests/python/test_mpm_particle_list.py:9-53, bls_test_template.py:148-201):
    pointer -> pointer -> {dense voxels, dynamic particle-id list}; a particle loop appends every particle to the list of
    its block (ti.append from thousands of threads), a struct-for over the dynamic SNode visits the ids, and
    `grid.deactivate_all(); build_pid()` repeated must not leak (the whole pool would be gone after ~110 rounds)."""
    import torch
    N, inv_dx = 10000, 128.0
    pid, grid_m = ti.field(ti.i32), ti.field(ti.f32)
    grid = ti.root.pointer(ti.ij, 32)
    block = grid.pointer(ti.ij, 16)
    voxel = block.dense(ti.ij, 8)
    voxel.place(grid_m)
    lists = block.dynamic(ti.k, 1024 * 1024, chunk_size=4096)
    lists.place(pid)
    g = torch.Generator().manual_seed(5)
    x = (torch.rand(N, 2, generator=g) * 0.5).cuda()
    k = ti.struct_for("""
    __device__ void base_of(const float *x, float inv_dx, int *b) {       // floor(x * inv_dx - 0.5) + 1
      b[0] = (int)floorf(x[0] * inv_dx - 0.5f) + 1;
      b[1] = (int)floorf(x[1] * inv_dx - 0.5f) + 1;
    }
    B2_RANGE_BODY(c) {                                                     // build_pid
      int b[2];
      base_of(c.arg<const float *>(0) + 2 * c.i, c.arg<float>(1), b);
      int J[B2_MAX_AXES] = {b[0] >> 3, b[1] >> 3};                         // rescale to the block
      c.append<int>(0, J, (int)c.i);
    }
    B2_BODY(c) {                                                           // for I in ti.grouped(pid): p = pid[I]
      const int p = c.read<int>(0, c.I);
      int b[2];
      base_of(c.arg<const float *>(0) + 2 * p, c.arg<float>(1), b);
      unsigned long long *acc = c.arg<unsigned long long *>(2);
      atomicAdd(acc + 0, 1ull);
      atomicAdd(acc + 1, (unsigned long long)p);
      atomicAdd(acc + 2, (unsigned long long)((b[0] >> 3) != c.I[0] || (b[1] >> 3) != c.I[1]));   // listed under a wrong block
      atomicAdd(acc + 3, (unsigned long long)(c.I[2] >= c.length(0, c.I)));                       // index beyond the list
      int V[B2_MAX_AXES] = {b[0], b[1]};
      atomicAdd(c.ref<float>(1, V), 1.0f);                                 // grid_m[base] += 1 (voxel resolution)
    }""", "particle_lists")
    rounds = 150
    for _ in range(rounds):
        grid.deactivate_all()
        k.range(0, N, fields=[pid, grid_m], args=[x, np.float32(inv_dx)])
    acc = torch.zeros(4, dtype=torch.int64, device="cuda")
    k(lists, fields=[pid, grid_m], args=[x, np.float32(inv_dx), acc])
    assert acc.tolist() == [N, N * (N - 1) // 2, 0, 0]
    base = np.floor(x.cpu().numpy() * np.float32(inv_dx) - np.float32(0.5)).astype(np.int64) + 1
    blk = base >> 3
    want_len = np.zeros((512, 512), np.int64)
    np.add.at(want_len, (blk[:, 0], blk[:, 1]), 1)
    occupied = np.argwhere(want_len > 0)
    assert np.array_equal(ti.length(lists, occupied), want_len[occupied[:, 0], occupied[:, 1]])
    assert np.array_equal(ti.active_indices(block), occupied)
    M = grid_m.to_numpy()
    want_m = np.zeros_like(M)
    np.add.at(want_m, (base[:, 0], base[:, 1]), 1.0)
    assert np.array_equal(M, want_m)
    # no leakage: nodes and list chunks were recycled through the GC round after round
    assert block._num_dynamically_allocated == len(occupied)
    assert lists._num_dynamically_allocated <= 2 * len(occupied)


@pytest.mark.gpu
def test_svd3_and_svd2_in_range_for(ti):
    """ti.svd inside a kernel body (test_svd.py:27-63) through b2_svd.cuh: U^T U = V^T V = I, U sig V^T = A at the
    reference's f32 tolerance 1e-5, singular values against numpy.linalg, det U = det V = +1."""
    import torch
    N = 50000
    rng = np.random.default_rng(21)
    A3 = rng.standard_normal((N, 3, 3)).astype(np.float32)
    A3[0] = [[1, 1, 3], [9, -3, 2], [-3, 4, 2]]
    A3[1] = 0
    A3[2] = np.eye(3)
    A3[3] = np.outer([1, 2, 3], [3, 2, 1])
    A2 = rng.standard_normal((N, 2, 2)).astype(np.float32)
    A2[0] = [[1, 1], [2, 3]]
    dummy = ti.field(ti.f32)
    ti.root.dense(ti.i, 4).place(dummy)
    k = ti.struct_for("""
    B2_RANGE_BODY(c) {
      const int dim = c.arg<int>(2);
      if (dim == 3) {
        const float *A = c.arg<const float *>(0) + 9 * c.i; float *o = c.arg<float *>(1) + 21 * c.i;
        float U[9], s[3], V[9];
        b2_svd3(A, U, s, V);
        for (int q = 0; q < 9; q++) { o[q] = U[q]; o[12 + q] = V[q]; }
        for (int q = 0; q < 3; q++) o[9 + q] = s[q];
      } else {
        const float *A = c.arg<const float *>(0) + 4 * c.i; float *o = c.arg<float *>(1) + 10 * c.i;
        float U[4], s[2], V[4];
        b2_svd2(A, U, s, V);
        for (int q = 0; q < 4; q++) { o[q] = U[q]; o[6 + q] = V[q]; }
        for (int q = 0; q < 2; q++) o[4 + q] = s[q];
      }
    }""", "svd")
    for n, A in ((3, A3), (2, A2)):
        dA = torch.from_numpy(A).cuda()
        out = torch.zeros(N, n * n * 2 + n, device="cuda")
        k.range(0, N, fields=[dummy], args=[dA, out, np.int32(n)])
        o = out.cpu().numpy()
        U, s, V = o[:, :n * n].reshape(N, n, n), o[:, n * n:n * n + n], o[:, n * n + n:].reshape(N, n, n)
        sc = np.maximum(1.0, np.abs(A).max(axis=(1, 2)))
        rec = np.einsum("nij,nj,nkj->nik", U, s, V)
        assert (np.abs(rec - A).max(axis=(1, 2)) / sc).max() < 1e-5
        eye = np.eye(n, dtype=np.float32)
        assert np.abs(np.einsum("nji,njk->nik", U, U) - eye).max() < 1e-5
        assert np.abs(np.einsum("nji,njk->nik", V, V) - eye).max() < 1e-5
        assert np.abs(np.linalg.det(U.astype(np.float64)) - 1).max() < 1e-4
        assert np.abs(np.linalg.det(V.astype(np.float64)) - 1).max() < 1e-4
        want = np.linalg.svd(A.astype(np.float64), compute_uv=False)
        assert (np.abs(-np.sort(-np.abs(s), axis=1) - want).max(axis=1) / sc).max() < 1e-5
        assert np.all(s[:, :-1] >= -1e-5 * sc[:, None])


@pytest.mark.gpu
def test_3d_constitutive_model_range_for(ti):
    """The F/SVD material update of mpm128.py:36-63 (hardening, snow plasticity clamp, liquid reset, fixed-corotated
    stress) written for 3-D matrices in a user range-for, Jp in a field; against a float64 NumPy evaluation of the
    same formulas on numpy.linalg.svd (the update is invariant to the SVD's ordering and sign conventions)."""
    import torch
    N = 30000
    rng = np.random.default_rng(33)
    dt, E, nu = 1e-4, 1e3, 0.2
    mu0, la0 = E / (2 * (1 + nu)), E * nu / ((1 + nu) * (1 - 2 * nu))
    q = np.linalg.qr(rng.standard_normal((N, 3, 3)))[0]
    q *= np.sign(np.linalg.det(q))[:, None, None]
    F0 = (q @ (np.eye(3) + 0.04 * rng.standard_normal((N, 3, 3)))).astype(np.float32)
    C0 = (20.0 * rng.standard_normal((N, 3, 3))).astype(np.float32)
    Jp0 = (1 + 0.05 * rng.standard_normal(N)).astype(np.float32)
    mat = rng.integers(0, 3, N).astype(np.int32)
    Jp = ti.field(ti.f32)
    ti.root.dense(ti.i, 32768).place(Jp)
    Jp.from_numpy(np.concatenate([Jp0, np.zeros(32768 - N, np.float32)]))
    k = ti.struct_for("""
    B2_RANGE_BODY(c) {
      const int p = c.i;
      float *Fp = c.arg<float *>(0) + 9 * p; const float *C = c.arg<const float *>(1) + 9 * p;
      const int material = c.arg<const int *>(2)[p]; float *out = c.arg<float *>(3) + 9 * p;
      const float dt = c.arg<float>(4), mu_0 = c.arg<float>(5), lambda_0 = c.arg<float>(6);
      float F[9];
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {      // F = (I + dt C) F
        float a = 0.f;
        for (int m = 0; m < 3; m++) a += ((i == m ? 1.f : 0.f) + dt * C[i * 3 + m]) * Fp[m * 3 + j];
        F[i * 3 + j] = a;
      }
      int I[B2_MAX_AXES] = {p};
      float jp = c.read<float>(0, I);
      float h = fmaxf(0.1f, fminf(5.f, expf(10.f * (1.f - jp))));
      if (material == 1) h = 0.3f;
      float mu = mu_0 * h; const float la = lambda_0 * h;
      if (material == 0) mu = 0.f;
      float U[9], sig[3], V[9];
      b2_svd3(F, U, sig, V);
      float J = 1.f;
      for (int d = 0; d < 3; d++) {
        float ns = sig[d];
        if (material == 2) ns = fminf(fmaxf(sig[d], 1.f - 2.5e-2f), 1.f + 4.5e-3f);
        jp *= sig[d] / ns; sig[d] = ns; J *= ns;
      }
      c.write<float>(0, I, jp);
      float R[9];
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
        R[i * 3 + j] = U[i * 3] * V[j * 3] + U[i * 3 + 1] * V[j * 3 + 1] + U[i * 3 + 2] * V[j * 3 + 2];
        if (material == 0) F[i * 3 + j] = i == j ? cbrtf(J) : 0.f;
        else if (material == 2)
          F[i * 3 + j] = U[i * 3] * sig[0] * V[j * 3] + U[i * 3 + 1] * sig[1] * V[j * 3 + 1] + U[i * 3 + 2] * sig[2] * V[j * 3 + 2];
      }
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {      // 2 mu (F - U V^T) F^T + la J (J - 1) I
        float a = 0.f;
        for (int m = 0; m < 3; m++) a += (F[i * 3 + m] - R[i * 3 + m]) * F[j * 3 + m];
        out[i * 3 + j] = 2.f * mu * a + (i == j ? la * J * (J - 1.f) : 0.f);
      }
      for (int q = 0; q < 9; q++) Fp[q] = F[q];
    }""", "material3d")
    dF, dC, dm = torch.from_numpy(F0).cuda(), torch.from_numpy(C0).cuda(), torch.from_numpy(mat).cuda()
    dS = torch.zeros(N, 9, device="cuda")
    k.range(0, N, fields=[Jp], args=[dF, dC, dm, dS, np.float32(dt), np.float32(mu0), np.float32(la0)])
    gotF, gotS, gotJp = dF.cpu().numpy().astype(np.float64), dS.cpu().numpy().reshape(N, 3, 3).astype(np.float64), Jp.to_numpy()[:N]
    # float64 evaluation of the same formulas
    F = (np.eye(3) + dt * C0.astype(np.float64)) @ F0.astype(np.float64)
    h = np.clip(np.exp(10 * (1 - Jp0.astype(np.float64))), 0.1, 5)
    h[mat == 1] = 0.3
    mu, la = mu0 * h, la0 * h
    mu[mat == 0] = 0
    U, sig, Vt = np.linalg.svd(F)
    ns = np.where((mat == 2)[:, None], np.clip(sig, 1 - 2.5e-2, 1 + 4.5e-3), sig)
    wantJp = Jp0 * np.prod(sig / ns, axis=1)
    J = np.prod(ns, axis=1)
    R = U @ Vt
    Fn = F.copy()
    Fn[mat == 0] = np.eye(3) * np.cbrt(J[mat == 0])[:, None, None]
    Fn[mat == 2] = (U * ns[:, None, :] @ Vt)[mat == 2]
    wantS = 2 * mu[:, None, None] * ((Fn - R) @ np.swapaxes(Fn, 1, 2)) + np.eye(3) * (la * J * (J - 1))[:, None, None]
    assert np.allclose(gotJp, wantJp, rtol=2e-5, atol=1e-6)
    assert np.abs(gotF - Fn).max() < 5e-6
    assert np.abs(gotS - wantS).max() < 2e-5 * 2 * mu0 * 5 + 1e-3      # |dS| <= 2 mu |dF| |F| at eps-level |dF|
    assert (mat == 2).sum() > 0 and np.abs(ns - sig)[mat == 2].max() > 1e-3   # the clamp was exercised


MPM3D_MATERIAL_SRC = """
__device__ void weights(const float *x, float inv_dx, int *base, float (*w)[3], float *fx) {
  for (int d = 0; d < 3; d++) {
    const float X = x[d] * inv_dx;
    base[d] = (int)(X - 0.5f);
    fx[d] = X - (float)base[d];
    w[0][d] = 0.5f * (1.5f - fx[d]) * (1.5f - fx[d]);
    w[1][d] = 0.75f - (fx[d] - 1.0f) * (fx[d] - 1.0f);
    w[2][d] = 0.5f * (fx[d] - 0.5f) * (fx[d] - 0.5f);
  }
}
struct Par { float dt, inv_dx, p_vol, p_mass, mu_0, lambda_0, gravity; int n_grid, bound, phase; };
B2_RANGE_BODY(c) {
  const int p = c.i;
  float *x = c.arg<float *>(0) + 3 * p, *v = c.arg<float *>(1) + 3 * p, *C = c.arg<float *>(2) + 9 * p;
  float *Fp = c.arg<float *>(3) + 9 * p, *Jp = c.arg<float *>(4) + p;
  const int material = c.arg<const int *>(5)[p];
  const Par P = *c.arg<const Par *>(6);
  const float dx = 1.0f / P.inv_dx;
  int base[3]; float w[3][3], fx[3];
  weights(x, P.inv_dx, base, w, fx);
  if (P.phase == 0) {                                  // ---- P2G (mpm128.py:31-71 in 3-D)
    float F[9];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
      float a = 0.f;
      for (int m = 0; m < 3; m++) a += ((i == m ? 1.f : 0.f) + P.dt * C[i * 3 + m]) * Fp[m * 3 + j];
      F[i * 3 + j] = a;
    }
    float jp = *Jp;
    float h = fmaxf(0.1f, fminf(5.f, expf(10.f * (1.f - jp))));
    if (material == 1) h = 0.3f;
    float mu = P.mu_0 * h; const float la = P.lambda_0 * h;
    if (material == 0) mu = 0.f;
    float U[9], sig[3], V[9];
    b2_svd3(F, U, sig, V);
    float J = 1.f;
    for (int d = 0; d < 3; d++) {
      float ns = sig[d];
      if (material == 2) ns = fminf(fmaxf(sig[d], 1.f - 2.5e-2f), 1.f + 4.5e-3f);
      jp *= sig[d] / ns; sig[d] = ns; J *= ns;
    }
    *Jp = jp;
    float R[9], A[9];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
      R[i * 3 + j] = U[i * 3] * V[j * 3] + U[i * 3 + 1] * V[j * 3 + 1] + U[i * 3 + 2] * V[j * 3 + 2];
      if (material == 0) F[i * 3 + j] = i == j ? cbrtf(J) : 0.f;
      else if (material == 2)
        F[i * 3 + j] = U[i * 3] * sig[0] * V[j * 3] + U[i * 3 + 1] * sig[1] * V[j * 3 + 1] + U[i * 3 + 2] * sig[2] * V[j * 3 + 2];
    }
    const float k = -P.dt * P.p_vol * 4.f * P.inv_dx * P.inv_dx;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
      float a = 0.f;
      for (int m = 0; m < 3; m++) a += (F[i * 3 + m] - R[i * 3 + m]) * F[j * 3 + m];
      A[i * 3 + j] = k * (2.f * mu * a + (i == j ? la * J * (J - 1.f) : 0.f)) + P.p_mass * C[i * 3 + j];
    }
    for (int q = 0; q < 9; q++) Fp[q] = F[q];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) for (int kk = 0; kk < 3; kk++) {
      const float dpos[3] = {((float)i - fx[0]) * dx, ((float)j - fx[1]) * dx, ((float)kk - fx[2]) * dx};
      const float wt = w[i][0] * w[j][1] * w[kk][2];
      int I[B2_MAX_AXES] = {base[0] + i, base[1] + j, base[2] + kk};
      for (int d = 0; d < 3; d++)
        atomicAdd(c.ref<float>(d, I), wt * (P.p_mass * v[d] + A[d * 3] * dpos[0] + A[d * 3 + 1] * dpos[1] + A[d * 3 + 2] * dpos[2]));
      atomicAdd(c.ref<float>(3, I), wt * P.p_mass);
    }
  } else {                                             // ---- G2P (mpm128.py:88-102 in 3-D)
    float nv[3] = {0.f, 0.f, 0.f}, nC[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) for (int kk = 0; kk < 3; kk++) {
      const float dpos[3] = {(float)i - fx[0], (float)j - fx[1], (float)kk - fx[2]};
      const float wt = w[i][0] * w[j][1] * w[kk][2];
      int I[B2_MAX_AXES] = {base[0] + i, base[1] + j, base[2] + kk};
      for (int d = 0; d < 3; d++) {
        const float g = c.read<float>(d, I);
        nv[d] += wt * g;
        for (int e = 0; e < 3; e++) nC[d * 3 + e] += 4.f * P.inv_dx * wt * g * dpos[e];
      }
    }
    for (int d = 0; d < 3; d++) { v[d] = nv[d]; x[d] += P.dt * nv[d]; }
    for (int q = 0; q < 9; q++) C[q] = nC[q];
  }
}
B2_BODY(c) {                                           // ---- grid op (mpm128.py:72-87 in 3-D)
  const Par P = *c.arg<const Par *>(6);
  const float m = c.read<float>(3, c.I);
  if (m > 0.f) {
    float v[3];
    for (int d = 0; d < 3; d++) v[d] = c.read<float>(d, c.I) / m;
    v[1] -= P.dt * P.gravity;
    for (int d = 0; d < 3; d++) {
      if (c.I[d] < P.bound && v[d] < 0.f) v[d] = 0.f;
      if (c.I[d] > P.n_grid - P.bound && v[d] > 0.f) v[d] = 0.f;
      c.write<float>(d, c.I, v[d]);
    }
  }
}
"""


def _np_mpm3d_material_substep(S, par):
    """float64 restatement of MPM3D_MATERIAL_SRC on numpy.linalg.svd."""
    x, v, C, F, Jp, mat = S["x"], S["v"], S["C"], S["F"], S["Jp"], S["mat"]
    n, dt, inv_dx = par["n_grid"], par["dt"], float(par["n_grid"])
    dx = 1.0 / inv_dx
    X = x * inv_dx
    base = (X - 0.5).astype(np.int64)
    fx = X - base
    w = [0.5 * (1.5 - fx) ** 2, 0.75 - (fx - 1.0) ** 2, 0.5 * (fx - 0.5) ** 2]
    F = (np.eye(3) + dt * C) @ F
    h = np.clip(np.exp(10 * (1 - Jp)), 0.1, 5)
    h[mat == 1] = 0.3
    mu, la = par["mu_0"] * h, par["lambda_0"] * h
    mu[mat == 0] = 0
    U, sig, Vt = np.linalg.svd(F)
    ns = np.where((mat == 2)[:, None], np.clip(sig, 1 - 2.5e-2, 1 + 4.5e-3), sig)
    Jp = Jp * np.prod(sig / ns, axis=1)
    J = np.prod(ns, axis=1)
    R = U @ Vt
    F[mat == 0] = np.eye(3) * np.cbrt(J[mat == 0])[:, None, None]
    F[mat == 2] = (U * ns[:, None, :] @ Vt)[mat == 2]
    stress = 2 * mu[:, None, None] * ((F - R) @ np.swapaxes(F, 1, 2)) + np.eye(3) * (la * J * (J - 1))[:, None, None]
    A = (-dt * par["p_vol"] * 4 * inv_dx * inv_dx) * stress + par["p_mass"] * C
    gv, gm = np.zeros((n, n, n, 3)), np.zeros((n, n, n))
    offs = [(i, j, k) for i in range(3) for j in range(3) for k in range(3)]
    for (i, j, k) in offs:
        o = np.array([i, j, k], np.float64)
        wt = w[i][:, 0] * w[j][:, 1] * w[k][:, 2]
        dpos = (o - fx) * dx
        idx = (base[:, 0] + i, base[:, 1] + j, base[:, 2] + k)
        np.add.at(gv, idx, wt[:, None] * (par["p_mass"] * v + np.einsum("nij,nj->ni", A, dpos)))
        np.add.at(gm, idx, wt * par["p_mass"])
    act = gm > 0
    gv[act] /= gm[act][:, None]
    gv[act, 1] -= dt * par["gravity"]
    I = np.indices((n, n, n)).transpose(1, 2, 3, 0)
    gv[act[..., None] & (I < par["bound"]) & (gv < 0)] = 0
    gv[act[..., None] & (I > n - par["bound"]) & (gv > 0)] = 0
    nv, nC = np.zeros_like(v), np.zeros_like(C)
    for (i, j, k) in offs:
        o = np.array([i, j, k], np.float64)
        wt = w[i][:, 0] * w[j][:, 1] * w[k][:, 2]
        g = gv[base[:, 0] + i, base[:, 1] + j, base[:, 2] + k]
        nv += wt[:, None] * g
        nC += 4 * inv_dx * wt[:, None, None] * np.einsum("ni,nj->nij", g, o - fx)
    S.update(x=x + dt * nv, v=nv, C=nC, F=F, Jp=Jp)
    return gm


@pytest.mark.gpu
def test_3d_mpm_with_materials_on_sparse_grid(ti):
    """the full mpm128 constitutive model (liquid / jelly / snow, ti.svd) carried to 3-D and written by
    the USER as run-time bodies over the sparse grid pointer -> bitmasked -> dense(4^3): P2G range-for with activating
    atomic adds, grid struct-for, G2P range-for, `deactivate_all` between substeps.  State after 6 substeps against the
    float64 NumPy restatement; grid mass per substep; active blocks = touched blocks."""
    import ctypes
    import torch
    from test_gpu_mpm import expected_blocks
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
    g = [ti.field(ti.f32) for _ in range(4)]                               # grid_v.x/y/z, grid_m
    top = ti.root.pointer(ti.ijk, 4)
    blk = top.bitmasked(ti.ijk, 4).dense(ti.ijk, 4)
    blk.place(*g)

    class Par(ctypes.Structure):
        _fields_ = [(k, ctypes.c_float) for k in ("dt", "inv_dx", "p_vol", "p_mass", "mu_0", "lambda_0", "gravity")] + \
                   [(k, ctypes.c_int32) for k in ("n_grid", "bound", "phase")]

    def par_dev(phase):
        pb = Par(par["dt"], float(n_grid), par["p_vol"], par["p_mass"], par["mu_0"], par["lambda_0"], par["gravity"],
                 n_grid, par["bound"], phase)
        return torch.frombuffer(bytearray(bytes(pb)), dtype=torch.uint8).cuda()
    P0, P1 = par_dev(0), par_dev(1)
    k = ti.struct_for(MPM3D_MATERIAL_SRC, "mpm3d_material")
    d = [torch.from_numpy(a.copy()).cuda() for a in (x0, v0, C0, F0, Jp0, mat)]
    S = dict(x=x0.astype(np.float64), v=v0.astype(np.float64), C=C0.astype(np.float64), F=F0.astype(np.float64),
             Jp=Jp0.astype(np.float64), mat=mat)
    for s in range(steps):
        x_before = d[0].cpu().numpy()
        top.deactivate_all()
        k.range(0, N, fields=g, args=d + [P0])
        got_m = g[3].to_numpy()
        assert np.array_equal(ti.active_indices(blk.parent()), expected_blocks(x_before, n_grid, halo=False))
        k(blk, fields=g, args=d + [P0])
        k.range(0, N, fields=g, args=d + [P1])
        want_m = _np_mpm3d_material_substep(S, par)
        assert np.allclose(got_m, want_m, rtol=2e-4, atol=par["p_mass"] * 1e-4), s
        assert abs(got_m.sum() - N * par["p_mass"]) < 1e-5 * N * par["p_mass"]
    got = [a.cpu().numpy().astype(np.float64) for a in d[:5]]
    assert np.abs(got[0] - S["x"]).max() < 2e-6
    assert np.abs(got[1] - S["v"]).max() < 2e-3 * max(1.0, np.abs(S["v"]).max())
    assert np.abs(got[2].reshape(N, 3, 3) - S["C"]).max() < 2e-3 * np.abs(S["C"]).max()
    assert np.abs(got[3].reshape(N, 3, 3) - S["F"]).max() < 2e-5
    assert np.allclose(got[4], S["Jp"], rtol=1e-4)
    assert np.abs(S["v"] - v0).max() > 0.05                                  # the substeps did something


STENCIL_SRC = """
B2_RANGE_BODY(c) {                                     // populate: for I in ti.grouped(ti.ndrange(*ndrange)): x[I] = sum I[i] ** (i + 1)
  const int dim = c.arg<int>(2), lo0 = c.arg<int>(4), lo1 = c.arg<int>(5), lo2 = c.arg<int>(6);
  const int n0 = c.arg<int>(7), n1 = c.arg<int>(8), n2 = c.arg<int>(9);
  int r = (int)c.i;
  int I[B2_MAX_AXES] = {0};
  if (dim == 3) { I[2] = lo2 + r % n2; r /= n2; }
  if (dim >= 2) { I[1] = lo1 + r % n1; r /= n1; }
  I[0] = lo0 + r;
  (void)n0;
  c.write<int>(0, I, I[0] + (dim >= 2 ? I[1] * I[1] : 0) + (dim == 3 ? I[2] * I[2] * I[2] : 0));
}
B2_BODY(c) {                                           // apply: for I in ti.grouped(x)
  const int *st = c.arg<const int *>(0); const int ns = c.arg<int>(1), dim = c.arg<int>(2), scatter = c.arg<int>(3);
  int s = 0;
  const int xv = scatter ? c.read<int>(0, c.I) : 0;
  for (int q = 0; q < ns; q++) {
    int J[B2_MAX_AXES] = {0};
    for (int d = 0; d < dim; d++) J[d] = c.I[d] + st[q * 3 + d];
    if (scatter) atomicAdd(c.ref<int>(1, J), xv);      // y[I + offset] += x[I]   (activates)
    else s += c.read<int>(0, J);                       // s += x[I + offset]      (never activates)
  }
  if (!scatter) c.write<int>(1, c.I, s);
}
"""


@pytest.mark.gpu
@pytest.mark.parametrize("dim,N,bs,stencil,scatter", [
    (1, 128, 32, [(0,)], False), (1, 128, 32, [(-1,), (0,)], False),
    (2, 128, 16, [(0, 0), (0, -1), (0, 1), (1, 0)], False), (2, 128, (4, 16), [(0, 0), (0, -1), (0, 1), (1, 0)], False),
    (3, 128, (4, 8, 16), [(-1, -1, -1), (2, 0, 1)], False),
    (1, 128, 32, [(0,)], True), (1, 128, 32, [(1,), (0,)], True),
    (2, 128, 16, [(0, 0), (0, -1), (0, 1), (1, 0)], True), (3, 128, (4, 8, 16), [(-1, -1, -1), (2, 0, 1)], True)])
def test_stencil_gather_scatter_over_pointer_blocks(ti, dim, N, bs, stencil, scatter):
    """The stencil cases of tests/python/test_bls.py:60-121 / bls_test_template.py:8-96 (1-D / 2-D / 3-D, non-square
    blocks (4, 16) and (4, 8, 16), gather and scatter, int32 => exact) as struct-for bodies over pointer -> dense(bs);
    where the reference compares block_local against plain, this compares with the NumPy answer.  (3-D runs at N = 128:
    with the reference's N = 64 the populated range [2 bs, N - 2 bs) is empty along the 16-wide axis.)"""
    import torch
    bs = [bs] * dim if isinstance(bs, int) else list(bs)
    axes = [ti.i, ti.ij, ti.ijk][dim - 1]
    grid = [N // b for b in bs]
    x, y = ti.field(ti.i32), ti.field(ti.i32)
    if scatter:
        block = ti.root.pointer(axes, grid)
        xs, ys = block.dense(axes, bs), block.dense(axes, bs)
    else:
        xs, ys = ti.root.pointer(axes, grid).dense(axes, bs), ti.root.pointer(axes, grid).dense(axes, bs)
    xs.place(x)
    ys.place(y)
    lo = [2 * b for b in bs] + [0] * (3 - dim)
    cnt = [N - 4 * b for b in bs] + [1] * (3 - dim)
    st = np.zeros((len(stencil), 3), np.int32)
    st[:, :dim] = np.array(stencil, np.int32)
    dst = torch.from_numpy(st).cuda()
    args = [dst, np.int32(len(stencil)), np.int32(dim), np.int32(int(scatter))] + [np.int32(v) for v in lo + cnt]
    k = ti.struct_for(STENCIL_SRC, "stencil")
    k.range(0, int(np.prod(cnt)), fields=[x, y], args=args)
    I = np.indices((N,) * dim)
    inside = np.ones((N,) * dim, bool)
    for d in range(dim):
        inside &= (I[d] >= lo[d]) & (I[d] < lo[d] + cnt[d])
    X = np.where(inside, sum(I[d].astype(np.int64) ** (d + 1) for d in range(dim)), 0).astype(np.int32)
    assert np.array_equal(x.to_numpy(), X)
    k(xs, fields=[x, y], args=args)
    want = np.zeros_like(X)
    for off in stencil:
        # gather: y[I] = sum x[I + off];  scatter: y[I + off] += x[I]
        want += np.roll(X, tuple(-o for o in off) if not scatter else tuple(off), axis=tuple(range(dim)))
    if not scatter:
        want = np.where(inside, want, 0)                                  # only cells of x's active blocks are visited
    assert np.array_equal(y.to_numpy(), want)
    assert np.array_equal(x.to_numpy(), X)                                # reads activated nothing in x
    nblk = int(np.prod([c // b for c, b in zip(cnt, bs)]))
    got_blk = len(ti.active_indices(xs.parent()))
    assert got_blk >= nblk if scatter else got_blk == nblk                # a scatter may open neighbouring blocks


@pytest.mark.gpu
def test_stencil_multiple_inputs(ti):
    """test_bls.py:124-157: w[i, j] = x[i, j-2] + y[i+2, j-1] + y[i-1, j] + z[i-1, j] + z[i+1, j] over x's active cells."""
    N, bs = 128, 8
    f = [ti.field(ti.i32) for _ in range(4)]
    blk = ti.root.pointer(ti.ij, N // bs).dense(ti.ij, bs)
    blk.place(*f)
    k = ti.struct_for("""
    B2_RANGE_BODY(c) {
      const int n = c.arg<int>(0), bs = c.arg<int>(1);
      const int i = bs + (int)c.i / n, j = bs + (int)c.i % n;
      int I[B2_MAX_AXES] = {i, j};
      c.write<int>(0, I, i - j); c.write<int>(1, I, i + j * j); c.write<int>(2, I, i * i - j);
    }
    B2_BODY(c) {
      const int i = c.I[0], j = c.I[1];
      int a[B2_MAX_AXES] = {i, j - 2}, b[B2_MAX_AXES] = {i + 2, j - 1}, d[B2_MAX_AXES] = {i - 1, j}, e[B2_MAX_AXES] = {i + 1, j};
      c.write<int>(3, c.I, c.read<int>(0, a) + c.read<int>(1, b) + c.read<int>(1, d) + c.read<int>(2, d) + c.read<int>(2, e));
    }""", "multi")
    n = N - 2 * bs
    k.range(0, n * n, fields=f, args=[np.int32(n), np.int32(bs)])
    k(blk, fields=f, args=[np.int32(n), np.int32(bs)])
    i, j = np.indices((N, N))
    ins = (i >= bs) & (i < N - bs) & (j >= bs) & (j < N - bs)
    X, Y, Z = [np.where(ins, v, 0).astype(np.int32) for v in (i - j, i + j * j, i * i - j)]
    sh = lambda A, di, dj: np.roll(A, (-di, -dj), axis=(0, 1))
    want = np.where(ins, sh(X, 0, -2) + sh(Y, 2, -1) + sh(Y, -1, 0) + sh(Z, -1, 0) + sh(Z, 1, 0), 0)
    assert np.array_equal(f[3].to_numpy(), want)


# --------------------------------------------------------------------------- ti.block_local + thread-local reductions
BLS_STENCIL_SRC = """
B2_RANGE_BODY(c) {                                   // fill x over [lo, lo + cnt)
  const int dim = c.arg<int>(0);
  int I[B2_MAX_AXES] = {0};
  long long r = c.i;
  for (int d = dim - 1; d >= 0; d--) { const int n = c.arg<int>(5 + d); I[d] = c.arg<int>(2 + d) + (int)(r % n); r /= n; }
  int v = 0;
  for (int d = 0, p = 1; d < dim; d++) { p = 1; for (int q = 0; q <= d; q++) p *= I[d]; v += p; }
  c.write<int>(0, I, v);
}
B2_BODY(c) {
  const int dim = c.arg<int>(0), ns = c.arg<int>(1), use_bls = c.arg<int>(8), scatter = c.arg<int>(9);
  const int *st = c.arg<const int *>(10);
  int acc = 0;
  const int mine = scatter ? c.read<int>(0, c.I) : 0;
  for (int s = 0; s < ns; s++) {
    int J[B2_MAX_AXES];
    for (int a = 0; a < B2_MAX_AXES; a++) J[a] = c.I[a] + (a < dim ? st[3 * s + a] : 0);
    if (scatter) {
      if (use_bls) c.bls_add<int>(0, J, mine); else atomicAdd(c.ref<int>(1, J), mine);
    } else {
      acc += use_bls ? c.bls_read<int>(0, J) : c.read<int>(0, J);
    }
  }
  if (!scatter) c.write<int>(1, c.I, acc);
}
"""


@pytest.mark.gpu
@pytest.mark.parametrize("dim,N,bs,stencil,scatter", [
    (1, 128, 32, [(-1,), (0,)], False), (2, 128, 16, [(0, 0), (0, -1), (0, 1), (1, 0)], False),
    (2, 128, (4, 16), [(0, 0), (0, -1), (0, 1), (1, 0)], False), (3, 128, (4, 8, 16), [(-1, -1, -1), (2, 0, 1)], False),
    (1, 128, 32, [(1,), (0,)], True), (2, 128, 16, [(0, 0), (0, -1), (0, 1), (1, 0)], True),
    (3, 128, (4, 8, 16), [(-1, -1, -1), (2, 0, 1)], True)])
def test_block_local_equals_plain_stencils(ti, dim, N, bs, stencil, scatter):
    """tests/python/test_bls.py:60-121 / bls_test_template.py:8-96: the same stencil body once through a block-local pad
    (ti.block_local(x) for gathers, ti.block_local(y) for scatters) and once through plain global accesses must agree
    exactly (int32), and both with NumPy; the pad geometry is ScratchPad's (block + offsets)."""
    import torch
    bs = [bs] * dim if isinstance(bs, int) else list(bs)
    axes = [ti.i, ti.ij, ti.ijk][dim - 1]
    grid = [N // b for b in bs]
    x, y, y2 = ti.field(ti.i32), ti.field(ti.i32), ti.field(ti.i32)
    block = ti.root.pointer(axes, grid)
    xs, ys, ys2 = block.dense(axes, bs), block.dense(axes, bs), block.dense(axes, bs)
    xs.place(x), ys.place(y), ys2.place(y2)
    lo = [2 * b for b in bs] + [0] * (3 - dim)
    cnt = [N - 4 * b for b in bs] + [1] * (3 - dim)
    st = np.zeros((len(stencil), 3), np.int32)
    st[:, :dim] = np.array(stencil, np.int32)
    dst = torch.from_numpy(st).cuda()

    def args(use_bls):
        return [np.int32(dim), np.int32(len(stencil))] + [np.int32(v) for v in lo + cnt] + [np.int32(use_bls), np.int32(int(scatter)), dst]
    k = ti.struct_for(BLS_STENCIL_SRC, "bls_stencil")
    k.range(0, int(np.prod(cnt)), fields=[x, y], args=args(0))
    off = np.array(stencil)
    pad = ti.bls(1 if scatter else 0, "add" if scatter else "read", off.min(0), off.max(0))
    k(xs, fields=[x, y], args=args(1), block_local=[pad])
    k(xs, fields=[x, y2], args=args(0))
    I = np.indices((N,) * dim)
    inside = np.ones((N,) * dim, bool)
    for d in range(dim):
        inside &= (I[d] >= lo[d]) & (I[d] < lo[d] + cnt[d])
    X = np.where(inside, sum(I[d].astype(np.int64) ** (d + 1) for d in range(dim)), 0).astype(np.int32)
    want = np.zeros_like(X)
    for o in stencil:
        want += np.roll(X, tuple(-q for q in o) if not scatter else tuple(o), axis=tuple(range(dim)))
    if not scatter:
        want = np.where(inside, want, 0)
    assert np.array_equal(y.to_numpy(), y2.to_numpy())       # BLS == non-BLS (bls_test_template.py:85-96)
    assert np.array_equal(y.to_numpy(), want)
    assert np.array_equal(x.to_numpy(), X)


@pytest.mark.gpu
@pytest.mark.parametrize("halo", [False, True])
def test_block_local_particle_grid(ti, halo):
    """bls_test_template.py:99-287 bls_particle_grid (N = 128, ppc = 10, block 8, pointer level 1): per-block particle-id
    lists (block.dynamic), struct-for over the lists, P2G scatter into ti.block_local(m), then G2P gather out of
    ti.block_local(m) -- BLS and plain agree within 1e-4 (float sums reorder), and with NumPy; halo=True activates the
    whole pad like the reference's CUDA epilogue, halo=False only touched cells."""
    import torch
    N, bs, ppc = 128, 8, 10
    m1, m2, pid = ti.field(ti.f32), ti.field(ti.f32), ti.field(ti.i32)
    block = ti.root.pointer(ti.ij, N // bs)
    block.dense(ti.ij, bs).place(m1)
    block.dense(ti.ij, bs).place(m2)
    plist = block.dynamic(ti.l, 4096, chunk_size=64)
    plist.place(pid)
    rng = np.random.default_rng(5)
    npart = ppc * (N // 2) ** 2
    xp = (rng.random((npart, 2)) * 0.5 + 0.2).astype(np.float32)
    dx = torch.from_numpy(xp).cuda()
    out = torch.zeros(2, npart, device="cuda")
    k = ti.struct_for("""
    B2_RANGE_BODY(c) {                                // build_pid: ti.append(pid.parent(), block index, p)
      const float2 x = c.arg<const float2 *>(0)[c.i];
      const int N = c.arg<int>(1), bs = c.arg<int>(2);
      const int bi = (int)floorf(x.x * N - 0.5f), bj = (int)floorf(x.y * N - 0.5f);
      int B[B2_MAX_AXES] = {bi / bs, bj / bs};
      c.append<int>(0, B, (int)c.i);
    }
    B2_BODY(c) {                                      // for I in pid: p = pid[I]; scatter / gather around base
      const int p = c.read<int>(0, c.I);
      const float2 x = c.arg<const float2 *>(0)[p];
      const int N = c.arg<int>(1), use_bls = c.arg<int>(3), phase = c.arg<int>(4);
      const int bi = (int)floorf(x.x * N - 0.5f), bj = (int)floorf(x.y * N - 0.5f);
      const float fx = x.x * N - bi, fy = x.y * N - bj;
      const float wx[3] = {0.5f * (1.5f - fx) * (1.5f - fx), 0.75f - (fx - 1.0f) * (fx - 1.0f), 0.5f * (fx - 0.5f) * (fx - 0.5f)};
      const float wy[3] = {0.5f * (1.5f - fy) * (1.5f - fy), 0.75f - (fy - 1.0f) * (fy - 1.0f), 0.5f * (fy - 0.5f) * (fy - 0.5f)};
      float acc = 0.0f;
      for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
          int J[B2_MAX_AXES] = {bi + i, bj + j};
          if (phase == 0) {
            if (use_bls) c.bls_add<float>(0, J, wx[i] * wy[j]); else atomicAdd(c.ref<float>(1, J), wx[i] * wy[j]);
          } else {
            acc += wx[i] * wy[j] * (use_bls ? c.bls_read<float>(0, J) : c.read<float>(1, J));
          }
        }
      if (phase == 1) c.arg<float *>(5)[p] = acc;
    }""", "particle_grid")
    base = [dx, np.int32(N), np.int32(bs)]
    k.range(0, npart, fields=[pid], args=base)
    assert int(ti.length(plist, ti.active_indices(block)[:, :2]).sum()) == npart
    for use_bls, m, row in ((1, m1, 0), (0, m2, 1)):
        a = base + [np.int32(use_bls), np.int32(0), out[row]]
        pads = [ti.bls(1, "add", (0, 0), (2, 2), extent=(bs, bs), halo=halo)] if use_bls else []
        k(plist, fields=[pid, m], args=a, block_local=pads)                     # P2G
        a[4] = np.int32(1)
        pads = [ti.bls(1, "read", (0, 0), (2, 2), extent=(bs, bs))] if use_bls else []
        k(plist, fields=[pid, m], args=a, block_local=pads)                     # G2P
    g1, g2 = m1.to_numpy(), m2.to_numpy()
    assert np.allclose(g1, g2, rtol=0, atol=1e-4)                               # bls_test_template.py:258-272 (abs 1e-4)
    assert np.allclose(out[0].cpu().numpy(), out[1].cpu().numpy(), rtol=0, atol=1e-4)
    X = xp.astype(np.float32) * np.float32(N)
    b = np.floor(X - np.float32(0.5)).astype(np.int32)
    f = X - b
    w = np.stack([0.5 * (1.5 - f) ** 2, 0.75 - (f - 1.0) ** 2, 0.5 * (f - 0.5) ** 2], 0).astype(np.float64)
    want = np.zeros((N, N))
    for i in range(3):
        for j in range(3):
            np.add.at(want, (b[:, 0] + i, b[:, 1] + j), w[i, :, 0] * w[j, :, 1])
    assert np.allclose(g1, want, rtol=1e-5, atol=1e-4)
    assert abs(g1.sum() - npart) < 1e-3 * npart                                 # partition of unity
    # activation: plain scatter opens exactly the touched cells' blocks; the pad with halo opens every block it overlaps
    act1 = set(map(tuple, ti.active_indices(block)[:, :2]))
    occupied = set(map(tuple, np.unique(b // bs, axis=0)))
    assert occupied <= act1


@pytest.mark.gpu
def test_thread_local_reductions(ti):
    """make_thread_local.cpp:100-213 / runtime.cpp:1235-1266: 0-D reductions from a struct-for body go through a thread-local
    accumulator and ONE atomic per CTA (c.reduce_add / reduce_min / reduce_max) instead of one atomic per cell."""
    import torch
    x = ti.field(ti.f32)
    blk = ti.root.pointer(ti.ij, 32).bitmasked(ti.ij, 8)
    blk.place(x)
    rng = np.random.default_rng(3)
    pts = rng.integers(0, 256, (5000, 2))
    pts = np.unique(pts, axis=0)
    vals = (rng.random(len(pts)) * 10 - 3).astype(np.float32)
    for (i, j), v in zip(pts[:400], vals[:400]):
        x[int(i), int(j)] = float(v)
    k = ti.struct_for("""B2_BODY(c) {
        const float v = c.read<float>(0, c.I);
        c.reduce_add<float>(0, v);
        c.reduce_add<int>(1, 1);
        c.reduce_max<float>(2, v);
        c.reduce_min<float>(3, v);
    }""", "tls")
    s = torch.zeros(1, device="cuda")
    n = torch.zeros(1, dtype=torch.int32, device="cuda")
    mx = torch.full((1,), -1e30, device="cuda")
    mn = torch.full((1,), 1e30, device="cuda")
    l0 = 0
    k(blk, fields=[x], args=[s, n, mx, mn])
    v = vals[:400].astype(np.float64)
    assert int(n.item()) == 400
    assert abs(float(s.item()) - v.sum()) < 1e-3
    assert float(mx.item()) == float(vals[:400].max()) and float(mn.item()) == float(vals[:400].min())


def _np_xorshift_streams(n_threads, n_draws, seed):
    """NumPy restatement of initialize_rand_state + rand_u32 (runtime.cpp:535-541,1790-1807): draws[k, i] = k-th value of state i."""
    i = (np.uint32(seed) * np.uint32(1048391) + np.arange(n_threads, dtype=np.uint32)).astype(np.uint32)
    with np.errstate(over="ignore"):
        x = (np.uint32(123456789) * i * np.uint32(1000000007)).astype(np.uint32)
        y = np.full(n_threads, 362436069, np.uint32)
        z = np.full(n_threads, 521288629, np.uint32)
        w = np.full(n_threads, 88675123, np.uint32)
        out = np.zeros((n_draws, n_threads), np.uint32)
        for k in range(n_draws):
            t = x ^ (x << np.uint32(11))
            x, y, z = y, z, w
            w = (w ^ (w >> np.uint32(19))) ^ (t ^ (t >> np.uint32(8)))
            out[k] = w * np.uint32(1000000007)
    return out


@pytest.mark.gpu
def test_rand_streams_of_a_range_for():
    """ti.random in a run-time body = the reference's xorshift128, one persistent state per linear thread id
    seeded from ti.init(random_seed=) (runtime.cpp:527-541,973-977,1790-1827).  A range-for of n <= grid x block indices maps
    index i to thread i, so the first launch must return draw 0 of every stream and the second launch draw 1."""
    if not have_gpu():
        pytest.skip("no CUDA device")
    import torch
    import taichi_b200 as ti
    seed, n = 7, 20000
    ti.init(arch=ti.cuda, device_memory_GB=0.25, random_seed=seed)
    try:
        f = ti.field(ti.i32)
        ti.root.dense(ti.i, 8).place(f)                       # (a tree must exist; the body only uses launch arguments)
        k = ti.struct_for("""
B2_RANGE_BODY(c) {
  unsigned *u = c.arg<unsigned *>(0); float *r = c.arg<float *>(1); double *d = c.arg<double *>(2);
  const int what = c.arg<int>(3);
  if (what == 0) u[c.i] = c.rand_u32();
  else if (what == 1) r[c.i] = c.rand_f32();
  else d[c.i] = c.rand_f64();
}
""", "rand_streams")
        u = torch.zeros(n, dtype=torch.int32, device="cuda")
        r = torch.zeros(n, dtype=torch.float32, device="cuda")
        d = torch.zeros(n, dtype=torch.float64, device="cuda")
        want = _np_xorshift_streams(n, 5, seed)
        for draw in range(2):
            k.range(0, n, fields=[f], args=[u, r, d, np.int32(0)])
            assert np.array_equal(u.cpu().numpy().view(np.uint32), want[draw]), draw
        k.range(0, n, fields=[f], args=[u, r, d, np.int32(1)])
        assert np.array_equal(r.cpu().numpy(), ((want[2] >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)))
        k.range(0, n, fields=[f], args=[u, r, d, np.int32(2)])
        u64 = (want[3].astype(np.uint64) << np.uint64(32)) + want[4].astype(np.uint64)
        assert np.array_equal(d.cpu().numpy(), (u64 >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0))
        assert 0.45 < float(r.mean()) < 0.55 and float(r.min()) >= 0.0 and float(r.max()) < 1.0
    finally:
        ti.reset()
