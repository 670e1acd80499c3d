"""Pins the oracle: oracle/snode_oracle.c (hand restatement) against
  (1) the reference's own known answers for this path, and
  (2) the reference's own runtime.cpp compiled for x64 (oracle/_ref/libref_runtime.so), on
      identical inputs: lists / struct-for coordinate sets / allocator counts must be identical.
Reference tests restated here (file:line under /root/reference):
  tests/python/test_bitmasked.py:5-137, tests/python/test_sparse_deactivate.py:5-107,
  tests/python/test_listgen.py:8-58, runtime_module/internal_functions.h:65-155,
  tests/cpp/codegen/refine_coordinates_test.cpp:149-165.
CPU only (no GPU, no /root/reference needed at run time: _ref is prebuilt)."""
import ctypes as C
import numpy as np
import pytest

from _desc import TreeBuilder, coords12, i32p, load_ref_runtime, load_oracle

ORC = load_oracle()
REF = load_ref_runtime()


class Impl:
    def __init__(self, lib, prefix):
        self.lib, self.p = lib, prefix

    def __getattr__(self, name):
        return getattr(self.lib, self.p + name)

    def create(self, tb):
        assert getattr(self.lib, self.p + "create")(tb.array(), len(tb)) == 0

    def activate(self, target, coords):
        c = coords12(coords)
        getattr(self.lib, self.p + "activate")(target, i32p(c), len(c))

    def deactivate(self, target, coords):
        c = coords12(coords)
        getattr(self.lib, self.p + "deactivate")(target, i32p(c), len(c))

    def write_i32(self, place, coords, vals):
        c = coords12(coords)
        v = np.asarray(vals, dtype=np.int32)
        getattr(self.lib, self.p + "write")(place, i32p(c), len(c), v.ctypes.data_as(C.c_void_p), 0)

    def read_i32(self, place, coords):
        c = coords12(coords)
        v = np.zeros(len(c), dtype=np.int32)
        getattr(self.lib, self.p + "read")(place, i32p(c), len(c), v.ctypes.data_as(C.c_void_p))
        return v

    def is_active(self, target, coords):
        c = coords12(coords)
        out = np.zeros(len(c), dtype=np.uint8)
        getattr(self.lib, self.p + "is_active")(target, i32p(c), len(c), out.ctypes.data_as(C.POINTER(C.c_uint8)))
        return out

    def list(self, target):
        n = getattr(self.lib, self.p + "listgen")(target)
        out = np.zeros((max(n, 1), 14), dtype=np.int32)
        getattr(self.lib, self.p + "get_list")(target, i32p(out), n)
        out = out[:n]
        return out[np.lexsort(out.T[::-1])] if n else out

    def visited(self, target):
        n = getattr(self.lib, self.p + "struct_for_coords")(target, None, 0)
        out = np.zeros((max(n, 1), 12), dtype=np.int32)
        getattr(self.lib, self.p + "struct_for_coords")(target, i32p(out), n)
        out = out[:n]
        return out[np.lexsort(out.T[::-1])] if n else out

    def stats(self, snode):
        st = np.zeros(4, dtype=np.int32)
        assert getattr(self.lib, self.p + "alloc_stats")(snode, i32p(st)) == 0
        return st

    def deactivate_all(self, snode):
        getattr(self.lib, self.p + "deactivate_all")(snode)

    def gc(self, snode):
        getattr(self.lib, self.p + "gc")(snode)


IMPLS = [Impl(ORC, "orc_")] + ([Impl(REF, "refrt_")] if REF is not None else [])
ids = ["oracle"] + (["reference"] if REF is not None else [])


def test_reference_runtime_is_present():
    # the prebuilt reference runtime must travel with the repo (oracle/_ref/ is not gpurun-ignored)
    assert REF is not None, "oracle/_ref/libref_runtime.so missing: run `make -C oracle ref`"


@pytest.mark.skipif(REF is None, reason="reference runtime not built")
def test_reference_internal_selftests():
    # internal_functions.h:65-155 run unchanged inside the reference runtime
    tb = TreeBuilder()
    tb.place(tb.dense(0, "i", 4))
    assert REF.refrt_create(tb.array(), len(tb)) == 0
    assert REF.refrt_selftest_list_manager() == 0
    assert REF.refrt_selftest_node_allocator() == 0
    assert REF.refrt_selftest_node_allocator_gc() == 0


@pytest.mark.parametrize("impl", IMPLS, ids=ids)
def test_bitmasked_basic(impl):  # test_bitmasked.py:5-31 -> c == 3, s == 42
    tb = TreeBuilder()
    b0 = tb.bitmasked(0, "ij", (3, 6))
    bm = tb.bitmasked(b0, "i", 8)
    x = tb.place(bm)
    impl.create(tb)
    impl.write_i32(x, [[5, 1], [9, 4], [0, 3]], [2, 20, 20])
    v = impl.visited(bm)
    assert len(v) == 3
    assert impl.read_i32(x, v[:, :2]).sum() == 42
    assert impl.is_active(bm, v[:, :2]).sum() == 3


@pytest.mark.parametrize("impl", IMPLS, ids=ids)
def test_bitmasked_then_dense_and_bitmasked(impl):  # test_bitmasked.py:34-79 -> 256 and 4
    for inner, expect in (("dense", 256), ("bitmasked", 4)):
        tb = TreeBuilder()
        b = tb.bitmasked(0, "i", 128)
        d = getattr(tb, inner)(b, "i", 128)
        x = tb.place(d)
        impl.create(tb)
        impl.write_i32(x, [[0], [127], [256], [257]], [1, 1, 1, 1])
        assert len(impl.visited(d)) == expect


@pytest.mark.parametrize("impl", IMPLS, ids=ids)
def test_huge_bitmasked(impl):  # test_bitmasked.py:82-106 (n scaled 1024 -> 256 for CPU time)
    n = 256
    tb = TreeBuilder()
    b = tb.bitmasked(0, "i", n)
    b2 = tb.bitmasked(b, "i", 2 * n)
    x = tb.place(b2)
    impl.create(tb)
    idx = np.arange(0, n * n * 2, 32)
    impl.activate(b2, idx[:, None])
    assert len(impl.visited(b2)) == (n * n * 2) // 32


@pytest.mark.parametrize("impl", IMPLS, ids=ids)
def test_bitmasked_listgen_prime(impl):  # test_bitmasked.py:109-137, n = 80173 (prime)
    n = 80173
    tb = TreeBuilder()
    d = tb.dense(0, "i", n)
    b = tb.bitmasked(d, "i", 1)
    tb.place(b)
    impl.create(tb)
    impl.activate(b, np.arange(n)[:, None])
    assert len(impl.visited(b)) == n
    # root listgen splits the dense container into ceil(n/1024) Elements (runtime.cpp:1313-1328)
    impl.lib  # noqa
    lst = impl.list(d)
    assert len(lst) == (n + 1023) // 1024 and lst[:, 13].max() == n


@pytest.mark.parametrize("impl", IMPLS, ids=ids)
def test_sparse_deactivate_pointer(impl):  # test_sparse_deactivate.py:5-33 and :36-70
    tb = TreeBuilder()
    ptr = tb.pointer(0, "i", 16)
    d = tb.dense(ptr, "i", 16)
    x = tb.place(d)
    impl.create(tb)
    impl.write_i32(x, [[0], [19]], [1, 1])
    assert len(impl.visited(d)) == 32
    impl.deactivate(ptr, [[0]])
    impl.gc(ptr)
    assert len(impl.visited(d)) == 16
    impl.write_i32(x, [[20], [45]], [1, 1])
    assert len(impl.visited(d)) == 32
    # re-activated cell 0 must read zero (recycled memory is zero-filled at GC)
    impl.activate(d, [[3]])
    assert impl.read_i32(x, [[0]])[0] == 0
    assert impl.stats(ptr)[0] == 3  # 3 nodes ever carved: cells {0,1}, then 2, then reuse of 0's node


@pytest.mark.parametrize("impl", IMPLS, ids=ids)
def test_gc_reuses_nodes(impl):  # test_gc.py:75-90: pointer -> _num_dynamically_allocated == 1
    tb = TreeBuilder()
    ptr = tb.pointer(0, "i", 16)
    d = tb.dense(ptr, "i", 16)
    x = tb.place(d)
    impl.create(tb)
    for it in range(20):
        impl.deactivate_all(ptr)
        impl.write_i32(x, [[(it * 7) % 256]], [1])
        assert impl.stats(ptr)[0] == 1
    assert len(impl.visited(d)) == 16


@pytest.mark.parametrize("impl", IMPLS, ids=ids)
def test_listgen_coverage_3d(impl):  # test_listgen.py:35-58 shape: 4 dense levels of 4 along ijk
    tb = TreeBuilder()
    p = 0
    for _ in range(3):
        p = tb.dense(p, "ijk", 4)
    x = tb.place(p)
    impl.create(tb)
    v = impl.visited(p)
    assert len(v) == 64**3
    lin = (v[:, 0].astype(np.int64) * 64 + v[:, 1]) * 64 + v[:, 2]
    assert np.array_equal(np.sort(lin), np.arange(64**3))  # every cell exactly once, correct coords


def test_refine_coordinates_known_answer():  # refine_coordinates_test.cpp:149-165
    tb = TreeBuilder()
    d = tb.dense(0, "ij", (4, 8))  # loop_index % 8 == j, / 8 == i
    tb.place(d)
    o = IMPLS[0]
    o.create(tb)
    v = o.visited(d)
    assert np.array_equal(v[:, 0] * 8 + v[:, 1], np.arange(32))


def _random_tree(rng, ndim):
    axes = "ijk"[:ndim]
    tb = TreeBuilder()
    kinds = [rng.choice(["pointer", "bitmasked", "dense"]) for _ in range(rng.integers(2, 4))]
    if all(k == "dense" for k in kinds):
        kinds[0] = "pointer"
    p = 0
    for k in kinds:
        sizes = [int(rng.choice([2, 3, 4, 8])) for _ in axes]
        p = getattr(tb, k)(p, axes, sizes)
    places = [tb.place(p, 4) for _ in range(rng.integers(1, 4))]
    extent = [1] * ndim
    q = p
    while q > 0:
        for a, s in zip(tb.nodes[q]["axes"], tb.nodes[q]["sizes"]):
            extent[a] *= s
        q = tb.nodes[q]["parent"]
    return tb, p, places, extent


@pytest.mark.skipif(REF is None, reason="reference runtime not built")
@pytest.mark.parametrize("seed", range(12))
def test_oracle_matches_reference_random_trees(seed):
    rng = np.random.default_rng(seed)
    ndim = int(rng.integers(1, 4))
    tb, leaf, places, extent = _random_tree(rng, ndim)
    o, r = IMPLS
    o.create(tb)
    r.create(tb)
    for lvl in range(len(tb)):
        assert ORC.orc_node_size(lvl) == REF.refrt_node_size(lvl)
    n = int(rng.integers(1, 200))
    coords = np.stack([rng.integers(0, e, n) for e in extent], axis=1)
    for impl in (o, r):
        impl.activate(leaf, coords)
    path = []
    q = leaf
    while q > 0:
        path.append(q)
        q = tb.nodes[q]["parent"]
    for lvl in path:
        assert np.array_equal(o.list(lvl), r.list(lvl))
        assert np.array_equal(o.visited(lvl), r.visited(lvl))
    # deactivate a random subset at a random sparse level, gc, re-activate others
    sparse = [q for q in path if tb.nodes[q]["type"] in (2, 3)]
    lvl = int(rng.choice(sparse))
    vis = o.visited(lvl)
    if len(vis):
        sub = vis[rng.random(len(vis)) < 0.5][:, :ndim]
        if len(sub):
            for impl in (o, r):
                impl.deactivate(lvl, sub)
                if tb.nodes[lvl]["type"] == 2:
                    impl.gc(lvl)
    coords2 = np.stack([rng.integers(0, e, n // 2 + 1) for e in extent], axis=1)
    for impl in (o, r):
        impl.activate(leaf, coords2)
    for q in path:
        assert np.array_equal(o.list(q), r.list(q))
        if tb.nodes[q]["type"] == 2:
            assert np.array_equal(o.stats(q), r.stats(q))
    top = path[-1]
    for impl in (o, r):
        impl.deactivate_all(top)
    for q in path:
        assert len(o.visited(q)) == len(r.visited(q))
        if tb.nodes[q]["type"] == 2:
            assert np.array_equal(o.stats(q), r.stats(q))
