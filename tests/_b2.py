"""GPU-side test helpers: the same `Impl` surface as tests/test_oracle_vs_reference.py, but
calling the product through its C ABI (ctypes -> libb200mpm.so)."""
import ctypes as C
import numpy as np

from _desc import coords12


def have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


_RT = None


def runtime():
    """One runtime per test process (the reference's runtime is a per-process singleton too)."""
    global _RT
    from taichi_b200 import _lib as L
    if _RT is None:
        L.lib.b2_set_pool_size(C.c_uint64(8 << 30))  # the C4 full-size case needs 32768 x 131 KB containers
        _RT = L.lib.ti_create_runtime(L.TI_ARCH_CUDA, 0)
        L.check()
        assert _RT
    return _RT


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


class B2Impl:
    """Same methods as test_oracle_vs_reference.Impl; every call crosses the C ABI."""

    def __init__(self):
        from taichi_b200 import _lib as L
        self.L = L
        self.lib = L.lib
        self.tree = None

    def create(self, tb):
        if self.tree:
            self.lib.b2_destroy_tree(self.tree)
        arr = tb.array()
        self.tree = self.lib.b2_create_tree(runtime(), C.cast(arr, C.POINTER(self.L.B2SNodeDesc)), len(tb))
        self.L.check()
        assert self.tree

    def close(self):
        if self.tree:
            self.lib.b2_destroy_tree(self.tree)
            self.tree = None

    def _ok(self, rc):
        self.L.check()
        assert rc == 0

    def activate(self, target, coords):
        c = coords12(coords)
        self._ok(self.lib.b2_activate(self.tree, target, vp(c), len(c)))

    def deactivate(self, target, coords):
        c = coords12(coords)
        self._ok(self.lib.b2_deactivate(self.tree, target, vp(c), len(c)))

    def write_i32(self, place, coords, vals):
        c = coords12(coords)
        v = np.ascontiguousarray(vals, dtype=np.int32)
        self._ok(self.lib.b2_write(self.tree, place, vp(c), len(c), vp(v), 0))

    def add_f32(self, place, coords, vals):
        c = coords12(coords)
        v = np.ascontiguousarray(vals, dtype=np.float32)
        self._ok(self.lib.b2_write(self.tree, place, vp(c), len(c), vp(v), 1))

    def read_i32(self, place, coords):
        c = coords12(coords)
        v = np.zeros(len(c), dtype=np.int32)
        self._ok(self.lib.b2_read(self.tree, place, vp(c), len(c), vp(v)))
        return v

    def read_f32(self, place, coords):
        return self.read_i32(place, coords).view(np.float32)

    def is_active(self, target, coords):
        c = coords12(coords)
        out = np.zeros(len(c), dtype=np.uint8)
        self._ok(self.lib.b2_is_active(self.tree, target, vp(c), len(c), vp(out)))
        return out

    def list(self, target):
        n = self.lib.b2_listgen(self.tree, target)
        self.L.check()
        out = np.zeros((max(n, 1), 14), dtype=np.int32)
        self.lib.b2_get_list(self.tree, target, vp(out), n)
        self.L.check()
        out = out[:n]
        return out[np.lexsort(out.T[::-1])] if n else out

    def visited(self, target):
        n = self.lib.b2_struct_for_coords(self.tree, target, None, 0)
        self.L.check()
        out = np.zeros((max(n, 1), 12), dtype=np.int32)
        n2 = self.lib.b2_struct_for_coords(self.tree, target, vp(out), n)
        self.L.check()
        assert n2 == n
        out = out[:n]
        return out[np.lexsort(out.T[::-1])] if n else out

    def stats(self, snode):
        st = np.zeros(4, dtype=np.int32)
        self._ok(self.lib.b2_alloc_stats(self.tree, snode, st.ctypes.data_as(self.L.i32p)))
        return st

    def deactivate_all(self, snode):
        self._ok(self.lib.b2_deactivate_all(self.tree, snode))

    def gc(self, snode):
        self._ok(self.lib.b2_gc(self.tree, snode))

    def append(self, dyn, coords, vals):
        c = coords12(coords)
        v = np.ascontiguousarray(vals, dtype=np.int32)
        self._ok(self.lib.b2_dynamic_append_i32(self.tree, dyn, vp(c), len(c), vp(v)))

    def length(self, dyn, coords):
        c = coords12(coords)
        out = np.zeros(len(c), dtype=np.int32)
        self._ok(self.lib.b2_dynamic_length(self.tree, dyn, vp(c), len(c), vp(out)))
        return out

    def runtime_error(self):
        buf = C.create_string_buffer(256)
        code = self.lib.b2_check_runtime_error(runtime(), buf, 256)
        return code, buf.value.decode()
