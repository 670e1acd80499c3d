"""Shared helpers for tests: SNode-tree descriptors (one byte layout for the
reference harness, the C oracle and the CUDA library) and ctypes loaders."""
import ctypes as C
import os
import numpy as np

ROOT, DENSE, POINTER, BITMASKED, DYNAMIC, PLACE = range(6)
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class SNodeDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("parent", C.c_int32), ("num_axes", C.c_int32),
                ("axes", C.c_int32 * 12), ("sizes", C.c_int32 * 12),
                ("dtype_bytes", C.c_int32), ("chunk_size", C.c_int32)]


class TreeBuilder:
    """tb = TreeBuilder(); p = tb.pointer(0, 'ij', 4); d = tb.dense(p, 'ij', 8); f = tb.place(d, 4)"""

    def __init__(self):
        self.nodes = [dict(type=ROOT, parent=-1, axes=[], sizes=[], dtype_bytes=0, chunk_size=0)]

    def _add(self, type_, parent, axes, sizes, dtype_bytes=0, chunk_size=0):
        if isinstance(axes, str):
            axes = ["ijklmnop".index(c) for c in axes]
        if isinstance(sizes, int):
            sizes = [sizes] * len(axes)
        self.nodes.append(dict(type=type_, parent=parent, axes=list(axes), sizes=list(sizes),
                               dtype_bytes=dtype_bytes, chunk_size=chunk_size))
        return len(self.nodes) - 1

    def dense(self, parent, axes, sizes):
        return self._add(DENSE, parent, axes, sizes)

    def pointer(self, parent, axes, sizes):
        return self._add(POINTER, parent, axes, sizes)

    def bitmasked(self, parent, axes, sizes):
        return self._add(BITMASKED, parent, axes, sizes)

    def dynamic(self, parent, axis, n, chunk_size):
        return self._add(DYNAMIC, parent, axis, [n], chunk_size=chunk_size)

    def place(self, parent, dtype_bytes=4):
        return self._add(PLACE, parent, [], [], dtype_bytes=dtype_bytes)

    def array(self):
        arr = (SNodeDesc * len(self.nodes))()
        for i, n in enumerate(self.nodes):
            arr[i].type, arr[i].parent = n["type"], n["parent"]
            arr[i].num_axes = len(n["axes"])
            for k, (a, s) in enumerate(zip(n["axes"], n["sizes"])):
                arr[i].axes[k], arr[i].sizes[k] = a, s
            arr[i].dtype_bytes, arr[i].chunk_size = n["dtype_bytes"], n["chunk_size"]
        return arr

    def __len__(self):
        return len(self.nodes)


def coords12(c):
    """(n, ndim) int array -> contiguous (n, 12) int32"""
    c = np.asarray(c, dtype=np.int32)
    if c.ndim == 1:
        c = c[:, None]
    out = np.zeros((c.shape[0], 12), dtype=np.int32)
    out[:, : c.shape[1]] = c
    return np.ascontiguousarray(out)


def i32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def load_ref_runtime():
    path = os.path.join(REPO, "oracle", "_ref", "libref_runtime.so")
    if not os.path.exists(path):
        return None
    lib = C.CDLL(path)
    for f in ("refrt_node_size", "refrt_cell_size", "refrt_num_cells"):
        getattr(lib, f).restype = C.c_int64
    return lib


def load_oracle():
    path = os.path.join(REPO, "oracle", "liboracle.so")
    lib = C.CDLL(path)
    for f in ("orc_node_size", "orc_cell_size", "orc_num_cells"):
        getattr(lib, f).restype = C.c_int64
    return lib
