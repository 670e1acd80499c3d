"""Taichi-like Python surface for the sparse-SNode path (names + argument meaning of
python/taichi/lang/snode.py:26-304, lang/misc.py:323, lang/impl.py:592-625, lang/field.py:121-351).

    import taichi_b200 as ti
    ti.init(arch=ti.cuda, device_memory_GB=2)
    x = ti.field(ti.f32)
    block = ti.root.pointer(ti.ijk, 16).bitmasked(ti.ijk, 8).dense(ti.ijk, 4)
    block.place(x)
    x[3, 4, 5] = 1.0                  # activates, like the reference's python-scope write
    for I in ti.active_indices(block):  # struct-for over active cells (host view of the kernel's loop)
        ...
    block.parent().deactivate_all()

There is no kernel compiler here: struct-fors run as the library's built-in bodies
(fill / reduce / to_numpy / deactivate / coordinate dump) and as the prebuilt MPM kernels.
"""
import ctypes as C
import numpy as np

from . import _lib as L

lib = L.lib

cuda = gpu = "cuda"
cpu = x64 = "x64"
f32, i32, f64, i64 = "f32", "i32", "f64", "i64"
_DT = {"f32": (np.float32, 4), "i32": (np.int32, 4), "f64": (np.float64, 8), "i64": (np.int64, 8),
       float: (np.float32, 4), int: (np.int32, 4)}


class _Axes(tuple):
    pass


i, j, k, l = _Axes((0,)), _Axes((1,)), _Axes((2,)), _Axes((3,))
ij, ijk, ijkl = _Axes((0, 1)), _Axes((0, 1, 2)), _Axes((0, 1, 2, 3))
jk, ik = _Axes((1, 2)), _Axes((0, 2))


class _Program:
    def __init__(self):
        self.rt = None
        self.nodes = None
        self.tree = None
        self.fields = []
        self.kernels = []


_prog = _Program()
root = None


# ti.init options that have no effect on this path and are accepted silently (lang/misc.py:323); anything else is an error
_INIT_ACCEPTED = {"default_fp": "f32", "default_ip": "i32", "random_seed": None, "kernel_profiler": None, "debug": None, "offline_cache": None,
                  "log_level": None, "verbose": None, "fast_math": None, "cpu_max_num_threads": None, "packed": None}


def init(arch=cuda, device_memory_GB=None, device_memory_fraction=None, device_index=0, **options):
    """ti.init: only arch=ti.cuda exists in this library (no CPU fallback).  Options this path cannot honour raise."""
    global root
    if arch not in (cuda,):
        raise RuntimeError("taichi_b200 implements arch=ti.cuda only (no ti.cpu fallback)")
    for k, v in options.items():
        if k not in _INIT_ACCEPTED:
            raise TypeError(f"ti.init(): option {k!r} is not supported by taichi_b200")
        if _INIT_ACCEPTED[k] is not None and v not in (_INIT_ACCEPTED[k], globals().get(_INIT_ACCEPTED[k])):
            raise ValueError(f"ti.init({k}={v!r}): only {_INIT_ACCEPTED[k]} is implemented")
    if device_memory_GB is None and device_memory_fraction is not None:
        import torch
        device_memory_GB = float(device_memory_fraction) * torch.cuda.get_device_properties(device_index).total_memory / (1 << 30)
    reset()
    if device_memory_GB is not None:
        lib.b2_set_pool_size(C.c_uint64(int(device_memory_GB * (1 << 30))))
    lib.b2_set_random_seed(int(options.get("random_seed") or 0))   # c.rand_* of run-time bodies (ti.random)
    _prog.rt = lib.ti_create_runtime(L.TI_ARCH_CUDA, device_index)
    L.check()
    _prog.nodes = [dict(type=L.ROOT, parent=-1, axes=[], sizes=[], dtype_bytes=0, chunk_size=0)]
    root = SNode(0)
    return _prog.rt


def reset():
    global root
    for k in list(_prog.kernels):
        k.destroy()
    _prog.kernels = []
    if _prog.tree:
        lib.b2_destroy_tree(_prog.tree)
    if _prog.rt:
        lib.ti_destroy_runtime(_prog.rt)
    _prog.rt = _prog.tree = _prog.nodes = None
    _prog.fields = []
    root = None


def sync():
    lib.b2_sync(_prog.rt)
    L.check()


def _materialize():
    """Program::add_snode_tree: first use of any field freezes the tree (program.cpp:238-255)."""
    if _prog.tree is None:
        n = len(_prog.nodes)
        arr = (L.B2SNodeDesc * n)()
        for q, nd in enumerate(_prog.nodes):
            arr[q].type, arr[q].parent, arr[q].num_axes = nd["type"], nd["parent"], len(nd["axes"])
            for t, (a, s) in enumerate(zip(nd["axes"], nd["sizes"])):
                arr[q].axes[t], arr[q].sizes[t] = a, s
            arr[q].dtype_bytes, arr[q].chunk_size = nd["dtype_bytes"], nd["chunk_size"]
        _prog.tree = lib.b2_create_tree(_prog.rt, arr, n)
        L.check()
    return _prog.tree


def _coords(idx, ndim=None):
    c = np.atleast_2d(np.asarray(idx, dtype=np.int32))
    out = np.zeros((c.shape[0], 12), np.int32)
    out[:, :c.shape[1]] = c
    return out


class SNode:
    def __init__(self, idx):
        self.id = idx

    def _child(self, type_, axes, sizes, chunk_size=0):
        if _prog.tree is not None:
            raise RuntimeError("the SNode tree is already materialized")
        axes = list(axes)
        sizes = [sizes] * len(axes) if isinstance(sizes, int) else list(sizes)
        _prog.nodes.append(dict(type=type_, parent=self.id, axes=axes, sizes=sizes, dtype_bytes=0, chunk_size=chunk_size))
        return SNode(len(_prog.nodes) - 1)

    def dense(self, axes, dimensions):
        return self._child(L.DENSE, axes, dimensions)

    def pointer(self, axes, dimensions):
        return self._child(L.POINTER, axes, dimensions)

    def bitmasked(self, axes, dimensions):
        return self._child(L.BITMASKED, axes, dimensions)

    def dynamic(self, axis, dimension, chunk_size=None):
        return self._child(L.DYNAMIC, axis, [dimension], chunk_size or dimension)

    def place(self, *fields):
        for f in fields:
            for comp in f._components():
                _prog.nodes.append(dict(type=L.PLACE, parent=self.id, axes=[], sizes=[], dtype_bytes=comp.itemsize, chunk_size=0))
                comp.place_id = len(_prog.nodes) - 1
                comp.parent_snode = self
        return self

    def parent(self, n=1):
        p = self
        for _ in range(n):
            p = SNode(_prog.nodes[p.id]["parent"])
        return p

    @property
    def shape(self):
        ext = {}
        q = self.id
        while q > 0:
            nd = _prog.nodes[q]
            for a, s in zip(nd["axes"], nd["sizes"]):
                ext[a] = ext.get(a, 1) * s
            q = nd["parent"]
        return tuple(ext[a] for a in sorted(ext))

    def _axes(self):
        ax = set()
        q = self.id
        while q > 0:
            ax.update(_prog.nodes[q]["axes"])
            q = _prog.nodes[q]["parent"]
        return sorted(ax)

    def deactivate_all(self):
        lib.b2_deactivate_all(_materialize(), self.id)
        L.check()

    @property
    def _num_dynamically_allocated(self):
        n = lib.b2_num_dynamically_allocated(_materialize(), self.id)
        L.check()
        return int(n)

    @property
    def _cell_size_bytes(self):
        return int(lib.b2_snode_cell_size(_materialize(), self.id))

    def _to_full(self, idx):
        c = np.atleast_2d(np.asarray(idx, dtype=np.int32))
        if not self._axes():                       # the root cell: ti.append(x.parent(), [], v)
            return np.zeros((max(c.shape[0], 1) if c.size else 1, 12), np.int32)
        out = np.zeros((c.shape[0], 12), np.int32)
        for t, a in enumerate(self._axes()):
            out[:, a] = c[:, t]
        return out


def activate(snode, idx):
    c = snode._to_full(idx)
    lib.b2_activate(_materialize(), snode.id, c.ctypes.data_as(C.c_void_p), len(c))
    L.check()


def deactivate(snode, idx):
    c = snode._to_full(idx)
    lib.b2_deactivate(_materialize(), snode.id, c.ctypes.data_as(C.c_void_p), len(c))
    L.check()


def is_active(snode, idx):
    c = snode._to_full(idx)
    out = np.zeros(len(c), np.uint8)
    lib.b2_is_active(_materialize(), snode.id, c.ctypes.data_as(C.c_void_p), len(c), out.ctypes.data_as(C.c_void_p))
    L.check()
    return out.astype(bool) if np.ndim(idx) > 1 else bool(out[0])


def append(snode, idx, val):
    c = snode.parent()._to_full(idx) if np.ndim(idx) else snode.parent()._to_full([idx])
    if not snode.parent()._axes() and np.ndim(val):  # root-level list: one append per value
        c = np.zeros((len(val), 12), np.int32)
    v = np.ascontiguousarray(np.broadcast_to(val, (len(c),)), dtype=np.int32)
    lib.b2_dynamic_append_i32(_materialize(), snode.id, c.ctypes.data_as(C.c_void_p), len(c), v.ctypes.data_as(C.c_void_p))
    L.check()


def length(snode, idx):
    c = snode.parent()._to_full(idx) if np.ndim(idx) else snode.parent()._to_full([idx])
    out = np.zeros(len(c), np.int32)
    lib.b2_dynamic_length(_materialize(), snode.id, c.ctypes.data_as(C.c_void_p), len(c), out.ctypes.data_as(C.c_void_p))
    L.check()
    return out if np.ndim(idx) > 1 else int(out[0])


def active_indices(snode):
    """Coordinates a struct-for over `snode` visits (n, ndim), sorted."""
    t = _materialize()
    n = lib.b2_struct_for_coords(t, snode.id, None, 0)
    L.check()
    out = np.zeros((max(n, 1), 12), np.int32)
    lib.b2_struct_for_coords(t, snode.id, out.ctypes.data_as(C.c_void_p), n)
    L.check()
    out = out[:n][:, snode._axes()]
    return out[np.lexsort(out.T[::-1])] if n else out


def memory_profiler_info():
    """ti.profiler.print_memory_profiler_info(): the report as a string."""
    t = _materialize()
    n = lib.b2_memory_profiler_info(t, None, 0)
    L.check()
    buf = C.create_string_buffer(int(n))
    lib.b2_memory_profiler_info(t, buf, n)
    L.check()
    return buf.value.decode()


def deactivate_all_snodes():
    for q, nd in enumerate(_prog.nodes):
        if nd["parent"] == 0 and nd["type"] != L.PLACE:
            SNode(q).deactivate_all()


class ScalarField:
    def __init__(self, dtype):
        self.np_dtype, self.itemsize = _DT[dtype]
        self.np_dtype = np.dtype(self.np_dtype)
        self.place_id = None
        self.parent_snode = None

    def _components(self):
        return [self]

    @property
    def shape(self):
        return self.parent_snode.shape

    def __setitem__(self, idx, val):
        c = self.parent_snode._to_full([idx] if np.ndim(idx) <= 1 else idx)
        v = np.ascontiguousarray(np.broadcast_to(val, (len(c),)), dtype=self.np_dtype)
        lib.b2_write(_materialize(), self.place_id, c.ctypes.data_as(C.c_void_p), len(c), v.ctypes.data_as(C.c_void_p), 0)
        L.check()

    def __getitem__(self, idx):
        c = self.parent_snode._to_full([idx] if np.ndim(idx) <= 1 else idx)
        v = np.zeros(len(c), self.np_dtype)
        lib.b2_read(_materialize(), self.place_id, c.ctypes.data_as(C.c_void_p), len(c), v.ctypes.data_as(C.c_void_p))
        L.check()
        return v if np.ndim(idx) > 1 else v[0]

    def atomic_add(self, idx, val):
        """ti.atomic_add(x[I], v) for a batch of indices (f32)."""
        c = self.parent_snode._to_full(idx)
        v = np.ascontiguousarray(np.broadcast_to(val, (len(c),)), dtype=np.float32)
        lib.b2_write(_materialize(), self.place_id, c.ctypes.data_as(C.c_void_p), len(c), v.ctypes.data_as(C.c_void_p), 1)
        L.check()

    def fill(self, val):
        raw = np.array([val], self.np_dtype).tobytes()                     # the value's bit pattern, any width up to 8 bytes
        bits = int.from_bytes(raw, "little")
        lib.b2_field_fill(_materialize(), self.place_id, C.c_uint64(bits))
        L.check()

    def to_numpy(self):
        out = np.zeros(self.shape, self.np_dtype)
        lib.b2_field_to_dense(_materialize(), self.place_id, out.ctypes.data_as(C.c_void_p))
        L.check()
        return out

    def from_numpy(self, arr):
        arr = np.ascontiguousarray(arr, dtype=self.np_dtype)
        assert arr.shape == self.shape
        lib.b2_field_from_dense(_materialize(), self.place_id, arr.ctypes.data_as(C.c_void_p))
        L.check()

    def to_sparse(self):
        """Checkpoint form of a sparse field: (coords [n, ndim] of the ACTIVE cells in sorted order, values [n]).
        Unlike to_numpy (dense image, `_kernels.py:40-60`) it keeps the active set, so `from_sparse` restores the
        field exactly -- same active cells, same values, nothing else activated."""
        idx = active_indices(self.parent_snode)
        vals = self[idx] if len(idx) else np.zeros(0, self.np_dtype)
        return idx, np.asarray(vals, self.np_dtype)

    def from_sparse(self, coords, values):
        coords = np.asarray(coords, np.int32)
        if len(coords):
            self[coords] = np.asarray(values, self.np_dtype)

    def to_torch(self, device="cuda"):
        import torch
        out = torch.zeros(self.shape, dtype=getattr(torch, self.np_dtype.name), device=device)
        lib.b2_field_to_dense(_materialize(), self.place_id, C.c_void_p(out.data_ptr()))
        L.check()
        sync()
        return out

    def from_torch(self, t):
        assert t.is_contiguous() and tuple(t.shape) == self.shape
        lib.b2_field_from_dense(_materialize(), self.place_id, C.c_void_p(t.data_ptr()))
        L.check()
        sync()

    def _reduce(self, op):
        out = np.zeros(1, self.np_dtype)
        lib.b2_field_reduce(_materialize(), self.place_id, op, 0 if self.np_dtype == np.float32 else 1, out.ctypes.data_as(C.c_void_p))
        L.check()
        return out[0]

    def sum(self):
        return self._reduce(0)

    def min(self):
        return self._reduce(1)

    def max(self):
        return self._reduce(2)


def field(dtype, shape=None):
    f = ScalarField(dtype)
    if shape is not None:
        shape = (shape,) if isinstance(shape, int) else tuple(shape)
        root.dense(_Axes(range(len(shape))), list(shape)).place(f)
    _prog.fields.append(f)
    return f


def block_local(*_fields):
    """ti.block_local(field) is a hint INSIDE a Python-compiled kernel in the reference (misc.py:480-496).  Here loop bodies are
    CUDA C++ given to ti.struct_for, and the pad is declared at the launch: kernel(snode, fields=[..], block_local=[ti.bls(..)])."""
    raise NotImplementedError("declare pads at the launch: kernel(snode, fields=[...], block_local=[ti.bls(field_index, 'read'|'add', lo, hi)]); "
                              "the prebuilt MPM solver takes Mpm3D(block_local=True)")


def no_activate(*_snodes):
    """ti.no_activate (misc.py:473): reads through c.read never activate; writes always do -- there is no per-SNode switch."""
    raise NotImplementedError("ti.no_activate is not available: c.read<T>() never activates, c.write / c.ref always do")


def loop_config(**_kw):
    """ti.loop_config(block_dim=..) (misc.py:612): launch shapes are chosen by the library."""
    raise NotImplementedError("ti.loop_config is not available: launch shapes are chosen by the library")


def bls(field, mode, lo, hi, axes=None, extent=None, halo=False):
    """One block-local pad for a struct-for launch: `field` = index into the launch's field list; mode 'read' (stencil gather)
    or 'add' (scatter with c.bls_add); lo / hi = smallest / largest offset the body reaches per pad axis (inclusive);
    axes = physical axes of the pad (default 0..len(lo)-1); extent = block size per axis in the field's index space
    (default: the loop SNode's own shape -- pass it for pid-list loops over a dynamic SNode); halo=True activates every pad
    cell in the epilogue like the reference's CUDA backend, False only the cells that received something (its x64 result)."""
    lo, hi = list(np.atleast_1d(lo)), list(np.atleast_1d(hi))
    axes = list(range(len(lo))) if axes is None else list(axes)
    extent = [0] * len(lo) if extent is None else list(np.atleast_1d(extent))
    return dict(field=int(field), mode=mode, lo=lo, hi=hi, axes=axes, extent=extent, halo=bool(halo))


# ---------------------------------------------------------------------------------------------
# struct-for with a run-time compiled body (the part of @ti.kernel this library offers):
# codegen_llvm.cpp:2089-2317 + jit_cuda.cpp in the reference, NVRTC -> sm_100a cubin here.
# ---------------------------------------------------------------------------------------------
def _arg_bits(a):
    """One 64-bit launch slot: device pointers (torch CUDA tensors / objects with data_ptr), numpy scalars by their
    dtype, python int -> i64, python float -> f32 (default_fp, compile_config.cpp:30)."""
    if hasattr(a, "data_ptr"):
        if hasattr(a, "is_cuda") and not a.is_cuda:
            raise TypeError("struct-for arguments are device pointers or scalars (got a CPU tensor)")
        return int(a.data_ptr())
    if isinstance(a, (bool, int, np.integer)) and not isinstance(a, (np.int32, np.uint32)):
        return int(np.array(a, np.int64).view(np.uint64))
    if isinstance(a, (np.int32, np.uint32)):
        return int(np.array([a, 0], np.int32).view(np.uint64)[0])
    if isinstance(a, np.float64):
        return int(np.array(a, np.float64).view(np.uint64))
    if isinstance(a, (float, np.floating)):
        return int(np.array([a, 0], np.float32).view(np.uint64)[0])
    raise TypeError(f"unsupported struct-for argument {type(a)}")


class StructFor:
    """`for I in ti.grouped(snode): <body>` with the body given as CUDA C++ (contract: csrc/b2_jit.cuh).

        scale = ti.struct_for('''B2_BODY(c) {
            float v = c.read<float>(0, c.I);            // x[I]
            c.write<float>(1, c.I, c.arg<float>(0) * v); // y[I] = a * x[I]   (activates y's cell)
        }''')
        scale(block, fields=[x, y], args=[2.0])

    Compiled once (NVRTC, sm_100a) at construction; a body that does not compile raises with the compiler log."""

    def __init__(self, source, name="b2_body"):
        if _prog.rt is None:
            raise RuntimeError("ti.init() first")
        self.name = name
        self.handle = lib.b2_jit_compile(_prog.rt, source.encode(), name.encode())
        L.check()
        self._rt = _prog.rt
        _prog.kernels.append(self)  # unloaded with the runtime (ti.reset)

    def __call__(self, snode, fields=(), args=(), gc=(), block_local=()):
        """gc: pointer / dynamic SNodes whose cells the body deactivates -- their GC runs right after the launch,
        which is where the reference's offloader puts it (offload.cpp:671-695).
        block_local: pads made with ti.bls(...) -- slot k of the body (c.bls_read<T>(k, J) / c.bls_add<T>(k, J, v))."""
        if self.handle is None or self._rt != _prog.rt:
            raise RuntimeError("struct-for kernel belongs to a runtime that was reset")
        t = _materialize()
        places = (C.c_int32 * max(1, len(fields)))(*[f.place_id for f in fields])
        slots = (C.c_uint64 * max(1, len(args)))(*[_arg_bits(a) for a in args])
        if block_local:
            specs = (L.B2BlsSpec * len(block_local))()
            for sp, d in zip(specs, block_local):
                f = fields[d["field"]]
                acc = {np.dtype(np.float32): 2, np.dtype(np.int32): 3, np.dtype(np.float64): 4}
                sp.field, sp.ndim, sp.halo = d["field"], len(d["lo"]), int(d["halo"])
                sp.mode = 1 if d["mode"] == "read" else acc[np.dtype(f.np_dtype)]
                for q in range(len(d["lo"])):
                    sp.axis[q], sp.extent[q], sp.lo[q], sp.hi[q] = d["axes"][q], int(d["extent"][q]), int(d["lo"][q]), int(d["hi"][q])
            lib.b2_jit_launch_bls(self.handle, t, snode.id, places, len(fields), slots, len(args), specs, len(block_local))
        else:
            lib.b2_jit_launch(self.handle, t, snode.id, places, len(fields), slots, len(args))
        L.check()
        for sn in gc:
            lib.b2_gc(t, sn.id)
            L.check()

    def range(self, begin, end, fields=(), args=(), gc=()):
        """`for i in range(begin, end)` over the source's B2_RANGE_BODY (particle loops; c.i = index)."""
        if self.handle is None or self._rt != _prog.rt:
            raise RuntimeError("kernel belongs to a runtime that was reset")
        t = _materialize()
        places = (C.c_int32 * max(1, len(fields)))(*[f.place_id for f in fields])
        slots = (C.c_uint64 * max(1, len(args)))(*[_arg_bits(a) for a in args])
        lib.b2_jit_launch_range(self.handle, t, int(begin), int(end), places, len(fields), slots, len(args))
        L.check()
        for sn in gc:
            lib.b2_gc(t, sn.id)
            L.check()

    def log(self):
        n = lib.b2_jit_log(self.handle, None, 0)
        buf = C.create_string_buffer(int(n))
        lib.b2_jit_log(self.handle, buf, n)
        return buf.value.decode()

    def destroy(self):
        if self.handle is not None and self._rt == _prog.rt:
            lib.b2_jit_destroy(self.handle)
        self.handle = None


def struct_for(source, name="b2_body"):
    return StructFor(source, name)
