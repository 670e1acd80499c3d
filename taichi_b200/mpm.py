"""Host-side mirrors of the reference's MPM entry points, over the C ABI.

* `AotModule` / `Kernel` / `NdArray` mirror the C++ wrappers of
  c_api/include/taichi/cpp/taichi.hpp (ti::Runtime, ti::AotModule, ti::Kernel, ti::NdArray) as
  used by c_api/tests/mpm88_test.cpp:18-88: kernels `substep_reset_grid`, `substep_p2g`,
  `substep_update_grid_v`, `substep_g2p`, `init_particles` with ndarray arguments.
* `Mpm3D` is the block-binned sparse solver (mpm3d.py:36-81 physics over
  pointer->bitmasked->dense(4^3) with ti.block_local P2G/G2P).
"""
import ctypes as C
import numpy as np

from . import _lib as L

lib = L.lib


class Runtime:
    """ti::Runtime (taichi.hpp) -- one per process and device."""

    def __init__(self, device_index=0, device_memory_GB=None):
        if device_memory_GB is not None:
            lib.b2_set_pool_size(C.c_uint64(int(device_memory_GB * (1 << 30))))
        self.handle = lib.ti_create_runtime(L.TI_ARCH_CUDA, device_index)
        L.check()
        self.memories = []

    def wait(self):
        lib.ti_wait(self.handle)
        L.check()

    def launches(self):
        return int(lib.b2_launch_count(self.handle))

    def allocate_ndarray(self, shape, elem_shape=(), dtype=np.float32, host_access=False):
        return NdArray(self, shape, elem_shape, dtype, host_access)

    def load_aot_module(self, path):
        return AotModule(self, path)

    def destroy(self):
        if self.handle:
            lib.ti_destroy_runtime(self.handle)
            self.handle = None


class NdArray:
    """ti::NdArray<T>: TiMemory + shape + elem_shape."""

    def __init__(self, rt, shape, elem_shape=(), dtype=np.float32, host_access=False):
        self.rt, self.shape, self.elem_shape, self.dtype = rt, tuple(shape), tuple(elem_shape), np.dtype(dtype)
        n = int(np.prod(self.shape + self.elem_shape, dtype=np.int64))
        info = L.TiMemoryAllocateInfo(size=max(n, 1) * self.dtype.itemsize, host_write=int(host_access),
                                      host_read=int(host_access), export_sharing=0, usage=1)
        self.memory = lib.ti_allocate_memory(rt.handle, C.byref(info))
        L.check()
        self.nbytes = n * self.dtype.itemsize

    def device_ptr(self):
        info = L.TiCudaMemoryInteropInfo()
        lib.ti_export_cuda_memory(self.rt.handle, self.memory, C.byref(info))
        L.check()
        return info.ptr

    def write(self, arr):
        arr = np.ascontiguousarray(arr, dtype=self.dtype)
        assert arr.nbytes == self.nbytes
        p = lib.ti_map_memory(self.rt.handle, self.memory)
        L.check()
        C.memmove(p, arr.ctypes.data, self.nbytes)
        lib.ti_unmap_memory(self.rt.handle, self.memory)
        L.check()

    def read(self):
        out = np.empty(self.shape + self.elem_shape, dtype=self.dtype)
        p = lib.ti_map_memory(self.rt.handle, self.memory)
        L.check()
        C.memmove(out.ctypes.data, p, self.nbytes)
        lib.ti_unmap_memory(self.rt.handle, self.memory)
        L.check()
        return out

    def argument(self):
        a = L.TiArgument()
        a.type = L.TI_ARGUMENT_TYPE_NDARRAY
        nd = a.value.ndarray
        nd.memory = self.memory
        nd.shape.dim_count = len(self.shape)
        for i, s in enumerate(self.shape):
            nd.shape.dims[i] = s
        nd.elem_shape.dim_count = len(self.elem_shape)
        for i, s in enumerate(self.elem_shape):
            nd.elem_shape.dims[i] = s
        nd.elem_type = L.TI_DATA_TYPE_F32 if self.dtype == np.float32 else L.TI_DATA_TYPE_I32
        return a


def f32_argument(v):
    a = L.TiArgument()
    a.type = L.TI_ARGUMENT_TYPE_F32
    a.value.f32 = float(v)
    return a


class Kernel:
    def __init__(self, rt, handle):
        self.rt, self.handle, self.args = rt, handle, {}

    def __setitem__(self, i, arg):
        self.args[i] = arg

    def launch(self, *args):
        if args:
            self.args = dict(enumerate(args))
        n = len(self.args)
        arr = (L.TiArgument * max(n, 1))()
        for i in range(n):
            a = self.args[i]
            arr[i] = a.argument() if isinstance(a, NdArray) else f32_argument(a)
        lib.ti_launch_kernel(self.rt.handle, self.handle, n, arr)
        L.check()


class AotModule:
    def __init__(self, rt, path):
        self.rt = rt
        self.handle = lib.ti_load_aot_module(rt.handle, path.encode())
        L.check()

    def get_kernel(self, name):
        h = lib.ti_get_aot_module_kernel(self.handle, name.encode())
        L.check()
        return Kernel(self.rt, h)


def _ptr(a):
    """numpy array / torch tensor / None -> void* (host or device; the library detects which)."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return C.c_void_p(a.ctypes.data)
    if hasattr(a, "data_ptr"):  # torch tensor: zero-copy, must be contiguous (kernel_impl.py:745-796)
        assert a.is_contiguous()
        return C.c_void_p(a.data_ptr())
    raise TypeError(type(a))


LAYOUTS = {"dense": 0, "pointer_bitmasked": 1, "pointer_pointer": 2}


class Mpm3D:
    """3-D MLS-MPM on a (sparse) SNode grid with block-binned particles.

    layout: 'dense'              ti.root.dense(ijk, n/4).dense(ijk, 4)
            'pointer_bitmasked'  ti.root.pointer(ijk, n/32).bitmasked(ijk, 8).dense(ijk, 4)
            'pointer_pointer'    ti.root.pointer(ijk, n/32).pointer(ijk, 8).dense(ijk, 4)
    """

    def __init__(self, rt, n_grid, max_particles, dt, layout="pointer_bitmasked", E=400.0, p_rho=1.0, gravity=9.8,
                 bound=3, block_local=True, bls_activate_halo=False, deactivate_every=0, max_bins=0, model="fluid", nu=0.2,
                 p_vol=0.0):
        """model: 'fluid' = mpm3d.py (weakly compressible, J);  'materials' = deformation gradient F, plastic Jp and a
        material id per particle with ti.svd (mpm128.py:36-63 in 3-D: 0 liquid, 1 jelly, 2 snow; E, nu -> mu_0, lambda_0)."""
        self.rt = rt
        self.n_grid = n_grid
        self.model = {"fluid": 0, "materials": 1}[model]
        p = L.B2MpmParams(n_grid=n_grid, layout=LAYOUTS[layout], dt=dt, E=E, p_rho=p_rho, gravity=gravity, bound=bound,
                          block_local=int(block_local), bls_activate_halo=int(bls_activate_halo), slab_lo=0, slab_hi=0,
                          deactivate_every=int(deactivate_every), max_particles=max_particles, max_bins=int(max_bins),
                          model=self.model, nu=nu, p_vol=p_vol)
        self.handle = lib.b2_mpm_create(rt.handle, C.byref(p))
        L.check()
        self.tree = lib.b2_mpm_tree(self.handle)
        self.leaf = lib.b2_mpm_leaf_snode(self.handle)

    def _ok(self, rc):
        L.check()
        assert rc == 0

    def set_particles(self, x, v=None, Cm=None, J=None):
        n = x.shape[0]
        self._ok(lib.b2_mpm_set_particles(self.handle, n, _ptr(x), _ptr(v), _ptr(Cm), _ptr(J)))

    def add_particles(self, x, v=None, Cm=None, J=None, ids=None):
        n = x.shape[0]
        self._ok(lib.b2_mpm_add_particles(self.handle, n, _ptr(x), _ptr(v), _ptr(Cm), _ptr(J), _ptr(ids)))

    def set_particles_model(self, x, v=None, Cm=None, F=None, Jp=None, material=None):
        """model='materials': F (n, 3, 3) float32, Jp (n,) float32, material (n,) int32; None = identity / 1 / 0."""
        self._ok(lib.b2_mpm_set_particles_model(self.handle, x.shape[0], _ptr(x), _ptr(v), _ptr(Cm), _ptr(F), _ptr(Jp), _ptr(material)))

    def get_particles_model(self):
        n = self.num_particles
        x, v, Cm = np.empty((n, 3), np.float32), np.empty((n, 3), np.float32), np.empty((n, 3, 3), np.float32)
        F, Jp, mat = np.empty((n, 3, 3), np.float32), np.empty(n, np.float32), np.empty(n, np.int32)
        self._ok(lib.b2_mpm_get_particles_model(self.handle, _ptr(x), _ptr(v), _ptr(Cm), _ptr(F), _ptr(Jp), _ptr(mat)))
        return x, v, Cm, F, Jp, mat

    @property
    def num_particles(self):
        return int(lib.b2_mpm_num_particles(self.handle))

    def get_particles(self, x=None, v=None, Cm=None, J=None, wait=True):
        """Fills the given arrays (numpy or torch, any may be None); allocates numpy ones if none given.
        wait=False: host arrays (page-locked) are filled by a copy stream after return -- call wait_io()
        before reading them; the download then overlaps whatever is enqueued next."""
        if x is None and v is None and Cm is None and J is None:
            n = self.num_particles
            x, v = np.empty((n, 3), np.float32), np.empty((n, 3), np.float32)
            Cm, J = np.empty((n, 3, 3), np.float32), np.empty(n, np.float32)
        fn = lib.b2_mpm_get_particles if wait else lib.b2_mpm_get_particles_async
        self._ok(fn(self.handle, _ptr(x), _ptr(v), _ptr(Cm), _ptr(J)))
        return x, v, Cm, J

    def wait_io(self):
        """Blocks until every particle upload / download issued through this solver has finished."""
        self._ok(lib.b2_mpm_wait_io(self.handle))

    def substep(self, n=1):
        self._ok(lib.b2_mpm_substep(self.handle, n))

    def clear_grid(self):
        self._ok(lib.b2_mpm_clear_grid(self.handle))

    def p2g(self):
        self._ok(lib.b2_mpm_p2g(self.handle))

    def grid_op(self):
        self._ok(lib.b2_mpm_grid_op(self.handle))

    def g2p(self):
        self._ok(lib.b2_mpm_g2p(self.handle))

    def grid_to_numpy(self):
        out = np.zeros((self.n_grid,) * 3 + (4,), np.float32)
        self._ok(lib.b2_mpm_grid_to_dense(self.handle, _ptr(out)))
        return out

    def grid_blocks(self):
        """Active 4^3 blocks as (keys[n] sorted ascending, cells[n][64][4]); key = (bx*nb+by)*nb+bz."""
        n = int(lib.b2_mpm_grid_blocks(self.handle, None, None, 0))
        L.check()
        keys, cells = np.zeros(max(n, 1), np.int32), np.zeros((max(n, 1), 64, 4), np.float32)
        lib.b2_mpm_grid_blocks(self.handle, _ptr(keys), _ptr(cells), n)
        L.check()
        order = np.argsort(keys[:n])
        return keys[:n][order], cells[:n][order]

    def active_blocks(self):
        """Sorted (nblocks, 3) block coordinates of the active 4^3 leaf blocks (listgen contents)."""
        n = lib.b2_listgen(self.tree, self.leaf)
        L.check()
        out = np.zeros((max(n, 1), 14), np.int32)
        lib.b2_get_list(self.tree, self.leaf, _ptr(out), n)
        L.check()
        b = out[:n, :3]
        return b[np.lexsort(b.T[::-1])] if n else b

    def stats(self):
        out = (C.c_double * 8)()
        self._ok(lib.b2_mpm_stats(self.handle, out))
        return dict(particles=int(out[0]), slots=int(out[1]), occupied_slots=int(out[2]), chunks=int(out[3]),
                    active_blocks=int(out[4]))

    DIAG_KEYS = ("particles", "px", "py", "pz", "kinetic_energy", "sum_J", "max_speed", "max_dJ")

    def diagnostics(self):
        """0-D reductions over the particles (momentum, kinetic energy, max |v| for the CFL check, ...)."""
        out = (C.c_double * 8)()
        self._ok(lib.b2_mpm_diagnostics(self.handle, out))
        return dict(zip(self.DIAG_KEYS, list(out)))

    def set_profiling(self, on):
        lib.b2_mpm_set_profiling(self.handle, int(on))

    def phase_ms(self):
        out = (C.c_float * 6)()
        lib.b2_mpm_phase_ms(self.handle, out)
        return dict(zip(("clear", "p2g", "listgen", "grid_op", "g2p", "deactivate_gc"), list(out)))

    def destroy(self):
        if self.handle:
            lib.b2_mpm_destroy(self.handle)
            self.handle = None
