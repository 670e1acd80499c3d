"""ctypes binding of libb200mpm.so (include/b200mpm.h).

This is the reference-side binding a maintainer would add: where the reference's Python
frontend calls into `taichi_python` (pybind, taichi/python/export_lang.cpp) for
SNode.dense/pointer/bitmasked/dynamic/place, Program.launch_kernel etc., this module
calls the C ABI.  The library is REQUIRED: there is no Python or CPU fallback -- if the
in-tree shared object is missing or cannot be loaded, importing this module raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2_LIB") or os.path.join(_HERE, "lib", "libb200mpm.so")  # B2_LIB: experiment builds of the same library

TI_ARCH_CUDA = 3
TI_ARGUMENT_TYPE_I32, TI_ARGUMENT_TYPE_F32, TI_ARGUMENT_TYPE_NDARRAY = 0, 1, 2
TI_DATA_TYPE_F32, TI_DATA_TYPE_I32 = 1, 5
ROOT, DENSE, POINTER, BITMASKED, DYNAMIC, PLACE = range(6)

i32p = C.POINTER(C.c_int32)
f32p = C.POINTER(C.c_float)


class B2SNodeDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("parent", C.c_int32), ("num_axes", C.c_int32),
                ("axes", C.c_int32 * 12), ("sizes", C.c_int32 * 12),
                ("dtype_bytes", C.c_int32), ("chunk_size", C.c_int32)]


class B2BlsSpec(C.Structure):
    _fields_ = [("field", C.c_int32), ("mode", C.c_int32), ("ndim", C.c_int32), ("halo", C.c_int32),
                ("axis", C.c_int32 * 3), ("extent", C.c_int32 * 3), ("lo", C.c_int32 * 3), ("hi", C.c_int32 * 3)]


class B2MpmParams(C.Structure):
    _fields_ = [("n_grid", C.c_int32), ("layout", C.c_int32), ("dt", C.c_float), ("E", C.c_float),
                ("p_rho", C.c_float), ("gravity", C.c_float), ("bound", C.c_int32),
                ("block_local", C.c_int32), ("bls_activate_halo", C.c_int32),
                ("slab_lo", C.c_int32), ("slab_hi", C.c_int32), ("deactivate_every", C.c_int32), ("max_particles", C.c_int64), ("max_bins", C.c_int64),
                ("model", C.c_int32), ("nu", C.c_float), ("p_vol", C.c_float), ("pad_", C.c_int32)]


class TiMemoryAllocateInfo(C.Structure):
    _fields_ = [("size", C.c_uint64), ("host_write", C.c_uint32), ("host_read", C.c_uint32),
                ("export_sharing", C.c_uint32), ("usage", C.c_uint32)]


class TiMemorySlice(C.Structure):
    _fields_ = [("memory", C.c_void_p), ("offset", C.c_uint64), ("size", C.c_uint64)]


class TiNdShape(C.Structure):
    _fields_ = [("dim_count", C.c_uint32), ("dims", C.c_uint32 * 16)]


class TiNdArray(C.Structure):
    _fields_ = [("memory", C.c_void_p), ("shape", TiNdShape), ("elem_shape", TiNdShape), ("elem_type", C.c_uint32)]


class _TiImageExtent(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("depth", C.c_uint32), ("array_layer_count", C.c_uint32)]


class _TiTexture(C.Structure):
    _fields_ = [("image", C.c_void_p), ("sampler", C.c_void_p), ("dimension", C.c_uint32),
                ("extent", _TiImageExtent), ("format", C.c_uint32)]


class _TiScalarValue(C.Union):
    _fields_ = [("x8", C.c_uint8), ("x16", C.c_uint16), ("x32", C.c_uint32), ("x64", C.c_uint64)]


class _TiScalar(C.Structure):
    _fields_ = [("type", C.c_uint32), ("value", _TiScalarValue)]


class _TiTensorValue(C.Union):
    _fields_ = [("x8", C.c_uint8 * 128), ("x16", C.c_uint16 * 64), ("x32", C.c_uint32 * 32), ("x64", C.c_uint64 * 16)]


class _TiTensorValueWithLength(C.Structure):
    _fields_ = [("length", C.c_uint32), ("data", _TiTensorValue)]


class _TiTensor(C.Structure):
    _fields_ = [("type", C.c_uint32), ("contents", _TiTensorValueWithLength)]


class TiArgumentValue(C.Union):
    _fields_ = [("i32", C.c_int32), ("f32", C.c_float), ("ndarray", TiNdArray), ("texture", _TiTexture),
                ("scalar", _TiScalar), ("tensor", _TiTensor)]


class TiArgument(C.Structure):
    _fields_ = [("type", C.c_uint32), ("value", TiArgumentValue)]


class TiCudaMemoryInteropInfo(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("size", C.c_uint64)]


# name -> (restype, argtypes); every symbol declared in include/b200mpm.h
SIGNATURES = {
    "ti_get_version": (C.c_uint32, []),
    "ti_get_available_archs": (None, [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "ti_get_last_error": (C.c_int32, [C.POINTER(C.c_uint64), C.c_char_p]),
    "ti_set_last_error": (None, [C.c_int32, C.c_char_p]),
    "ti_create_runtime": (C.c_void_p, [C.c_uint32, C.c_uint32]),
    "ti_destroy_runtime": (None, [C.c_void_p]),
    "ti_allocate_memory": (C.c_void_p, [C.c_void_p, C.POINTER(TiMemoryAllocateInfo)]),
    "ti_free_memory": (None, [C.c_void_p, C.c_void_p]),
    "ti_map_memory": (C.c_void_p, [C.c_void_p, C.c_void_p]),
    "ti_unmap_memory": (None, [C.c_void_p, C.c_void_p]),
    "ti_copy_memory_device_to_device": (None, [C.c_void_p, C.POINTER(TiMemorySlice), C.POINTER(TiMemorySlice)]),
    "ti_launch_kernel": (None, [C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(TiArgument)]),
    "ti_launch_compute_graph": (None, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    "ti_flush": (None, [C.c_void_p]),
    "ti_set_runtime_capabilities_ext": (None, [C.c_void_p, C.c_uint32, C.c_void_p]),
    "ti_get_runtime_capabilities": (None, [C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p]),
    "ti_create_aot_module": (C.c_void_p, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "ti_allocate_image": (C.c_void_p, [C.c_void_p, C.c_void_p]),
    "ti_free_image": (None, [C.c_void_p, C.c_void_p]),
    "ti_create_sampler": (C.c_void_p, [C.c_void_p, C.c_void_p]),
    "ti_destroy_sampler": (None, [C.c_void_p, C.c_void_p]),
    "ti_copy_image_device_to_device": (None, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "ti_track_image_ext": (None, [C.c_void_p, C.c_void_p, C.c_uint32]),
    "ti_transition_image": (None, [C.c_void_p, C.c_void_p, C.c_uint32]),
    "ti_wait": (None, [C.c_void_p]),
    "ti_load_aot_module": (C.c_void_p, [C.c_void_p, C.c_char_p]),
    "ti_destroy_aot_module": (None, [C.c_void_p]),
    "ti_get_aot_module_kernel": (C.c_void_p, [C.c_void_p, C.c_char_p]),
    "ti_get_aot_module_compute_graph": (C.c_void_p, [C.c_void_p, C.c_char_p]),
    "ti_export_cuda_memory": (None, [C.c_void_p, C.c_void_p, C.POINTER(TiCudaMemoryInteropInfo)]),
    "ti_import_cuda_memory": (C.c_void_p, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "ti_set_cuda_stream": (None, [C.c_void_p]),
    "ti_get_cuda_stream": (None, [C.POINTER(C.c_void_p)]),
    "b2_create_tree": (C.c_void_p, [C.c_void_p, C.POINTER(B2SNodeDesc), C.c_int32]),
    "b2_destroy_tree": (None, [C.c_void_p]),
    "b2_snode_num_cells": (C.c_int64, [C.c_void_p, C.c_int32]),
    "b2_snode_cell_size": (C.c_int64, [C.c_void_p, C.c_int32]),
    "b2_snode_node_size": (C.c_int64, [C.c_void_p, C.c_int32]),
    "b2_activate": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]),
    "b2_deactivate": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]),
    "b2_is_active": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "b2_listgen": (C.c_int64, [C.c_void_p, C.c_int32]),
    "b2_listgen_async": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b2_get_list": (C.c_int64, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]),
    "b2_struct_for_coords": (C.c_int64, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]),
    "b2_deactivate_all": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b2_gc": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b2_alloc_stats": (C.c_int32, [C.c_void_p, C.c_int32, i32p]),
    "b2_num_dynamically_allocated": (C.c_int64, [C.c_void_p, C.c_int32]),
    "b2_write": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32]),
    "b2_read": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "b2_dynamic_append_i32": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "b2_dynamic_length": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p]),
    "b2_field_fill": (C.c_int32, [C.c_void_p, C.c_int32, C.c_uint64]),
    "b2_field_to_dense": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p]),
    "b2_field_from_dense": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p]),
    "b2_field_reduce": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "b2_reduce": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]),
    "b2_reduce_device": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]),
    "b2_reduce_noalloc": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]),
    "b2_sync": (None, [C.c_void_p]),
    "b2_device_index": (C.c_int32, [C.c_void_p]),
    "b2_pool_bytes_used": (C.c_int64, [C.c_void_p]),
    "b2_set_pool_size": (None, [C.c_uint64]),
    "b2_set_random_seed": (None, [C.c_int32]),
    "b2_launch_count": (C.c_uint64, [C.c_void_p]),
    "b2_check_runtime_error": (C.c_int32, [C.c_void_p, C.c_char_p, C.c_uint64]),
    "b2_mpm_create": (C.c_void_p, [C.c_void_p, C.POINTER(B2MpmParams)]),
    "b2_mpm_destroy": (None, [C.c_void_p]),
    "b2_mpm_tree": (C.c_void_p, [C.c_void_p]),
    "b2_mpm_leaf_snode": (C.c_int32, [C.c_void_p]),
    "b2_mpm_set_particles": (C.c_int32, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2_mpm_add_particles": (C.c_int32, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2_mpm_num_particles": (C.c_int64, [C.c_void_p]),
    "b2_mpm_set_particles_model": (C.c_int32, [C.c_void_p, C.c_int64] + [C.c_void_p] * 6),
    "b2_mpm_get_particles_model": (C.c_int32, [C.c_void_p] + [C.c_void_p] * 6),
    "b2_mpm_get_particles": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2_mpm_get_particles_async": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2_mpm_wait_io": (C.c_int32, [C.c_void_p]),
    "b2_mpm_substep": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b2_mpm_clear_grid": (C.c_int32, [C.c_void_p]),
    "b2_mpm_p2g": (C.c_int32, [C.c_void_p]),
    "b2_mpm_grid_op": (C.c_int32, [C.c_void_p]),
    "b2_mpm_g2p": (C.c_int32, [C.c_void_p]),
    "b2_mpm_grid_to_dense": (C.c_int32, [C.c_void_p, C.c_void_p]),
    "b2_mpm_grid_blocks": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "b2_mpm_stats": (C.c_int32, [C.c_void_p, C.POINTER(C.c_double)]),
    "b2_mpm_invalidate_lists": (C.c_int32, [C.c_void_p]),
    "b2_memory_profiler_info": (C.c_int64, [C.c_void_p, C.c_char_p, C.c_int64]),
    "b2_mpm_diagnostics": (C.c_int32, [C.c_void_p, C.POINTER(C.c_double)]),
    "b2_mpm_diagnostics_device": (C.c_int32, [C.c_void_p, C.c_void_p]),
    "b2_slab_diagnostics": (C.c_int32, [C.c_void_p, C.POINTER(C.c_double)]),
    "b2_jit_compile": (C.c_void_p, [C.c_void_p, C.c_char_p, C.c_char_p]),
    "b2_jit_check": (C.c_int64, [C.c_char_p, C.c_char_p]),
    "b2_jit_destroy": (None, [C.c_void_p]),
    "b2_jit_log": (C.c_int64, [C.c_void_p, C.c_char_p, C.c_int64]),
    "b2_jit_launch": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_uint64), C.c_int32]),
    "b2_jit_launch_bls": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_uint64), C.c_int32,
                                      C.POINTER(B2BlsSpec), C.c_int32]),
    "b2_bls_pad_geometry": (C.c_int64, [C.c_int32, C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                        C.POINTER(C.c_int32)]),
    "b2_jit_launch_range": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_uint64), C.c_int32]),
    "b2_mpm_set_profiling": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b2_mpm_set_deactivate_every": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b2_mpm_set_compact_every": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b2_mpm_phase_ms": (C.c_int32, [C.c_void_p, f32p]),
    "b2_mpm_pack_blocks": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]),
    "b2_mpm_unpack_blocks": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32]),
    "b2_mpm_extract_migrants": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]),
    "b2_mpm_insert_records": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64]),
    "b2_mpm_extract_migrants2": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]),
    "b2_mpm_insert_records2": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]),
    "b2_mpm_begin_substep": (C.c_int32, [C.c_void_p]),
    "b2_mpm_end_substep": (C.c_int32, [C.c_void_p]),
    "b2_mpm_zero_headers": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),

    "b2_mpm_runtime": (C.c_void_p, [C.c_void_p]),
    "b2_comm_unique_id": (C.c_int32, [C.c_void_p]),
    "b2_comm_init": (C.c_void_p, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]),
    "b2_comm_loopback": (C.c_void_p, [C.c_void_p, C.c_int32]),
    "b2_comm_destroy": (None, [C.c_void_p]),
    "b2_slab_create": (C.c_void_p, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int64]),
    "b2_slab_destroy": (None, [C.c_void_p]),
    "b2_slab_substep": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b2_slab_substep_lockstep": (C.c_int32, [C.POINTER(C.c_void_p), C.c_int32, C.c_int32]),
    "b2_slab_tune": (C.c_int32, [C.c_void_p]),
    "b2_mpm_export": (C.c_int64, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "b2_mpm_export_async": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
}


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(make -C taichi_b200/csrc). taichi_b200 has no fallback path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()


class B2RuntimeError(RuntimeError):
    pass


def last_error():
    size = C.c_uint64(0)
    code = lib.ti_get_last_error(C.byref(size), None)
    if code == 0:
        return 0, ""
    buf = C.create_string_buffer(int(size.value) + 1)
    size = C.c_uint64(len(buf))
    code = lib.ti_get_last_error(C.byref(size), buf)
    return code, buf.value.decode(errors="replace")


def check():
    """Raise if the last C-ABI call recorded an error (the reference's TI_CHECK idiom, then clear)."""
    code, msg = last_error()
    if code != 0:
        lib.ti_set_last_error(0, None)
        raise B2RuntimeError(f"TiError {code}: {msg}")
