"""The drop-in boundary checked against the REFERENCE's own headers, not against itself:
 * tests/capi/layout_probe.c is compiled once with /root/reference/c_api/include/taichi/{taichi_core.h,taichi_cuda.h} and
   once with include/b200mpm.h; the sizeof / offsetof / enumerator tables must be identical;
 * tests/capi/mpm88_client.cpp -- the flow of c_api/tests/mpm88_test.cpp:18-88 without gtest -- compiles against the
   reference's header-only C++ wrapper (cpp/taichi.hpp) and links with libb200mpm.so with no unresolved ti_* symbol
   (it is RUN on a GPU box by tests/test_gpu_capi_client.py);
 * every ti_* function the reference headers declare for this path is exported by the library.
CPU only; needs /root/reference (absent on the GPU box: skipped there, the prebuilt client binary travels)."""
import os
import re
import subprocess

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_INC = "/root/reference/c_api/include/taichi"
CAPI = os.path.join(REPO, "tests", "capi")
LIB = os.path.join(REPO, "taichi_b200", "lib", "libb200mpm.so")
needs_ref = pytest.mark.skipif(not os.path.isdir(REF_INC), reason="reference tree not present")


def exported():
    out = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True, check=True).stdout
    return {l.split()[-1] for l in out.splitlines() if l.strip()}


@needs_ref
def test_struct_layouts_equal_the_reference_headers():
    subprocess.run(["make", "-C", CAPI, "bin/layout_b2", "bin/layout_ref"], check=True, capture_output=True)
    ref = subprocess.run([os.path.join(CAPI, "bin", "layout_ref")], capture_output=True, text=True, check=True).stdout
    b2 = subprocess.run([os.path.join(CAPI, "bin", "layout_b2")], capture_output=True, text=True, check=True).stdout
    assert len(ref.splitlines()) > 80
    assert ref == b2


@needs_ref
def test_reference_cpp_wrapper_client_links_against_the_library():
    r = subprocess.run(["make", "-C", CAPI, "-B", "bin/mpm88_client"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    binary = os.path.join(CAPI, "bin", "mpm88_client")
    und = subprocess.run(["nm", "-D", "--undefined-only", binary], capture_output=True, text=True, check=True).stdout
    wanted = {l.split()[-1] for l in und.splitlines() if re.search(r"\b(ti_|b2_)", l)}
    assert {"ti_create_runtime", "ti_load_aot_module", "ti_launch_kernel", "ti_allocate_memory", "ti_wait", "b2_launch_count"} <= wanted
    assert wanted <= exported(), sorted(wanted - exported())


@needs_ref
def test_every_reference_entry_point_is_exported():
    names = set()
    for h in ("taichi_core.h", "taichi_cuda.h"):
        text = open(os.path.join(REF_INC, h)).read()
        names |= set(re.findall(r"TI_API_CALL\s+(ti_[a-z0-9_]+)\s*\(", text))
    assert len(names) >= 30
    missing = names - exported()
    assert not missing, sorted(missing)


def test_entry_points_of_the_reference_wrapper_are_exported():
    """The same list, committed (taichi_core.h:881-1106, taichi_cuda.h:12-32): holds where the reference tree is absent."""
    names = """ti_get_version ti_get_available_archs ti_get_last_error ti_set_last_error ti_create_runtime ti_destroy_runtime
    ti_set_runtime_capabilities_ext ti_get_runtime_capabilities ti_allocate_memory ti_free_memory ti_map_memory ti_unmap_memory
    ti_allocate_image ti_free_image ti_create_sampler ti_destroy_sampler ti_copy_memory_device_to_device
    ti_copy_image_device_to_device ti_track_image_ext ti_transition_image ti_launch_kernel ti_launch_compute_graph ti_flush
    ti_wait ti_load_aot_module ti_create_aot_module ti_destroy_aot_module ti_get_aot_module_kernel
    ti_get_aot_module_compute_graph ti_export_cuda_memory ti_import_cuda_memory ti_set_cuda_stream ti_get_cuda_stream""".split()
    assert set(names) <= exported(), sorted(set(names) - exported())
