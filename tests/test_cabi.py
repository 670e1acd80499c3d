"""CPU-side checks of the boundary: the shared library loads, exports every symbol that
include/b200mpm.h declares (and nothing undeclared), keeps the reference's struct sizes, and
fails loudly -- no CPU fallback -- when there is no CUDA device.  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(REPO, "include", "b200mpm.h")


def declared():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b((?:ti|b2)_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from taichi_b200 import _lib
    names = declared()
    assert len(names) > 60
    for n in names:
        assert hasattr(_lib.lib, n), n
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l and re.match(r"(ti|b2)_", l.split()[-1]))
    assert exported == names  # nothing undeclared leaks out either


def test_struct_layouts_match_reference_header():
    """sizeof values of c_api/include/taichi/taichi_core.h structs (x86-64): TiNdShape 68, TiNdArray 152,
    TiArgument 160 (largest union member TiNdArray, 8-aligned), TiMemoryAllocateInfo 24."""
    from taichi_b200 import _lib as L
    assert C.sizeof(L.TiNdShape) == 68
    assert C.sizeof(L.TiNdArray) == 152
    assert C.sizeof(L.TiArgument) == 160
    assert C.sizeof(L.TiMemoryAllocateInfo) == 24
    assert C.sizeof(L.TiMemorySlice) == 24
    assert C.sizeof(L.B2SNodeDesc) == 116
    src = f'#include "{HDR}"\n#include <stdio.h>\nint main(){{printf("%zu %zu %zu %zu %zu %zu", sizeof(TiNdShape), sizeof(TiNdArray), sizeof(TiArgument), sizeof(TiMemoryAllocateInfo), sizeof(B2SNodeDesc), sizeof(B2MpmParams));}}'
    exe = "/tmp/_b2_sizes"
    subprocess.run(["gcc", "-x", "c", "-", "-o", exe], input=src.encode(), check=True)
    got = subprocess.check_output([exe]).decode().split()
    assert got == ["68", "152", "160", "24", "116", str(C.sizeof(L.B2MpmParams))]


def test_version_and_error_channel_without_gpu():
    from taichi_b200 import _lib as L
    assert L.lib.ti_get_version() == 1007000
    L.lib.ti_set_last_error(-4, b"hello")
    assert L.last_error() == (-4, "hello")
    L.lib.ti_set_last_error(0, None)
    assert L.last_error() == (0, "")


def test_no_cpu_fallback():
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present")
    except ImportError:
        pass
    from taichi_b200 import _lib as L
    assert L.lib.ti_create_runtime(L.TI_ARCH_CUDA, 0) is None
    code, msg = L.last_error()
    assert code == -1 and "no CPU fallback" in msg
    L.lib.ti_set_last_error(0, None)
    assert L.lib.ti_create_runtime(4, 0) is None  # TI_ARCH_X64 is not provided by this library
    L.lib.ti_set_last_error(0, None)
    import taichi_b200 as ti
    with pytest.raises(RuntimeError):
        ti.init(arch=ti.cpu)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(REPO, "taichi_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in txt and "oracle/" not in txt.replace("oracle/_ref", "oracle/").replace("# oracle/", "") or f == "nope", f


def test_jit_body_compiles_without_a_device():
    """b2_jit_check: NVRTC turns a B2_BODY source (+ the embedded b2_internal.cuh / b2_jit.cuh) into an sm_100a
    cubin on the host alone; a broken body reports the compiler log through the last-error channel."""
    from taichi_b200 import _lib as L
    ok = L.lib.b2_jit_check(b"B2_BODY(c) { c.write<float>(0, c.I, 2.0f * c.read<float>(0, c.I) + c.arg<float>(0)); }", b"k")
    if ok < 0 and L.last_error()[0] == -1:
        import pytest
        L.lib.ti_set_last_error(0, None)
        pytest.skip("libnvrtc not installed on this host")
    assert ok > 1000, L.last_error()
    assert L.lib.b2_jit_check(b"B2_BODY(c) { nonsense(); }", b"bad_body") == -1
    code, msg = L.last_error()
    L.lib.ti_set_last_error(0, None)
    assert code == -4 and "nonsense" in msg and "bad_body" in msg


def test_bls_pad_geometry_known_answers():
    """ScratchPad bounds pinned by the reference's own unit tests: bls_analyzer_test.cpp:51-77 (block 8x8, accesses x[i+1, j]
    and x[i, j-3]: bounds[0] = (0, 1 + 8), bounds[1] = (-3, 8)), :79-110 (offsets (0,0), (0,-3): (0, 8), (-3, 8)) and
    make_block_local_test.cpp:91-120 (block 4x4, accesses (0,0) and (-1,-3): pad 5 x 7)."""
    from taichi_b200 import _lib as L
    i32 = C.c_int32

    def geom(block, offsets):
        nd = len(block)
        off = (i32 * (nd * len(offsets)))(*[v for o in offsets for v in o])
        low, high, size = (i32 * nd)(), (i32 * nd)(), (i32 * nd)()
        n = L.lib.b2_bls_pad_geometry(nd, (i32 * nd)(*block), len(offsets), off, low, high, size)
        return list(low), list(high), list(size), n
    assert geom([8, 8], [(1, 0), (0, -3)]) == ([0, -3], [9, 8], [9, 11], 99)
    assert geom([8, 8], [(0, 0), (0, -3)]) == ([0, -3], [8, 8], [8, 11], 88)
    assert geom([4, 4], [(0, 0), (-1, -3)]) == ([-1, -3], [4, 4], [5, 7], 35)
    assert geom([4, 4, 4], [(0, 0, 0), (2, 2, 2)])[2:] == ([6, 6, 6], 216)          # the MPM quadratic stencil: (4 + 2)^3


def test_jit_check_block_local_and_reductions():
    from taichi_b200 import _lib as L
    ok = L.lib.b2_jit_check(b"""B2_BODY(c) {
        int J[B2_MAX_AXES]; c.offset(J, 0, 1);
        c.bls_add<float>(1, J, c.bls_read<float>(0, c.I));
        c.reduce_add<float>(0, 1.0f); c.reduce_max<int>(1, c.I[0]); c.reduce_min<double>(2, 0.5);
    }""", b"bls_body")
    if ok < 0 and "libnvrtc" in L.last_error()[1]:
        pytest.skip("libnvrtc not available here")
    assert ok > 0, L.last_error()
