"""b2_reduce_noalloc: the allocation-free variant of b2_reduce_device (static __device__ scratch, include/b200mpm.h).
Collected last on purpose.  Both entry points are called through the C ABI on the
same inputs; the test passes alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

from _b2 import have_gpu, runtime

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not have_gpu(), reason="needs a CUDA device")]

MAX_N = 1 << 28  # B2_REDUCE_NOALLOC_MAX_N


def _both(L, rt, t, op, dtype, out_a, out_b):
    import torch
    p, n = C.c_void_p(t.data_ptr()), t.numel()
    torch.cuda.synchronize()  # torch filled the inputs on its own stream; the library enqueues on the runtime's
    assert L.lib.b2_reduce_device(rt, p, n, op, dtype, C.c_void_p(out_a.data_ptr())) == 0
    L.check()
    assert L.lib.b2_reduce_noalloc(rt, p, n, op, dtype, C.c_void_p(out_b.data_ptr())) == 0
    L.check()
    L.lib.b2_sync(rt)
    torch.cuda.synchronize()
    return out_a.cpu().numpy()[0], out_b.cpu().numpy()[0]


def test_noalloc_matches_reduce_device_twice_in_a_row():
    import torch
    from taichi_b200 import _lib as L
    rt = runtime()
    g = torch.Generator(device="cuda").manual_seed(11)
    # two calls in a row per case: the second (a different, shorter input) sees the array as the first left it
    for n1, n2 in ((8 * 1024 * 1024 + 3, 1000), (5, 300001), (0, 257)):
        for dtype in (0, 1):
            for op in (0, 1, 2):
                for n in (n1, n2):
                    if dtype == 0:
                        t = torch.rand(n + 1, device="cuda", generator=g)[1:] - 0.25  # + 1, [1:]: a 4-byte aligned base too
                        oa, ob = torch.zeros(1, device="cuda"), torch.full((1,), 7.0, device="cuda")
                    else:
                        t = torch.randint(-1000, 1000, (n,), device="cuda", dtype=torch.int32, generator=g)
                        oa, ob = torch.zeros(1, device="cuda", dtype=torch.int32), torch.full((1,), 7, device="cuda", dtype=torch.int32)
                    a, b = _both(L, rt, t, op, dtype, oa, ob)
                    if dtype == 1 or op != 0:
                        assert a == b, (n, dtype, op, a, b)  # integer sums and min / max: bit-exact
                    else:
                        ref = float(t.double().sum().item())
                        # fp32 sums differ by summation order only (b2_reduce_device's own atomics are unordered);
                        # both are held to 2e-5 relative to the fp64 sum, and to each other
                        assert abs(float(a) - float(b)) <= 2e-5 * max(1.0, abs(ref)), (n, a, b, ref)
                        assert abs(float(b) - ref) <= 2e-5 * max(1.0, abs(ref)), (n, b, ref)
    # deterministic: the same input twice gives the same bits
    t = torch.rand(3_000_017, device="cuda", generator=g)
    o1, o2 = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
    torch.cuda.synchronize()
    for o in (o1, o2):
        assert L.lib.b2_reduce_noalloc(rt, C.c_void_p(t.data_ptr()), t.numel(), 0, 0, C.c_void_p(o.data_ptr())) == 0
    L.lib.b2_sync(rt)
    torch.cuda.synchronize()
    assert o1.cpu().numpy().tobytes() == o2.cpu().numpy().tobytes()


def test_noalloc_rejects_sizes_beyond_the_scratch():
    import torch
    from taichi_b200 import _lib as L
    rt = runtime()
    t, o = torch.zeros(16, device="cuda"), torch.zeros(1, device="cuda")
    rc = L.lib.b2_reduce_noalloc(rt, C.c_void_p(t.data_ptr()), MAX_N + 1, 0, 0, C.c_void_p(o.data_ptr()))
    assert rc != 0
    with pytest.raises(L.B2RuntimeError, match="exceeds the static scratch"):
        L.check()
    # and the entry point still works afterwards
    assert L.lib.b2_reduce_noalloc(rt, C.c_void_p(t.data_ptr()), 16, 2, 0, C.c_void_p(o.data_ptr())) == 0
    L.check()
    L.lib.b2_sync(rt)
    torch.cuda.synchronize()
    assert float(o.item()) == 0.0
