import sys, os, ctypes as C
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from taichi_b200 import _lib as L
from _b2 import vp
from _desc import coords12, TreeBuilder
lib = L.lib
coords = coords12(np.arange(4096, dtype=np.int32)[:, None] * 65536 + 7)
vals = np.full(4096, 1234567, np.int32)
buf = C.create_string_buffer(256)
for pool_mb in (48, 256):
    for n in (1, 8, 64, 4096):
        lib.b2_set_pool_size(C.c_uint64(pool_mb << 20))
        rt = lib.ti_create_runtime(L.TI_ARCH_CUDA, 0); L.check()
        tb = TreeBuilder(); p = tb.pointer(0, "i", 4096); d = tb.dense(p, "i", 65536); x = tb.place(d, 4)
        arr = tb.array()
        tree = lib.b2_create_tree(rt, C.cast(arr, C.POINTER(L.B2SNodeDesc)), len(tb)); L.check()
        u0 = lib.b2_pool_bytes_used(rt)
        lib.b2_write(tree, x, vp(coords[:n].copy()), n, vp(vals[:n].copy()), 0)
        e = lib.b2_check_runtime_error(rt, buf, 256)
        act = np.zeros(4096, np.uint8)
        lib.b2_is_active(tree, p, vp(coords12(np.arange(4096, dtype=np.int32)[:, None] * 65536)), 4096, vp(act)); L.check()
        print("pool", pool_mb, "n", n, "err", e, buf.value[:40] if e else b"", "used", u0, "->", lib.b2_pool_bytes_used(rt), "n_act", int(act.sum()), flush=True)
        lib.b2_destroy_tree(tree); lib.ti_destroy_runtime(rt)
