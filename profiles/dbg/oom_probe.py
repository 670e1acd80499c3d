import sys, os, ctypes as C
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
from taichi_b200 import _lib as L
from _b2 import vp
from _desc import coords12, TreeBuilder
lib = L.lib
coords = coords12(np.arange(4096, dtype=np.int32)[:, None] * 65536 + 7)
cells = coords12(np.arange(4096, dtype=np.int32))
vals = np.full(4096, 1234567, np.int32)
buf = C.create_string_buffer(256)
for pool_mb in (48,):
    for n in (1, 64, 300, 4096):
        lib.b2_set_pool_size(C.c_uint64(pool_mb << 20))
        rt = lib.ti_create_runtime(L.TI_ARCH_CUDA, 0); L.check()
        tb = TreeBuilder(); p = tb.pointer(0, "i", 4096); d = tb.dense(p, "i", 65536); x = tb.place(d, 4)
        arr = tb.array()
        tree = lib.b2_create_tree(rt, C.cast(arr, C.POINTER(L.B2SNodeDesc)), len(tb)); L.check()
        cn, vn = coords[:n].copy(), vals[:n].copy()
        lib.b2_write(tree, x, vp(cn), n, vp(vn), 0)
        e = lib.b2_check_runtime_error(rt, buf, 256)
        act = np.zeros(4096, np.uint8)
        lib.b2_is_active(tree, p, vp(cells), 4096, vp(act)); L.check()
        got = np.zeros(4096, np.int32)
        lib.b2_read(tree, x, vp(coords), 4096, vp(got)); L.check()
        st = np.zeros(4, np.int32); lib.b2_alloc_stats(tree, p, st.ctypes.data_as(L.i32p))
        print("n", n, "err", e, "used", lib.b2_pool_bytes_used(rt), "n_act", int(act.sum()), "first act", np.flatnonzero(act)[:6], "nonzero reads", int((got != 0).sum()), "stats", st, flush=True)
        lib.b2_destroy_tree(tree); lib.ti_destroy_runtime(rt)
