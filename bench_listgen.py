#!/usr/bin/env python
"""bench_listgen.py -- config[3] of BASELINE.json: listgen / activation microbenchmark on
ti.root.pointer(ijk,32).bitmasked(ijk,8).dense(ijk,4).place(f32)  (1024^3, 16.7M blocks) at 1 % and
10 % block occupancy, uniform-random and clustered, incl. deactivate_all + GC.

Prints one JSON line per case:
  listgen_blocks_per_s   leaf Elements emitted / time of (clear + listgen of all 3 levels)
  cycle_blocks_per_s     active blocks / time of (deactivate_all + GC + activate + listgen)
  roofline: listgen algorithmic bytes = 64 B/Element + 8 B/pointer slot + 4 B/mask word
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))


def block_set(kind, frac, nb, seed=4):
    rng = np.random.default_rng(seed)
    total = nb ** 3
    n = int(round(total * frac))
    if kind == "uniform":
        ids = rng.choice(total, size=n, replace=False)
        b = np.stack([ids // (nb * nb), (ids // nb) % nb, ids % nb], 1)
    else:  # clustered ball with the same block count
        r = (3 * n / (4 * np.pi)) ** (1 / 3)
        g = np.arange(nb) - nb / 2 + 0.5
        X, Y, Z = np.meshgrid(g, g, g, indexing="ij")
        d2 = (X * X + Y * Y + Z * Z).ravel()
        ids = np.argpartition(d2, n)[:n]
        b = np.stack([ids // (nb * nb), (ids // nb) % nb, ids % nb], 1)
    return b.astype(np.int32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--n-grid", type=int, default=1024)
    args = ap.parse_args()
    import torch
    from _desc import TreeBuilder, coords12
    from taichi_b200 import _lib as L
    lib = L.lib
    lib.b2_set_pool_size(C.c_uint64(10 << 30))
    rt = lib.ti_create_runtime(L.TI_ARCH_CUDA, 0)
    L.check()
    n = args.n_grid
    nb = n // 4
    tb = TreeBuilder()
    p = tb.pointer(0, "ijk", n // 32)
    b = tb.bitmasked(p, "ijk", 8)
    d = tb.dense(b, "ijk", 4)
    x = tb.place(d, 4)
    peak = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(REPO, "MEASURED_PEAKS.json")) else 6650.0
    for frac in (0.01, 0.10):
        for kind in ("uniform", "clustered"):
            arr = tb.array()
            tree = lib.b2_create_tree(rt, C.cast(arr, C.POINTER(L.B2SNodeDesc)), len(tb))
            L.check()
            blocks = block_set(kind, frac, nb)
            coords = torch.from_numpy(coords12(blocks * 4)).cuda()
            cptr = C.c_void_p(coords.data_ptr())
            nblk = len(blocks)

            def ev():
                return torch.cuda.Event(enable_timing=True)

            def timeit(fn, reps):
                fn()
                torch.cuda.synchronize()
                ts = []
                for _ in range(reps):
                    e0, e1 = ev(), ev()
                    e0.record()
                    fn()
                    e1.record()
                    e1.synchronize()
                    ts.append(e0.elapsed_time(e1))
                return float(np.median(ts))

            lib.b2_activate(tree, d, cptr, nblk)
            L.check()
            assert lib.b2_listgen(tree, d) == nblk
            n_ptr = int(lib.b2_listgen(tree, b))       # active pointer cells = bitmasked containers
            l0 = lib.b2_launch_count(rt)
            ms_listgen = timeit(lambda: lib.b2_listgen_async(tree, d), args.reps)

            def cycle():
                lib.b2_deactivate_all(tree, p)
                lib.b2_activate(tree, d, cptr, nblk)
                lib.b2_listgen_async(tree, d)
            ms_cycle = timeit(cycle, max(3, args.reps // 2))
            ms_deact = timeit(lambda: (lib.b2_deactivate_all(tree, p), lib.b2_activate(tree, d, cptr, nblk))[0], 3) - 0.0
            launches = lib.b2_launch_count(rt) - l0
            assert lib.b2_listgen(tree, d) == nblk
            st = np.zeros(4, np.int32)
            lib.b2_alloc_stats(tree, p, st.ctypes.data_as(L.i32p))
            L.check()
            buf = C.create_string_buffer(256)
            assert lib.b2_check_runtime_error(rt, buf, 256) == 0, buf.value
            alg = 64 * nblk + (n // 32) ** 3 * 8 + n_ptr * (64 + 16 * 4) + ((n // 32) ** 3 // 1024 + 1) * 64
            line = {"metric": "listgen blocks/s", "case": f"{kind} {frac:.0%}", "n_grid": n, "active_blocks": nblk,
                    "active_pointer_cells": n_ptr, "listgen_ms": ms_listgen, "listgen_blocks_per_s": nblk / (ms_listgen * 1e-3),
                    "cycle_ms": ms_cycle, "cycle_blocks_per_s": nblk / (ms_cycle * 1e-3), "deactivate_gc_activate_ms": ms_deact,
                    "nodes_allocated": int(st[0]), "gpu_launches": int(launches),
                    "roofline": {"bound": "hbm", "alg_bytes": alg, "achieved": alg / (ms_listgen * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                 "frac": alg / (ms_listgen * 1e-3) / 1e9 / peak}}
            print(json.dumps(line), flush=True)
            lib.b2_destroy_tree(tree)


if __name__ == "__main__":
    main()
