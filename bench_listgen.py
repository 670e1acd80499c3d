#!/usr/bin/env python
"""bench_listgen.py -- config[3] of BASELINE.json: listgen / activation microbenchmark on
ti.root.pointer(ijk,32).bitmasked(ijk,8).dense(ijk,4).place(f32)  (1024^3, 16.7M blocks) at 1 % and
10 % block occupancy, uniform-random and clustered, incl. deactivate_all + GC.

Prints one JSON line per case:
  listgen_blocks_per_s   leaf Elements emitted / time of (clear + listgen of all 3 levels)
  cycle_blocks_per_s     active blocks / time of (deactivate_all + GC + activate + listgen)
  roofline               listgen algorithmic bytes = 64 B/Element + 8 B/pointer slot + 4 B/mask word
  launch_floor_ms        three empty kernels back to back on the same stream (listgen = 3 launches: the latency floor)
  cpu_baseline           the reference's own runtime.cpp (oracle/_ref, x64: serial by construction, runtime.cpp:1299-1302)
                         -- or the C restatement where _ref is absent -- generating the same lists on one host core
  parity                 leaf element list == the CPU side's as sorted sets
bench.py imports run_case() for its `listgen` block.
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

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


def l1024_tree(n):
    from _desc import TreeBuilder
    tb = TreeBuilder()
    p = tb.pointer(0, "ijk", n // 32)
    b = tb.bitmasked(p, "ijk", 8)
    d = tb.dense(b, "ijk", 4)
    tb.place(d, 4)
    return tb, p, b, d


def cpu_listgen(blocks, n, reps=3):
    """Serial CPU listgen of the same tree and block set: (ms per clear+listgen of the 3 levels, sorted leaf list, kind)."""
    from _desc import coords12, i32p, load_oracle, load_ref_runtime
    ref = load_ref_runtime()
    lib, pre, kind = (ref, "refrt_", "reference") if ref is not None else (load_oracle(), "orc_", "port")
    tb, p, b, d = l1024_tree(n)
    assert getattr(lib, pre + "create")(tb.array(), len(tb)) == 0
    c = coords12(blocks * 4)
    getattr(lib, pre + "activate")(d, i32p(c), len(c))
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        cnt = getattr(lib, pre + "listgen")(d)
        ts.append(time.perf_counter() - t0)
    out = np.zeros((max(cnt, 1), 14), np.int32)
    getattr(lib, pre + "get_list")(d, i32p(out), cnt)
    out = out[:cnt]
    from _desc import TreeBuilder
    getattr(lib, pre + "create")(TreeBuilder().array(), 1)  # release the host-side tree
    return float(np.median(ts)) * 1e3, out[np.lexsort(out.T[::-1])], kind


def run_case(lib, L, rt, n, kind, frac, reps=10, cpu=True, peak=None):
    import torch
    from _desc import coords12
    nb = n // 4
    tb, p, b, d = l1024_tree(n)
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
    ms_listgen = timeit(lambda: lib.b2_listgen_async(tree, d), reps)
    launches_listgen = (lib.b2_launch_count(rt) - l0) // (reps + 1)
    # three empty launches on the same stream: what any 3-kernel sequence costs at least (SURVEY 8d asks for this floor)
    z = torch.zeros(1, device="cuda")
    ms_floor = timeit(lambda: (z.add_(0), z.add_(0), z.add_(0)), reps)

    def cycle():
        lib.b2_deactivate_all(tree, p)
        lib.b2_activate(tree, d, cptr, nblk)
        lib.b2_listgen_async(tree, d)
    ms_cycle = timeit(cycle, max(3, reps // 2))
    ms_deact = timeit(lambda: (lib.b2_deactivate_all(tree, p), lib.b2_activate(tree, d, cptr, nblk))[0], 3)
    assert lib.b2_listgen(tree, d) == nblk
    st = np.zeros(4, np.int32)
    lib.b2_alloc_stats(tree, p, st.ctypes.data_as(L.i32p))
    L.check()
    buf = C.create_string_buffer(256)
    assert lib.b2_check_runtime_error(rt, buf, 256) == 0, buf.value
    alg = 64 * nblk + (n // 32) ** 3 * 8 + n_ptr * (64 + 16 * 4) + ((n // 32) ** 3 // 1024 + 1) * 64
    if peak is None:
        pk = os.path.join(REPO, "MEASURED_PEAKS.json")
        peak = json.load(open(pk))["hbm_gbs"] if os.path.exists(pk) else 6650.0
    line = {"metric": "listgen blocks/s", "case": f"{kind} {frac:.0%}", "n_grid": n, "active_blocks": nblk,
            "active_pointer_cells": n_ptr, "listgen_ms": ms_listgen, "listgen_blocks_per_s": nblk / (ms_listgen * 1e-3),
            "launch_floor_ms": ms_floor, "listgen_launches": int(launches_listgen),
            "cycle_ms": ms_cycle, "cycle_blocks_per_s": nblk / (ms_cycle * 1e-3), "deactivate_gc_activate_ms": ms_deact,
            "nodes_allocated": int(st[0]),
            "roofline": {"bound": "hbm", "alg_bytes": alg, "achieved": alg / (ms_listgen * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": alg / (ms_listgen * 1e-3) / 1e9 / peak}}
    if cpu:
        cpu_ms, cpu_list, cpu_kind = cpu_listgen(blocks, n)
        got = np.zeros((nblk, 14), np.int32)
        lib.b2_get_list(tree, d, got.ctypes.data_as(C.c_void_p), nblk)
        L.check()
        got = got[np.lexsort(got.T[::-1])]
        line["parity"] = bool(np.array_equal(got, cpu_list))
        line["cpu_baseline"] = {"value": nblk / (cpu_ms * 1e-3), "unit": "blocks/s", "cores": 1, "kind": cpu_kind, "listgen_ms": cpu_ms,
                                "sample": "same tree and block set; the reference's x64 listgen is serial (runtime.cpp:1299-1302)"}
        assert line["parity"], "leaf element list differs from the CPU side"
    lib.b2_destroy_tree(tree)
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--n-grid", type=int, default=1024)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    from taichi_b200 import _lib as L
    lib = L.lib
    lib.b2_set_pool_size(C.c_uint64(10 << 30))
    rt = lib.ti_create_runtime(L.TI_ARCH_CUDA, 0)
    L.check()
    for frac in (0.01, 0.10):
        for kind in ("uniform", "clustered"):
            print(json.dumps(run_case(lib, L, rt, args.n_grid, kind, frac, args.reps, cpu=not args.no_cpu)), flush=True)


if __name__ == "__main__":
    main()
