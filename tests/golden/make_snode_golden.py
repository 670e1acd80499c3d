#!/usr/bin/env python
"""Generates tests/golden/snode_ref_*.json from the REFERENCE runtime itself: oracle/_ref/libref_runtime.so is the
reference's own runtime.cpp (+ node_*.h) compiled for x64 by oracle/Makefile from /root/reference, driven through
oracle/ref_runtime_harness.cpp.  Each fixture records, for a seeded scenario, what the reference produced:
node sizes, the element list of every level (sorted records: pcoord[12], lo, hi), the struct-for coordinate set
of the leaf, allocator counts before / after deactivate_all + GC, and the sets after re-activating a subset.
Run here (needs the built _ref):  python tests/golden/make_snode_golden.py
The fixtures travel to the GPU box; tests/test_golden.py checks the oracle (CPU) and the CUDA library (GPU)
against them, so parity with the reference does not depend on _ref being present there."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from _desc import POINTER, TreeBuilder, load_ref_runtime  # noqa: E402
from test_oracle_vs_reference import Impl  # noqa: E402


def scenario(name):
    """(TreeBuilder, leaf id, pointer ids (outermost first), coords to activate, coords to re-activate)"""
    rng = np.random.default_rng(sum(map(ord, name)))
    tb = TreeBuilder()
    if name == "pointer_dense_2d":
        p = tb.pointer(0, "ij", (8, 8))
        leaf = tb.dense(p, "ij", (4, 4))
        ext = (32, 32)
    elif name == "pointer_bitmasked_dense_3d":   # the L512 layout family, small
        p = tb.pointer(0, "ijk", (4, 4, 4))
        b = tb.bitmasked(p, "ijk", (4, 4, 4))
        leaf = tb.dense(b, "ijk", (2, 2, 2))
        ext = (32, 32, 32)
    elif name == "pointer_pointer_dense_3d":     # test_gc.py:17-18 / test_mpm_particle_list.py:19-28 idiom
        p = tb.pointer(0, "ijk", (4, 4, 4))
        b = tb.pointer(p, "ijk", (2, 2, 2))
        leaf = tb.dense(b, "ijk", (4, 4, 4))
        ext = (32, 32, 32)
    elif name == "non_pot_bitmasked_2d":         # non power-of-two shapes (test_struct_for_non_pot.py)
        p = tb.bitmasked(0, "ij", (3, 6))
        leaf = tb.bitmasked(p, "ij", (5, 3))
        ext = (15, 18)
    elif name == "big_element_split_1d":         # containers above 1024 cells are split into several Elements
        p = tb.pointer(0, "i", 4)
        leaf = tb.dense(p, "i", 4096)
        ext = (16384,)
    else:
        raise KeyError(name)
    tb.place(leaf, 4)
    n = {"big_element_split_1d": 9}.get(name, 150)
    coords = np.stack([rng.integers(0, e, n) for e in ext], axis=1)
    again = coords[rng.permutation(n)[: max(1, n // 5)]]
    return tb, leaf, coords, again


NAMES = ["pointer_dense_2d", "pointer_bitmasked_dense_3d", "pointer_pointer_dense_3d", "non_pot_bitmasked_2d", "big_element_split_1d"]


def snapshot(impl, tb, leaf):
    out = {"lists": {}, "stats": {}}
    q = leaf
    while q > 0:
        out["lists"][str(q)] = impl.list(q).tolist()
        if tb.nodes[q]["type"] == POINTER:
            out["stats"][str(q)] = impl.stats(q).tolist()
        q = tb.nodes[q]["parent"]
    v = impl.visited(leaf)
    out["visited_n"] = int(len(v))
    out["visited_sha1"] = hashlib.sha1(np.ascontiguousarray(v, dtype=np.int32).tobytes()).hexdigest()  # sorted (n, 12) i32
    if len(v) <= 1500:
        out["visited"] = v.tolist()
    return out


def run(impl, name, node_size):
    tb, leaf, coords, again = scenario(name)
    impl.create(tb)
    rec = {"name": name, "leaf": leaf, "coords": coords.tolist(), "again": again.tolist(),
           "node_size": [int(node_size(q)) for q in range(len(tb))]}
    impl.activate(leaf, coords)
    rec["after_activate"] = snapshot(impl, tb, leaf)
    top = leaf
    while tb.nodes[top]["parent"] > 0:
        top = tb.nodes[top]["parent"]
    impl.deactivate_all(top)      # snode.deactivate_all: children first, GC after (snode.py:286-304, offload.cpp:671-695)
    rec["after_deactivate_all"] = snapshot(impl, tb, leaf)
    impl.activate(leaf, again)
    rec["after_reactivate"] = snapshot(impl, tb, leaf)
    return rec


if __name__ == "__main__":
    REF = load_ref_runtime()
    assert REF is not None, "build oracle/_ref first: make -C oracle ref"
    for name in NAMES:
        rec = run(Impl(REF, "refrt_"), name, REF.refrt_node_size)
        rec["generator"] = "reference runtime.cpp (x64) via oracle/_ref/libref_runtime.so"
        path = os.path.join(HERE, f"snode_ref_{name}.json")
        json.dump(rec, open(path, "w"), separators=(",", ":"))
        print(name, os.path.getsize(path), "bytes; leaf elements", len(rec["after_activate"]["lists"][str(rec["leaf"])]),
              "visited", rec["after_activate"]["visited_n"])
