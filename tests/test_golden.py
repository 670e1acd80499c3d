"""Golden fixtures produced by the REFERENCE runtime (tests/golden/snode_ref_*.json, generator:
tests/golden/make_snode_golden.py over oracle/_ref = the reference's runtime.cpp compiled for x64).
 * CPU: the oracle restatement (oracle/snode_oracle.c) reproduces every fixture: node sizes, element lists of all
   levels, struct-for coordinate sets, allocator counts through activate -> deactivate_all + GC -> re-activate.
 * GPU: the CUDA library reproduces the same through the C ABI (lists / coordinate sets bit-exact as sorted sets,
   allocator counts equal; node sizes are NOT compared: the CUDA layout rounds nodes to 128 B by design).
Nothing here reads /root/reference or needs oracle/_ref at run time."""
import glob
import hashlib
import json
import os

import numpy as np
import pytest

from _b2 import B2Impl, have_gpu
from golden.make_snode_golden import NAMES, run, scenario
from test_oracle_vs_reference import Impl, ORC

HERE = os.path.dirname(os.path.abspath(__file__))


def load(name):
    return json.load(open(os.path.join(HERE, "golden", f"snode_ref_{name}.json")))


def test_fixtures_are_committed():
    assert sorted(os.path.basename(p) for p in glob.glob(os.path.join(HERE, "golden", "snode_ref_*.json"))) == \
        sorted(f"snode_ref_{n}.json" for n in NAMES)


def check(got, want, node_sizes):
    assert got["coords"] == want["coords"] and got["again"] == want["again"]  # same seeded inputs
    if node_sizes:
        assert got["node_size"] == want["node_size"]
    for stage in ("after_activate", "after_deactivate_all", "after_reactivate"):
        g, w = got[stage], want[stage]
        assert g["lists"].keys() == w["lists"].keys()
        for lvl in w["lists"]:
            assert g["lists"][lvl] == w["lists"][lvl], f"{stage}: element list of SNode {lvl}"
        assert g["visited_n"] == w["visited_n"] and g["visited_sha1"] == w["visited_sha1"], stage
        assert g["stats"] == w["stats"], f"{stage}: allocator counts"


@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_reference_fixture(name):
    got = run(Impl(ORC, "orc_"), name, ORC.orc_node_size)
    check(got, load(name), node_sizes=True)


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_reproduces_reference_fixture(name):
    if not have_gpu():
        pytest.skip("no CUDA device")
    g = B2Impl()
    try:
        got = run(g, name, lambda q: 0)
        check(got, load(name), node_sizes=False)
        code, msg = g.runtime_error()
        assert code == 0, msg
    finally:
        g.close()


def test_fixture_hash_is_of_sorted_coordinates():
    d = load("non_pot_bitmasked_2d")["after_activate"]
    v = np.asarray(d["visited"], np.int32)
    assert np.array_equal(v, v[np.lexsort(v.T[::-1])])
    assert hashlib.sha1(v.tobytes()).hexdigest() == d["visited_sha1"]
