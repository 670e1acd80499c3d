"""Host logic of the slab decomposition (taichi_b200/multigpu.py) on the CPU: the per-slab compute
is the NumPy stand-in (tests/_np_slab.py), the transport is (a) the in-process lockstep runner and
(b) torch.distributed with the gloo backend, world_size 2.  The decomposed run must reproduce the
single-domain run (same arithmetic; only the order of fp32 grid additions differs)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from _np_slab import NumpySlab
from taichi_b200 import multigpu as mg

N_GRID, DT, STEPS, N = 32, 4e-4, 12, 2500


def cloud():
    rng = np.random.default_rng(4)
    x = (rng.random((N, 3)) * np.array([0.5, 0.3, 0.3]) + np.array([0.25, 0.3, 0.3])).astype(np.float32)
    v = np.tile(np.array([2.5, -1.0, 0.5], np.float32), (N, 1))   # fast along i: forces migration across slabs
    v[N // 2:, 0] *= -1
    return x, v, np.zeros((N, 3, 3), np.float32), np.ones(N, np.float32), np.arange(N, dtype=np.int32)


def run_single():
    s = NumpySlab(N_GRID, DT, 0, N_GRID // 4)
    s.set_particles(*cloud())
    d = mg.SlabDriver(s, 0, 1, cap_blocks=64, cap_particles=64)
    mg.substep_local([d], STEPS)
    return s.x, s.v, s.J, s.ids


def make_rank(rank, world):
    x, v, Cm, J, ids = cloud()
    owner = mg.partition_particles(x, N_GRID, world)
    lo, hi = mg.slab_bounds(N_GRID // 4, world, rank)
    s = NumpySlab(N_GRID, DT, lo, hi)
    sel = owner == rank
    s.set_particles(x[sel], v[sel], Cm[sel], J[sel], ids[sel])
    return s, mg.SlabDriver(s, rank, world, cap_blocks=128, cap_particles=1024)


def check_against_single(parts):
    x = np.concatenate([p[0] for p in parts])
    v = np.concatenate([p[1] for p in parts])
    J = np.concatenate([p[2] for p in parts])
    ids = np.concatenate([p[3] for p in parts])
    assert len(ids) == N and len(np.unique(ids)) == N          # nobody lost, nobody duplicated
    o = np.argsort(ids)
    xs, vs, Js, ids_s = run_single()
    os_ = np.argsort(ids_s)
    assert np.allclose(x[o], xs[os_], rtol=0, atol=2e-6)
    assert np.allclose(v[o], vs[os_], rtol=1e-4, atol=1e-4)
    assert np.allclose(J[o], Js[os_], rtol=1e-5)


def test_slab_bounds_cover_domain():
    for nb, world in ((8, 2), (8, 4), (128, 8), (256, 8), (256, 3), (12, 5)):
        b = [mg.slab_bounds(nb, world, r) for r in range(world)]
        assert b[0][0] == 0 and b[-1][1] == nb
        assert all(b[r][1] == b[r + 1][0] for r in range(world - 1))
        assert all(hi > lo for lo, hi in b)
        if nb % 8 == 0 and nb // 8 >= world:
            assert all(lo % 8 == 0 for lo, _ in b)             # pointer-cell aligned


def test_weighted_slab_bounds():
    """Slabs cut by particle count: contiguous cover, aligned to pointer cells, loads within one unit's weight of even."""
    rng = np.random.default_rng(3)
    w = np.zeros(256)
    for c in (40, 90, 130, 170, 215):
        w += 1e6 * np.exp(-0.5 * ((np.arange(256) - c) / 9.0) ** 2)
    for world in (2, 3, 4, 8):
        b = mg.weighted_slab_bounds(w, world)
        assert b[0][0] == 0 and b[-1][1] == 256
        assert all(b[r][1] == b[r + 1][0] for r in range(world - 1))
        assert all(lo % 8 == 0 and hi > lo for lo, hi in b)
        load = np.array([w[lo:hi].sum() for lo, hi in b])
        assert load.max() - w.sum() / world <= w.reshape(32, 8).sum(1).max() + 1e-6
        eq = np.array([w[slice(*mg.slab_bounds(256, world, r))].sum() for r in range(world)])
        assert load.max() <= eq.max() + 1e-6
    # cuts at single block layers balance better than cuts at pointer cells (what bench.py --slabs weighted uses)
    for world in (4, 8):
        fine = mg.weighted_slab_bounds(w, world, unit=1)
        coarse = mg.weighted_slab_bounds(w, world, unit=8)
        assert fine[0][0] == 0 and fine[-1][1] == 256 and all(fine[r][1] == fine[r + 1][0] for r in range(world - 1))
        lf = max(w[lo:hi].sum() for lo, hi in fine)
        lc = max(w[lo:hi].sum() for lo, hi in coarse)
        assert lf <= lc + 1e-6 and lf - w.sum() / world <= w.max() + 1e-6
    # degenerate: all the load in one unit still leaves every rank a slab
    w = np.zeros(64); w[3] = 5.0
    b = mg.weighted_slab_bounds(w, 4)
    assert all(hi > lo for lo, hi in b) and b[-1][1] == 64
    x = rng.random((1000, 3)).astype(np.float32)
    own = mg.partition_particles(x, 256, 4, bounds=[(0, 8), (8, 16), (16, 40), (40, 64)])
    bx = ((x[:, 0] * np.float32(256) - np.float32(0.5)).astype(np.int32)) >> 2
    assert np.array_equal(own, np.digitize(bx, [8, 16, 40]))


@pytest.mark.parametrize("world", [2, 4])
def test_lockstep_slabs_match_single_domain(world):
    ranks = [make_rank(r, world) for r in range(world)]
    moved0 = [len(s.ids) for s, _ in ranks]
    mg.substep_local([d for _, d in ranks], STEPS)
    for s, _ in ranks:                                           # ownership invariant after migration
        if len(s.x):
            bx = (s.x[:, 0] / s.dx - np.float32(0.5)).astype(np.int32) >> 2
            assert ((bx >= s.lo) & (bx < s.hi)).all()
    assert [len(s.ids) for s, _ in ranks] != moved0              # particles really migrated
    check_against_single([(s.x, s.v, s.J, s.ids) for s, _ in ranks])


def _gloo_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s, d = make_rank(rank, world)
    mg.substep_distributed(d, STEPS)
    diag = mg.diagnostics_distributed(s)        # SUM / MAX all-reduce of the slabs' 0-D reductions
    gathered = [None] * world
    dist.all_gather_object(gathered, (s.x, s.v, s.J, s.ids, diag))
    if rank == 0:
        torch.save(gathered, out)
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_matches_single_domain(tmp_path):
    out = str(tmp_path / "g.pt")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_gloo_worker, args=(2, port, out), nprocs=2, join=True)
    parts = torch.load(out, weights_only=False)
    check_against_single([p[:4] for p in parts])
    # whole-domain diagnostics: every rank holds the same vector, equal to the reduction over all particles
    assert np.array_equal(parts[0][4], parts[1][4])
    v = np.concatenate([p[1] for p in parts]).astype(np.float64)
    J = np.concatenate([p[2] for p in parts])
    m = float(NumpySlab(N_GRID, DT, 0, 1).p_mass)
    want = np.array([N, *(m * v.sum(0)), 0.5 * m * (v * v).sum(), J.astype(np.float64).sum(),
                     np.sqrt((v * v).sum(1)).max(), np.abs(J - 1).max()])
    assert np.allclose(parts[0][4], want, rtol=1e-12, atol=1e-15)
