"""Slab decomposition on the GPU: every CUDA kernel of the exchange path (pack/unpack block layers,
migrant extraction/insertion, compact export) through the C ABI.
 * several slabs emulated on ONE GPU in one process (lockstep phases, messages handed over by copy);
 * with >= 2 GPUs: the same over torch.distributed / NCCL, one process per GPU.
The decomposed result must match the single-domain CUDA solver and the CPU oracle."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from _b2 import have_gpu, runtime
from test_gpu_mpm import oracle_mpm3d

pytestmark = pytest.mark.gpu
N_GRID, DT, STEPS, N = 64, 2e-4, 16, 30000


def cloud():
    rng = np.random.default_rng(8)
    x = (rng.random((N, 3)) * np.array([0.6, 0.3, 0.3]) + np.array([0.2, 0.3, 0.3])).astype(np.float32)
    v = np.tile(np.array([3.0, -1.0, 0.5], np.float32), (N, 1))
    v[N // 2:, 0] *= -1
    return x, v, np.zeros((N, 3, 3), np.float32), np.ones(N, np.float32), np.arange(N, dtype=np.int32)


def get_rt():
    import taichi_b200 as tb
    r = tb.Runtime.__new__(tb.Runtime)
    r.handle = runtime()
    r.memories = []
    return r


def compare(parts, ref):
    x = np.concatenate([p[0] for p in parts])
    v = np.concatenate([p[1] for p in parts])
    J = np.concatenate([p[3] for p in parts])
    ids = np.concatenate([p[4] for p in parts])
    assert len(ids) == N and len(np.unique(ids)) == N
    o = np.argsort(ids)
    xs, vs, Cs, Js = ref
    assert np.allclose(x[o], xs, rtol=0, atol=2e-6)
    assert np.allclose(v[o], vs, rtol=1e-4, atol=2e-4)
    assert np.allclose(J[o], Js, rtol=1e-4)


@pytest.mark.parametrize("world", [2, 3])
def test_emulated_slabs_on_one_gpu(world):
    if not have_gpu():
        pytest.skip("no CUDA device")
    from taichi_b200 import multigpu as mg
    rt = get_rt()
    x, v, Cm, J, ids = cloud()
    owner = mg.partition_particles(x, N_GRID, world)
    slabs, drivers = [], []
    for r in range(world):
        lo, hi = mg.slab_bounds(N_GRID // 4, world, r)
        s = mg.CudaSlab(rt, N_GRID, N, DT, lo, hi)
        sel = owner == r
        s.set_particles(x[sel], v[sel], Cm[sel], J[sel], ids[sel])
        slabs.append(s)
        drivers.append(mg.SlabDriver(s, r, world, cap_blocks=1024, cap_particles=8192))
    before = [len(s.export(N)[4]) for s in slabs]
    mg.substep_local(drivers, STEPS)
    parts = [s.export(N) for s in slabs]
    assert [len(p[4]) for p in parts] != before                   # particles migrated
    for s, p in zip(slabs, parts):                                # ownership invariant
        if len(p[0]):
            bx = (p[0][:, 0] * np.float32(N_GRID) - np.float32(0.5)).astype(np.int32) >> 2
            assert ((bx >= s.lo) & (bx < s.hi)).all()
    xs, vs, Cs, Js, *_ = oracle_mpm3d(N_GRID, x, v, Cm, J, DT, STEPS)
    compare(parts, (xs, vs, Cs, Js))
    buf = C.create_string_buffer(256)
    from taichi_b200 import _lib as L
    assert L.lib.b2_check_runtime_error(rt.handle, buf, 256) == 0, buf.value
    for s in slabs:
        s.destroy()


@pytest.mark.parametrize("world", [2, 3])
def test_native_driver_loopback(world):
    """The production C++ driver (b2_slab.cu: kernels + three exchanges per substep, message tuning, whole-domain
    diagnostics) at world = 2 and 3 on ONE GPU: the loopback communicator replaces every ncclSend/ncclRecv pair by a
    device-to-device copy into the peer's receive buffer, everything else is the code an NCCL job runs.  The decomposed
    result must match the CPU oracle (and so the single-domain solver) particle by particle."""
    if not have_gpu():
        pytest.skip("no CUDA device")
    from taichi_b200 import multigpu as mg
    from taichi_b200 import _lib as L
    rt = get_rt()
    x, v, Cm, J, ids = cloud()
    owner = mg.partition_particles(x, N_GRID, world)
    slabs = []
    for r in range(world):
        lo, hi = mg.slab_bounds(N_GRID // 4, world, r)
        s = mg.CudaSlab(rt, N_GRID, N, DT, lo, hi)
        sel = owner == r
        s.set_particles(x[sel], v[sel], Cm[sel], J[sel], ids[sel])
        slabs.append(s)
    g = mg.LoopbackSlabGroup(slabs, cap_blocks=1024, cap_particles=8192)
    before = [len(s.export(N)[4]) for s in slabs]
    g.substep(STEPS // 2)
    g.tune()                          # messages shrink to 2x the observed counts, then the run continues
    g.substep(STEPS - STEPS // 2)
    g.tune()                          # a second tune never shrinks a message
    parts = [s.export(N) for s in slabs]
    assert [len(p[4]) for p in parts] != before                   # particles migrated
    for s, p in zip(slabs, parts):                                # ownership invariant
        bx = (p[0][:, 0] * np.float32(N_GRID) - np.float32(0.5)).astype(np.int32) >> 2
        assert ((bx >= s.lo) & (bx < s.hi)).all()
    xs, vs, Cs, Js, *_ = oracle_mpm3d(N_GRID, x, v, Cm, J, DT, STEPS)
    compare(parts, (xs, vs, Cs, Js))
    d = g.diagnostics()
    local = np.stack([s.diagnostics() for s in slabs])
    assert d[0] == N and np.allclose(d[:6], local[:, :6].sum(0), rtol=1e-12) and np.array_equal(d[6:], local[:, 6:].max(0))
    buf = C.create_string_buffer(256)
    assert L.lib.b2_check_runtime_error(rt.handle, buf, 256) == 0, buf.value
    g.destroy()
    for s in slabs:
        s.destroy()


def test_full_migrant_message_keeps_the_particles():
    """A migrant message that is too small must not destroy particles: the bin that does not fit stays where
    it is, the published count covers written records only, the overflow raises the device error flag, and once the
    capacity is there the particles travel."""
    if not have_gpu():
        pytest.skip("no CUDA device")
    from taichi_b200 import multigpu as mg
    from taichi_b200 import _lib as L
    rt = get_rt()
    x, v, Cm, J, ids = cloud()
    world = 2
    owner = mg.partition_particles(x, N_GRID, world)
    slabs = []
    for r in range(world):
        lo, hi = mg.slab_bounds(N_GRID // 4, world, r)
        s = mg.CudaSlab(rt, N_GRID, N, DT, lo, hi)
        sel = owner == r
        s.set_particles(x[sel], v[sel], Cm[sel], J[sel], ids[sel])
        slabs.append(s)
    g = mg.LoopbackSlabGroup(slabs, cap_blocks=1024, cap_particles=1)    # one record: far too few for this cloud
    g.substep(STEPS)
    buf = C.create_string_buffer(256)
    assert L.lib.b2_check_runtime_error(rt.handle, buf, 256) != 0 and b"overflow" in buf.value.lower()
    parts = [s.export(N) for s in slabs]
    all_ids = np.concatenate([p[4] for p in parts])
    assert len(all_ids) == N and len(np.unique(all_ids)) == N          # nobody lost, nobody duplicated
    g.destroy()
    for s in slabs:
        s.destroy()


def test_native_driver_world1_equals_plain_substep():
    """b2_slab_substep with world = 1 is the plain substep sequence (no NCCL touched)."""
    if not have_gpu():
        pytest.skip("no CUDA device")
    from taichi_b200 import multigpu as mg
    rt = get_rt()
    x, v, Cm, J, ids = cloud()
    s = mg.CudaSlab(rt, N_GRID, N, DT, 0, N_GRID // 4)
    s.set_particles(x, v, Cm, J, ids)
    d = mg.NativeSlabDriver(s, 0, 1)
    d.substep(STEPS)
    d.tune()
    xs, vs, Cs, Js, *_ = oracle_mpm3d(N_GRID, x, v, Cm, J, DT, STEPS)
    compare([s.export(N)], (xs, vs, Cs, Js))
    d.destroy()
    s.destroy()


def test_export_async_equals_blocking_export():
    """The pipelined form of the compact export (device staging + copy stream, pinned host arrays) returns the
    same record set and count as the blocking one."""
    if not have_gpu():
        pytest.skip("no CUDA device")
    import torch
    from taichi_b200 import multigpu as mg
    rt = get_rt()
    x, v, Cm, J, ids = cloud()
    s = mg.CudaSlab(rt, N_GRID, 2 * N, DT, 0, N_GRID // 4)
    s.set_particles(x, v, Cm, J, ids)
    d = mg.NativeSlabDriver(s, 0, 1)
    d.substep(4)
    ref = s.export(2 * N)
    cap = N + N // 4
    hx, hv = torch.empty(cap, 3).pin_memory(), torch.empty(cap, 3).pin_memory()
    hi, cnt = torch.empty(cap, dtype=torch.int32).pin_memory(), torch.zeros(1, dtype=torch.int64).pin_memory()
    s.export_async(cap, cnt, x=hx, v=hv, ids=hi)
    s.wait_io()
    n = int(cnt.item())
    assert n == len(ref[4]) == N
    o, r = np.argsort(hi.numpy()[:n]), np.argsort(ref[4])
    assert np.array_equal(hi.numpy()[:n][o], ref[4][r])
    assert np.array_equal(hx.numpy()[:n][o], ref[0][r]) and np.array_equal(hv.numpy()[:n][o], ref[1][r])
    d.destroy()
    s.destroy()


def test_diagnostics_vs_numpy():
    """k_mpm_diag (fp64 sums from thread-local accumulators + warp shuffles, maxima as ordered bit patterns)
    against NumPy on the exported particles; world = 1 slab driver returns the same vector."""
    if not have_gpu():
        pytest.skip("no CUDA device")
    from taichi_b200 import multigpu as mg
    rt = get_rt()
    x, v, Cm, J, ids = cloud()
    s = mg.CudaSlab(rt, N_GRID, N, DT, 0, N_GRID // 4)
    s.set_particles(x, v, Cm, J, ids)
    d = mg.NativeSlabDriver(s, 0, 1)
    d.substep(5)
    xe, ve, Ce, Je, ie = s.export(N)
    m = (1.0 / N_GRID * 0.5) ** 2
    v64 = ve.astype(np.float64)
    want = np.array([len(ie), *(np.float32(m) * v64.sum(0)), 0.5 * np.float32(m) * (v64 * v64).sum(), Je.astype(np.float64).sum(),
                     np.sqrt((v64 * v64).sum(1)).max(), np.abs(Je - np.float32(1)).max()])
    got = s.diagnostics()
    assert got[0] == N
    assert np.allclose(got[:6], want[:6], rtol=1e-9, atol=1e-12)
    assert np.allclose(got[6:], want[6:], rtol=1e-6)
    again = d.diagnostics()                    # fp64 atomics: the order of the adds differs from launch to launch
    assert np.allclose(again[:6], got[:6], rtol=1e-12, atol=0) and np.array_equal(again[6:], got[6:])
    d.destroy()
    s.destroy()


def _nccl_worker(rank, world, port, out, native):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch
    import torch.distributed as dist
    import taichi_b200 as tb
    from taichi_b200 import multigpu as mg
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    rt = tb.Runtime(rank, device_memory_GB=1)
    x, v, Cm, J, ids = cloud()
    owner = mg.partition_particles(x, N_GRID, world)
    lo, hi = mg.slab_bounds(N_GRID // 4, world, rank)
    s = mg.CudaSlab(rt, N_GRID, N, DT, lo, hi)
    sel = owner == rank
    s.set_particles(x[sel], v[sel], Cm[sel], J[sel], ids[sel])
    if native:
        d = mg.NativeSlabDriver(s, rank, world, cap_blocks=1024, cap_particles=8192)
        d.substep(STEPS // 2)
        d.tune()                      # shrink the messages to 2x the observed counts, then continue
        d.substep(STEPS - STEPS // 2)
    else:
        d = mg.SlabDriver(s, rank, world, cap_blocks=1024, cap_particles=8192)
        mg.substep_distributed(d, STEPS)
    torch.cuda.synchronize()
    buf = C.create_string_buffer(256)
    assert tb._lib.lib.b2_check_runtime_error(rt.handle, buf, 256) == 0, buf.value
    # whole-domain diagnostics: ncclAllReduce from C++ (native) / torch.distributed all_reduce (python driver)
    diag = d.diagnostics() if native else mg.diagnostics_distributed(s)
    gathered = [None] * world
    dist.all_gather_object(gathered, tuple(s.export(N)) + (diag, s.diagnostics()))
    if rank == 0:
        torch.save(gathered, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("native", [True, False])
def test_nccl_two_gpus(tmp_path, native):
    import torch
    if not have_gpu() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    out = str(tmp_path / "g.pt")
    mp.spawn(_nccl_worker, args=(2, 29700 + os.getpid() % 1000 + int(native), out, native), nprocs=2, join=True)
    x, v, Cm, J, ids = cloud()
    xs, vs, Cs, Js, *_ = oracle_mpm3d(N_GRID, x, v, Cm, J, DT, STEPS)
    parts = torch.load(out, weights_only=False)
    compare([p[:5] for p in parts], (xs, vs, Cs, Js))
    # every rank got the same reduced vector = SUM / MAX of the slabs' local vectors
    assert np.array_equal(parts[0][5], parts[1][5])
    local = np.stack([p[6] for p in parts])
    assert np.allclose(parts[0][5][:6], local[:, :6].sum(0), rtol=1e-12, atol=1e-18)
    assert np.array_equal(parts[0][5][6:], local[:, 6:].max(0))
    assert parts[0][5][0] == N
