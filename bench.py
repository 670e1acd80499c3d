#!/usr/bin/env python
"""bench.py -- MPM particle-updates/s of the sparse 3-D MLS-MPM hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3|c2]

One "step" = one frame = SUBSTEPS substeps (mpm3d.py:11 `steps = 25`) of the full substep
(grid clear + P2G + listgen + grid-op + G2P with re-binning).  Workload at N=1 = config[2] of
BASELINE.json, the one north_star quotes the metric/target on: sparse 512^3
pointer->bitmasked->dense(4^3) grid, 8M particles (SURVEY 8d C3 cloud).  `value` is timed with the
particles resident in HBM; `e2e` re-uploads the particle state from pinned host memory through
the C ABI (b2_mpm_set_particles) and reads the positions back (b2_mpm_get_particles) every step.
N>1 (torchrun): one process per GPU, weak scaling (8M particles per rank).
--impl reference: the CPU restatement of the reference's kernels on a SPARSE grid (oracle/mpm_oracle.c
orc_mpm3d_sparse_substeps: blocked active set, clear / grid-op over active blocks only, multi-threaded fp32 CAS
atomics as the reference's x64 backend, the same deactivate_every life cycle) on all host cores, same `config`;
each step executes a bounded sample (CPU_SAMPLE_SUBSTEPS of the step's 25 substeps) so that the run ends in minutes.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

SUBSTEPS = 25
CPU_SAMPLE_SUBSTEPS = 2   # substeps a CPU step really executes (the metric is a rate: updates/s)
DEACTIVATE_EVERY = 1      # the reference's sparse-MPM idiom: grid.deactivate_all() every substep (test_mpm_particle_list.py:9-53)
WORKLOADS = {
    # name: (n_grid, particles, dt, layout, description)
    "c3": (512, 8 * 1024 * 1024, 2e-5, "pointer_bitmasked",
           "sparse 3D MLS-MPM pointer->bitmasked->dense(4^3) 512^3, 8M particles, ti.block_local P2G/G2P"),
    "c2": (128, 1024 * 1024, 8e-5, "dense", "mpm3d dense 128^3, 1M particles, fp32"),
}


def make_cloud(name, n_particles, seed, rank=0):
    """C3 = three balls r=0.08, ~8 particles/cell, v=(0,-1,0); C2 = mpm3d.py:87 cube."""
    rng = np.random.default_rng(seed + 1000 * rank)
    if name == "c2":
        x = (rng.random((n_particles, 3), dtype=np.float32) * 0.4 + 0.15).astype(np.float32)
        v = np.zeros((n_particles, 3), np.float32)
    else:
        centres = np.array([[0.3, 0.6, 0.5], [0.5, 0.3, 0.5], [0.7, 0.6, 0.5]], np.float32)
        d = rng.standard_normal((n_particles, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        r = 0.08 * rng.random(n_particles, dtype=np.float32) ** (1.0 / 3.0)
        x = (centres[np.arange(n_particles) % 3] + d * r[:, None]).astype(np.float32)
        v = np.tile(np.array([0.0, -1.0, 0.0], np.float32), (n_particles, 1))
    return x, v, np.zeros((n_particles, 3, 3), np.float32), np.ones(n_particles, np.float32)


def make_slab_cloud(world, rank, per_rank, n_grid):
    """SURVEY 8d C5 (weak-scaling form): every rank owns exactly ~per_rank particles: one ball in the
    middle of its slab plus half of a ball sitting ON each interior slab boundary (radius 41 cells,
    like C3's 0.08 at 512^3), so ghost-layer exchange and migration are exercised at every boundary
    and the load is even.  v = (0.5,-1,0): particles drift across slab boundaries."""
    from taichi_b200 import multigpu as mg
    nb = n_grid // 4
    radius = 41.0 / n_grid
    n_b = per_rank // 4                      # particles each side receives from one boundary ball

    def ball(seed, n, cx, cy, id0):
        rng = np.random.default_rng(seed)
        d = rng.standard_normal((n, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        r = radius * rng.random(n, dtype=np.float32) ** (1.0 / 3.0)
        return (np.array([cx, cy, 0.5], np.float32) + d * r[:, None]).astype(np.float32), np.arange(id0, id0 + n, dtype=np.int32)

    lo, hi = mg.slab_bounds(nb, world, rank)
    n_bound = (rank > 0) + (rank + 1 < world)
    n_c = per_rank - n_bound * n_b
    stride = 2 * per_rank                    # id space reserved per rank / boundary
    xs, ids = [], []
    x, i = ball(7000 + rank, n_c, 0.5 * (lo + hi) / nb, 0.45, rank * stride)
    xs.append(x), ids.append(i)
    for side, bnd in ((0, lo), (1, hi)):
        if (side == 0 and rank == 0) or (side == 1 and rank + 1 == world):
            continue
        bidx = rank if side == 1 else rank - 1          # boundary index = lower rank
        x, i = ball(9000 + bidx, 2 * n_b, bnd / nb, 0.6, bidx * stride + per_rank)
        xs.append(x), ids.append(i)
    x, ids = np.concatenate(xs), np.concatenate(ids)
    keep = mg.partition_particles(x, n_grid, world) == rank
    x, ids = np.ascontiguousarray(x[keep]), np.ascontiguousarray(ids[keep])
    n = len(x)
    v = np.tile(np.array([0.5, -1.0, 0.0], np.float32), (n, 1))
    return x, v, np.zeros((n, 3, 3), np.float32), np.ones(n, np.float32), ids


def c5_ball(b, total, n_grid):
    per_ball = total // 6
    radius = (3.0 * per_ball / 9.7 / (4.0 * np.pi)) ** (1.0 / 3.0) / n_grid
    rng = np.random.default_rng(5 + b)
    n = per_ball if b < 5 else total - 5 * per_ball
    d = rng.standard_normal((n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.float32(radius) * rng.random(n, dtype=np.float32) ** np.float32(1.0 / 3.0)
    return (np.array([0.1 + 0.8 * (b + 0.5) / 6.0, 0.6 if b % 2 == 0 else 0.35, 0.5], np.float32) + d * r[:, None]).astype(np.float32)


def c5_bounds(world, total, n_grid, unit=1):
    """Static slabs cut by the initial particle count per block layer (--slabs weighted); unit = alignment of the cuts in
    block layers (8 = pointer cells; the driver needs none: the loopback test runs unaligned slabs at world 3)."""
    from taichi_b200 import multigpu as mg
    w = np.zeros(n_grid // 4, np.float64)
    for b in range(6):
        x = c5_ball(b, total, n_grid)
        bx = ((x[:, 0] * np.float32(n_grid) - np.float32(0.5)).astype(np.int32)) >> 2
        w += np.bincount(bx, minlength=n_grid // 4)[:n_grid // 4]
    return mg.weighted_slab_bounds(w, world, unit=unit)


def make_c5_cloud(world, rank, total, n_grid, bounds=None):
    """the C3 cloud scaled x8 -- SIX balls spread along x (seed 5), same particle density as C3 (ball radius
    64 cells at 1024^3 for 64M particles), v = (0.5,-1,0).  Slabs are static, so the ranks' loads differ (SURVEY hard part 7)."""
    from taichi_b200 import multigpu as mg
    per_ball = total // 6
    xs, ids = [], []
    for b in range(6):
        x = c5_ball(b, total, n_grid)
        keep = mg.partition_particles(x, n_grid, world, bounds) == rank
        xs.append(np.ascontiguousarray(x[keep]))
        ids.append(np.flatnonzero(keep).astype(np.int32) + np.int32(b * per_ball))
        del x, keep
    x, ids = np.concatenate(xs), np.concatenate(ids)
    n = len(x)
    v = np.tile(np.array([0.5, -1.0, 0.0], np.float32), (n, 1))
    return x, v, np.zeros((n, 3, 3), np.float32), np.ones(n, np.float32), ids


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.rows, self.stop_flag, self.index = [], False, index
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop_flag = True
        self.t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 2 + k and r[2 + k].startswith("Active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(self.rows[0][1]) if self.rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def config_of(name, deactivate_every):
    """The workload both arms run (identical dict in both JSON lines)."""
    n_grid, n_p, dt, layout, desc = WORKLOADS[name]
    return {"workload": desc, "n_grid": n_grid, "particles_per_gpu": n_p, "substeps_per_step": SUBSTEPS, "dt": dt,
            "deactivate_every": deactivate_every if layout != "dense" else 0,
            "l2": "particle state (>=512 MiB/substep) exceeds the 126 MB L2: no flush needed", "parallelism": "single GPU"}


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def cpu_reference(name, steps, warmup, deactivate_every):
    """The reference's CPU path restated on a sparse grid (oracle/mpm_oracle.c mode 1), all host cores.  A step runs
    CPU_SAMPLE_SUBSTEPS substeps of the same cloud (state carried from step to step)."""
    from _desc import load_oracle
    orc = load_oracle()
    n_grid, n_p, dt, layout, desc = WORKLOADS[name]
    cores = os.cpu_count() or 1
    threads = min(cores, 256)
    x, v, Cm, J = make_cloud(name, n_p, seed=3)
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))  # noqa: E731
    cap = min((n_grid // 4) ** 3, 1 << 18)
    nblk = C.c_int32(0)
    k = deactivate_every if layout != "dense" else 0

    def one():
        rc = orc.orc_mpm3d_sparse_substeps(n_grid, C.c_int64(n_p), fp(x), fp(v), fp(Cm), fp(J), C.c_float(dt), C.c_float(400.0),
                                           C.c_float(1.0), C.c_float(9.8), 3, CPU_SAMPLE_SUBSTEPS, 1, threads, k, cap, C.byref(nblk),
                                           None, None, None)
        assert rc == 0
    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    dt_s = (time.perf_counter() - t0) / max(steps, 1)
    sample = (f"{CPU_SAMPLE_SUBSTEPS} of the {SUBSTEPS} substeps of every step on the full cloud ({n_p} particles, sparse blocked grid, "
              f"{nblk.value} active 4^3 blocks, deactivate_every={k}), {threads} threads, fp32 CAS atomics (atomic.h:75)")
    return n_p * CPU_SAMPLE_SUBSTEPS / dt_s, dt_s * 1e3, threads, sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, ms, threads, sample = cpu_reference(args.workload, max(1, args.steps), args.warmup, args.deactivate_every)
    line = {"impl": "reference", "metric": "MPM particle-updates/s", "value": val, "unit": "particle-updates/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_of(args.workload, args.deactivate_every),
            "cpu_baseline": {"value": val, "unit": "particle-updates/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "particle-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_ours(args):
    import torch
    import torch.distributed as dist
    import taichi_b200 as tb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if world > 1 or args.cloud == "c5" or args.force_slabs:
        return run_slabs(args, world, rank, local)
    n_grid, n_p, dt, layout, desc = WORKLOADS[args.workload]
    rt = tb.Runtime(local, device_memory_GB=10 if args.workload == "c3" else 1)  # (the C4 listgen block needs 4.3 GB of containers)
    k_life = args.deactivate_every if layout != "dense" else 0
    solver = tb.Mpm3D(rt, n_grid, n_p, dt, layout=layout, deactivate_every=k_life)
    x, v, Cm, J = make_cloud(args.workload, n_p, seed=3, rank=rank)
    # pinned host buffers (e2e) and the initial upload
    hx, hv, hC, hJ = (torch.from_numpy(a).pin_memory() for a in (x, v, Cm, J))
    hx_out = torch.empty_like(hx).pin_memory()
    solver.set_particles(hx, hv, hC, hJ)
    rt.wait()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, finish=None):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        if finish is not None:
            finish()  # host-blocking: the copy streams are drained before the closing event is recorded
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def step_resident():
        solver.substep(SUBSTEPS)

    # Every e2e step uploads its particle state from pinned host memory, runs the frame and downloads the
    # positions into one of two pinned result buffers.  The calls are asynchronous: upload k+1 (copy stream,
    # double-buffered staging) and download k overlap the substeps of the neighbouring steps; all copies of
    # all K steps lie inside the timed region, which ends after wait_io().
    hx_outs = [hx_out, torch.empty_like(hx).pin_memory()]
    e2e_i = [0]

    def step_e2e():
        solver.set_particles(hx, hv, None, hJ)      # C = 0 at injection (mpm3d.py:87-90) is not uploaded: 28 B per particle
        solver.substep(SUBSTEPS)
        solver.get_particles(x=hx_outs[e2e_i[0] & 1], wait=False)
        e2e_i[0] += 1

    for _ in range(args.warmup):
        step_resident()
    solver.set_profiling(True)
    l0 = rt.launches()
    with ClockSampler(local) as clocks:
        total_ms = timed(step_resident, args.steps)
    launches = rt.launches() - l0
    phases = solver.phase_ms()
    solver.set_profiling(False)
    st = solver.stats()
    # e2e through the C ABI with host buffers
    for _ in range(2):  # untimed: first-touch of the pinned buffers and of the staging allocations
        step_e2e()
    solver.wait_io()
    e2e_ms = timed(step_e2e, max(1, args.steps), finish=solver.wait_io)
    rt.wait()
    x_ref = solver.get_particles(x=hx_out)[0]  # blocking read of the same state: the pipelined result must match
    assert torch.equal(x_ref, hx_outs[(e2e_i[0] - 1) & 1]), "pipelined download differs from the blocking one"
    buf = C.create_string_buffer(256)
    err = tb._lib.lib.b2_check_runtime_error(rt.handle, buf, 256)
    assert err == 0, buf.value

    ms_per_step = total_ms / args.steps
    value = world * n_p * SUBSTEPS / (ms_per_step * 1e-3)
    e2e_value = world * n_p * SUBSTEPS / (e2e_ms / max(1, args.steps) * 1e-3)
    cells = st["active_blocks"] * 64
    peak, peak_kind = hbm_peak()
    extras = {} if args.no_extras else bench_extras(args, rt, tb, solver, timed, peak)
    alg = {"p2g": 64 * n_p + 32 * cells, "g2p": 80 * n_p + 12 * cells, "clear": 16 * cells, "grid_op": 28 * cells}
    dom = max(("p2g", "g2p"), key=lambda k: phases[k])
    traffic = None  # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (C3 only)
    try:
        if args.workload == "c3":
            traffic = json.load(open(os.path.join(REPO, "profiles", "r2_traffic.json"))).get("k_mpm_" + dom)
    except (OSError, ValueError):
        pass
    ach = alg[dom] / (phases[dom] * 1e-3) / 1e9 if phases[dom] > 0 else 0.0
    sub_ms = sum(phases[k] for k in ("deactivate_gc", "clear", "p2g", "listgen", "grid_op", "g2p"))
    sub_alg = 144 * n_p + 88 * cells
    line = {
        "metric": "MPM particle-updates/s", "value": value, "unit": "particle-updates/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": config_of(args.workload, args.deactivate_every),
        "state": {"active_cells": cells, "active_blocks": st["active_blocks"], "occupied_bins": st["occupied_slots"]},
        "e2e": {"value": e2e_value, "unit": "particle-updates/s", "h2d_bytes_per_step": int(n_p * 28), "d2h_bytes_per_step": int(n_p * 12),
                "uploads": "x, v, J (28 B/particle; C = 0 at injection)",
                "returns": "positions only (12 of the 64 B/particle state); uploads and downloads of neighbouring steps are pipelined"},
        "gpu_launches": int(launches),
        "clocks": clocks.summary(),
        "roofline": {"bound": "hbm", "kernel": "k_mpm_" + dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                     "traffic": traffic, "peak_kind": peak_kind, "alg_bytes_per_launch": alg[dom], "ms_per_launch": phases[dom]},
        "roofline_substep": {"alg_bytes": sub_alg, "ms": sub_ms, "achieved": sub_alg / (sub_ms * 1e-3) / 1e9 if sub_ms else 0,
                             "frac": (sub_alg / (sub_ms * 1e-3) / 1e9 / peak) if sub_ms else 0},
        "phase_ms": phases,
    }
    line.update(extras)
    if rank == 0:
        if world == 1 and not args.no_cpu:
            cv, cms, threads, sample = cpu_reference(args.workload, 3, 1, args.deactivate_every)
            line["cpu_baseline"] = {"value": cv, "unit": "particle-updates/s", "cores": threads, "kind": "port", "sample": sample}
        emit(line)
    solver.destroy()
    if world > 1:
        dist.destroy_process_group()


def bench_extras(args, rt, tb, solver, timed, peak):
    """Numbers BASELINE.json names beside the headline: the sparse life cycle at K = 1 / 25 / never, the C4 listgen
    microbenchmark (one case; bench_listgen.py runs all four) and the one figure the reference publishes for this path
    (8M-f32 max-reduction: 57 us on an RTX 3090, docs/lang/articles/performance_tuning/performance.md:125-127)."""
    import torch
    lib = tb._lib.lib
    out = {}
    n_grid, n_p, dt, layout, desc = WORKLOADS[args.workload]
    if layout != "dense":
        life = {}
        for k in (1, SUBSTEPS, 0):
            lib.b2_mpm_set_deactivate_every(solver.handle, k)
            solver.substep(SUBSTEPS)
            solver.set_profiling(True)
            ms = timed(lambda: solver.substep(SUBSTEPS), 2) / 2
            ph = solver.phase_ms()
            solver.set_profiling(False)
            life["never" if k == 0 else f"every_{k}"] = {"value": n_p * SUBSTEPS / (ms * 1e-3), "ms_per_step": ms, "phase_ms": ph,
                                                          "active_blocks": solver.stats()["active_blocks"]}
        lib.b2_mpm_set_deactivate_every(solver.handle, args.deactivate_every)
        out["lifecycle"] = life
    if args.workload == "c3":
        import bench_listgen
        out["listgen"] = bench_listgen.run_case(lib, tb._lib, rt.handle, 1024, "uniform", 0.01, reps=10, cpu=not args.no_cpu, peak=peak)
        n = 8 * 1024 * 1024
        data = torch.rand(n, device="cuda")
        res = torch.zeros(1, device="cuda")
        fn = lambda: lib.b2_reduce_device(rt.handle, C.c_void_p(data.data_ptr()), n, 2, 0, C.c_void_p(res.data_ptr()))  # noqa: E731
        fn()
        torch.cuda.synchronize()
        assert float(res.item()) == float(data.max().item())
        big = torch.empty(256 * 1024 * 1024 // 4, device="cuda")   # 256 MiB written between the reps: the 32 MB input never sits in L2
        ts = []
        for _ in range(20):
            big.fill_(1.0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        us = float(np.median(ts))
        out["reduce_8m_f32_us"] = {"value": us, "unit": "us", "op": "max", "launches": 2, "achieved_gbs": 4 * n / us / 1e3, "frac": 4 * n / us / 1e3 / peak,
                                   "reference_published_us": 57.0, "reference_hw": "RTX 3090 (performance.md:125-127)", "l2": "flushed between reps"}
    return out


def run_slabs(args, world, rank, local):
    """N > 1: config[4] family -- 1024^3 pointer->bitmasked->dense(4^3), 8M particles per GPU (64M at 8),
    slab-decomposed along i with ghost-layer exchange and particle migration over NCCL (weak scaling)."""
    import torch
    import torch.distributed as dist
    import taichi_b200 as tb
    from taichi_b200 import multigpu as mg
    n_grid, per_rank, dt = 1024, 8 * 1024 * 1024, 1e-5
    total_c5 = 64 * 1024 * 1024 if args.strong else per_rank * world
    desc = ("sparse 3D MLS-MPM 1024^3 pointer->bitmasked->dense(4^3), " +
            ("64M particles in total (strong scaling), " if args.strong else "8M particles/GPU (weak scaling), ") +
            ("SURVEY C5 cloud: six balls along x, static slabs " + ("cut by initial particle count, " if args.slabs == "weighted" else "(equal width: unbalanced), ") if args.cloud == "c5" else
             "one ball per slab + half a ball on every slab boundary (balanced), ") + "slabs along i, halo exchange + migration")
    rt = tb.Runtime(local, device_memory_GB=6)
    lo, hi = mg.slab_bounds(n_grid // 4, world, rank)
    if args.cloud == "c5":
        bounds = c5_bounds(world, total_c5, n_grid, args.slab_unit) if args.slabs == "weighted" else None
        if bounds is not None:
            lo, hi = bounds[rank]
        x, v, Cm, J, ids = make_c5_cloud(world, rank, total_c5, n_grid, bounds)
    else:
        assert not args.strong, "--strong uses --cloud c5"
        x, v, Cm, J, ids = make_slab_cloud(world, rank, per_rank, n_grid)
    n_local = len(x)
    slab = mg.CudaSlab(rt, n_grid, max(2 * per_rank, n_local + n_local // 4 + (1 << 20)), dt, lo, hi)
    hx, hv, hC, hJ, hid = (torch.from_numpy(a).pin_memory() for a in (x, v, Cm, J, ids))
    slab.set_particles(hx, hv, hC, hJ, hid)
    drv = mg.NativeSlabDriver(slab, rank, world, cap_blocks=8192, cap_particles=131072)
    counts = torch.zeros(world, device="cuda", dtype=torch.int64)
    counts[rank] = n_local
    if world > 1:
        dist.all_reduce(counts)
    n_total = int(counts.sum().item())
    per_rank_counts = [int(c) for c in counts.tolist()]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, finish=None):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        if finish is not None:
            finish()  # host-blocking: the copy streams are drained before the closing event is recorded
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def step_resident():
        drv.substep(SUBSTEPS)

    # e2e: every step uploads the slab's particle records from pinned host memory, runs the frame with its
    # neighbour exchange and downloads the positions of the particles the slab then owns (records compacted
    # on the device; out_cap entries are copied, the live count comes with them).  Asynchronous calls: upload
    # k+1 and download k overlap the substeps around them; all copies lie inside the timed region.
    out_cap = max(per_rank, n_local) + max(per_rank, n_local) // 4
    hx_outs = [torch.empty(out_cap, 3).pin_memory() for _ in range(2)]
    hcounts = [torch.zeros(1, dtype=torch.int64).pin_memory() for _ in range(2)]
    e2e_i = [0]

    def step_e2e():
        slab.set_particles(hx, hv, None, hJ, hid)   # C = 0 at injection (mpm3d.py:87-90) is not uploaded: 32 B per particle
        drv.substep(SUBSTEPS)
        b = e2e_i[0] & 1
        slab.export_async(out_cap, hcounts[b], x=hx_outs[b])
        e2e_i[0] += 1

    for _ in range(args.warmup):
        step_resident()
    drv.tune()  # messages shrink from the fixed capacity to 2x the observed record counts
    l0 = rt.launches()
    with ClockSampler(local) as clocks:
        total_ms = timed(step_resident, args.steps)
    launches = rt.launches() - l0
    for _ in range(2):  # untimed: first-touch of the pinned buffers and of the staging allocations
        step_e2e()
    slab.wait_io()
    e2e_ms = timed(step_e2e, max(1, args.steps), finish=slab.wait_io)
    n_out = int(hcounts[(e2e_i[0] - 1) & 1].item())
    assert 0 < n_out <= out_cap, f"export of rank {rank}: {n_out} records for a buffer of {out_cap}"
    buf = C.create_string_buffer(256)
    err = tb._lib.lib.b2_check_runtime_error(rt.handle, buf, 256)
    assert err == 0, buf.value
    ms_per_step = total_ms / args.steps
    value = n_total * SUBSTEPS / (ms_per_step * 1e-3)
    e2e_value = n_total * SUBSTEPS / (e2e_ms / max(1, args.steps) * 1e-3)
    peak, peak_kind = hbm_peak()
    sub_alg = 144 * n_total / world  # per GPU, grid term omitted (active cells stay on the device)
    sub_ms = ms_per_step / SUBSTEPS
    ach = sub_alg / (sub_ms * 1e-3) / 1e9
    line = {
        "metric": "MPM particle-updates/s", "value": value, "unit": "particle-updates/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if args.strong else "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "n_grid": n_grid, "particles_total": n_total, "particles_per_rank": per_rank_counts, "cloud": args.cloud, "slabs": args.slabs,
                   "substeps_per_step": SUBSTEPS, "dt": dt,
                   "l2": "particle state per GPU (>=512 MiB/substep) exceeds the 126 MB L2: no flush needed",
                   "parallelism": f"slab{world}: 1 process/GPU, ncclSend/ncclRecv ghost-layer + migrant exchange enqueued from C++, no data-path collective"},
        "e2e": {"value": e2e_value, "unit": "particle-updates/s", "h2d_bytes_per_step": int(n_total * 32), "d2h_bytes_per_step": int(world * (out_cap * 12 + 8)),
                "uploads": "x, v, J, id (32 B/particle; C = 0 at injection)", "returns": "positions only; uploads and downloads of neighbouring steps are pipelined"},
        "gpu_launches": int(launches) * world,
        "clocks": clocks.summary(),
        "roofline": {"bound": "hbm", "kernel": "whole substep (per GPU)", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                     "traffic": None, "peak_kind": peak_kind, "alg_bytes_per_launch": sub_alg, "ms_per_launch": sub_ms},
    }
    if rank == 0:
        emit(line)
    drv.destroy()
    slab.destroy()
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the real stdout; fd 1 itself is routed to stderr for the whole run
    because libraries (NCCL's version banner) write to it directly."""
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--deactivate-every", type=int, default=DEACTIVATE_EVERY,
                    help="sparse life cycle: deactivate_all + GC every K-th substep (0 = never; the reference idiom is 1)")
    ap.add_argument("--no-extras", action="store_true", help="skip the lifecycle / listgen / reduction blocks")
    ap.add_argument("--cloud", default="balanced", choices=["balanced", "c5"],
                    help="N > 1: balanced = one ball per slab (+ boundary balls), c5 = SURVEY C5: six balls along x under static slabs")
    ap.add_argument("--slabs", default="equal", choices=["equal", "weighted"],
                    help="--cloud c5: equal-width slabs, or static slabs cut so every rank starts with the same particle count")
    ap.add_argument("--slab-unit", type=int, default=1, help="--slabs weighted: alignment of the slab cuts in 4^3-block layers")
    ap.add_argument("--force-slabs", action="store_true", help="N = 1 through the slab driver on the N > 1 workload (what one rank of a weak-scaling run does, minus the exchange)")
    ap.add_argument("--strong", action="store_true", help="N > 1 with --cloud c5: 64M particles in total at every N (also N = 1 via torchrun)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
