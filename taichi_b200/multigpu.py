"""Single-box multi-GPU layer: 1-D slab decomposition of the sparse MPM grid (new -- the reference
is single-GPU: rhi/cuda/cuda_context.cpp:22, c_api/src/taichi_core_impl.cpp:270-297).

One process per GPU.  Rank r owns the 4^3-block layers lo_r <= bx < hi_r and the particles whose
base cell lies in them.  Stencils reach base..base+2, so only ONE foreign layer (bx == hi_r) is
ever written or read.  Per substep:

    clear, P2G            local
    A: ghost layer hi_r   -> r+1 adds it into its own layer lo_{r+1}     (activating)
    grid-op               local
    B: owned layer lo_r   -> r-1 overwrites its ghost layer hi_{r-1}
    G2P + re-binning      local
    C: particles now binned outside [lo_r, hi_r) -> r-1 / r+1 insert them

Messages are fixed-capacity device buffers {u32 count; pad; records}; the receiver reads the count
on the device, so the host never synchronises.  The transport is torch.distributed P2P
(NCCL over NVLink between GPUs; gloo in the CPU tests), or an in-process hand-over when several
slabs are emulated on one GPU.  The per-slab compute goes through a small backend interface:
`CudaSlab` (the C ABI) in production, a NumPy stand-in in the CPU tests of the host logic.
"""
import ctypes as C

import numpy as np

BLOCK_REC = 1040
PART_REC = 80
HEADER = 16


def slab_bounds(nb, world, rank):
    """Contiguous, pointer-cell aligned (8 blocks) slabs along i."""
    unit = 8 if (nb % 8 == 0 and nb // 8 >= world) else 1
    cells = nb // unit
    lo = (cells * rank) // world * unit
    hi = (cells * (rank + 1)) // world * unit
    return lo, hi


def weighted_slab_bounds(layer_weight, world, unit=8):
    """Static slabs cut by load: [(lo, hi)] over block layers along i such that every rank's share of sum(layer_weight)
    is as even as pointer-cell alignment (unit layers) allows; every rank keeps at least one unit."""
    w = np.asarray(layer_weight, np.float64)
    nb = len(w)
    if nb % unit != 0 or nb // unit < world:
        unit = 1
    cells = nb // unit
    cw = np.concatenate([[0.0], np.cumsum(w.reshape(cells, unit).sum(1))])
    cuts = [0]
    for r in range(1, world):
        c = int(np.searchsorted(cw, cw[-1] * r / world, side="left"))
        if c > 0 and abs(cw[c - 1] - cw[-1] * r / world) <= abs(cw[min(c, cells)] - cw[-1] * r / world):
            c -= 1
        cuts.append(min(max(c, cuts[-1] + 1), cells - (world - r)))
    cuts.append(cells)
    return [(cuts[r] * unit, cuts[r + 1] * unit) for r in range(world)]


class CudaSlab:
    """Backend over libb200mpm.so: one Mpm3D restricted to [lo, hi) block layers."""

    def __init__(self, rt, n_grid, max_particles, dt, lo, hi, layout="pointer_bitmasked", **kw):
        import torch
        from . import _lib as L
        from .mpm import LAYOUTS
        self.torch, self.L, self.lib = torch, L, L.lib
        self.lo, self.hi, self.nb, self.n_grid = lo, hi, n_grid // 4, n_grid
        p = L.B2MpmParams(n_grid=n_grid, layout=LAYOUTS[layout], dt=dt, E=kw.get("E", 400.0), p_rho=kw.get("p_rho", 1.0),
                          gravity=kw.get("gravity", 9.8), bound=kw.get("bound", 3), block_local=1, bls_activate_halo=0,
                          slab_lo=lo, slab_hi=hi, deactivate_every=0, max_particles=max_particles, max_bins=0)
        self.handle = self.lib.b2_mpm_create(rt.handle, C.byref(p))
        L.check()
        self.device = torch.device("cuda", self.lib.b2_device_index(rt.handle))

    def _ok(self, rc):
        self.L.check()
        assert rc == 0

    def new_buffer(self, cap, rec):
        return self.torch.zeros(HEADER + cap * rec, dtype=self.torch.uint8, device=self.device)

    def set_particles(self, x, v, Cm, J, ids):
        from .mpm import _ptr
        self._ok(self.lib.b2_mpm_set_particles(self.handle, 0, None, None, None, None))
        if len(x):
            self._ok(self.lib.b2_mpm_add_particles(self.handle, len(x), _ptr(x), _ptr(v), _ptr(Cm), _ptr(J), _ptr(ids)))

    def export_async(self, cap, count, x=None, v=None, Cm=None, J=None, ids=None):
        """Compacted live records into page-locked host arrays of `cap` entries, filled by the solver's copy
        stream after return; `count` (pinned int64[1]) receives the record count. Valid after wait_io()."""
        from .mpm import _ptr
        self._ok(self.lib.b2_mpm_export_async(self.handle, _ptr(x), _ptr(v), _ptr(Cm), _ptr(J), _ptr(ids), cap, _ptr(count)))

    def wait_io(self):
        self._ok(self.lib.b2_mpm_wait_io(self.handle))

    def diagnostics(self):
        """Local (this slab's particles) 8-vector: count, momentum xyz, kinetic energy, sum J, max |v|, max |J-1|."""
        out = (C.c_double * 8)()
        self._ok(self.lib.b2_mpm_diagnostics(self.handle, out))
        return np.array(list(out), np.float64)

    def export(self, cap):
        from .mpm import _ptr
        x, v = np.empty((cap, 3), np.float32), np.empty((cap, 3), np.float32)
        Cm, J, ids = np.empty((cap, 3, 3), np.float32), np.empty(cap, np.float32), np.empty(cap, np.int32)
        n = self.lib.b2_mpm_export(self.handle, _ptr(x), _ptr(v), _ptr(Cm), _ptr(J), _ptr(ids), cap)
        self.L.check()
        assert 0 <= n <= cap
        return x[:n], v[:n], Cm[:n], J[:n], ids[:n]

    def clear_grid(self):
        self._ok(self.lib.b2_mpm_clear_grid(self.handle))

    def p2g(self):
        self._ok(self.lib.b2_mpm_p2g(self.handle))

    def grid_op(self):
        self._ok(self.lib.b2_mpm_grid_op(self.handle))

    def g2p(self):
        self._ok(self.lib.b2_mpm_g2p(self.handle))

    def pack_blocks(self, layer, buf, cap):
        self._ok(self.lib.b2_mpm_pack_blocks(self.handle, layer, C.c_void_p(buf.data_ptr()), cap))

    def unpack_blocks(self, buf, cap, mode):
        self._ok(self.lib.b2_mpm_unpack_blocks(self.handle, C.c_void_p(buf.data_ptr()), cap, mode))

    def extract_migrants(self, side, buf, cap):
        self._ok(self.lib.b2_mpm_extract_migrants(self.handle, side, C.c_void_p(buf.data_ptr()), cap))

    def insert_records(self, buf, cap):
        self._ok(self.lib.b2_mpm_insert_records(self.handle, C.c_void_p(buf.data_ptr()), cap))

    def destroy(self):
        if self.handle:
            self.lib.b2_mpm_destroy(self.handle)
            self.handle = None


class NativeSlabDriver:
    """The production driver: b2_slab_substep runs the whole decomposed substep from C++ (kernels +
    ncclSend/ncclRecv on one stream).  torch.distributed is only used once, to hand rank 0's NCCL
    unique id to the other ranks."""

    def __init__(self, slab, rank, world, cap_blocks=8192, cap_particles=131072, comm=None):
        L = slab.L
        self.slab, self.L, self.lib = slab, L, L.lib
        self.comm, self.own_comm = comm, comm is None
        if world > 1 and comm is None:
            import torch
            import torch.distributed as dist
            idt = torch.zeros(128, dtype=torch.uint8)
            if rank == 0:
                buf = (C.c_uint8 * 128)()
                assert self.lib.b2_comm_unique_id(buf) == 0, L.last_error()
                idt = torch.tensor(list(buf), dtype=torch.uint8)
            idt = idt.cuda()
            dist.broadcast(idt, 0)
            raw = (C.c_uint8 * 128)(*idt.cpu().tolist())
            self.comm = self.lib.b2_comm_init(self.lib.b2_mpm_runtime(slab.handle), raw, rank, world)
            L.check()
        self.handle = self.lib.b2_slab_create(slab.handle, self.comm, rank, world, slab.lo, slab.hi, cap_blocks, cap_particles)
        L.check()

    def substep(self, n=1):
        rc = self.lib.b2_slab_substep(self.handle, n)
        self.L.check()
        assert rc == 0

    def diagnostics(self):
        """Whole-domain diagnostics: local reduction + ncclAllReduce (SUM / MAX) enqueued from C++; every rank
        gets the same 8-vector (count, momentum xyz, kinetic energy, sum J, max |v|, max |J-1|)."""
        out = (C.c_double * 8)()
        rc = self.lib.b2_slab_diagnostics(self.handle, out)
        self.L.check()
        assert rc == 0
        return np.array(list(out), np.float64)

    def tune(self):
        rc = self.lib.b2_slab_tune(self.handle)
        self.L.check()
        assert rc == 0

    def destroy(self):
        if self.handle:
            self.lib.b2_slab_destroy(self.handle)
            self.handle = None
        if self.comm and self.own_comm:
            self.lib.b2_comm_destroy(self.comm)
        self.comm = None


class LoopbackSlabGroup:
    """All slabs of a domain in ONE process on one GPU, advanced by the native driver (b2_slab_substep_lockstep): the
    NCCL send/recv pairs become device-to-device copies, everything else is the production path."""

    def __init__(self, slabs, cap_blocks=8192, cap_particles=131072):
        L = slabs[0].L
        self.L, self.lib, self.world = L, L.lib, len(slabs)
        self.comm = self.lib.b2_comm_loopback(self.lib.b2_mpm_runtime(slabs[0].handle), self.world)
        L.check()
        self.drivers = [NativeSlabDriver(s, r, self.world, cap_blocks, cap_particles, comm=self.comm) for r, s in enumerate(slabs)]
        self.handles = (C.c_void_p * self.world)(*[d.handle for d in self.drivers])

    def substep(self, n=1):
        rc = self.lib.b2_slab_substep_lockstep(self.handles, self.world, n)
        self.L.check()
        assert rc == 0

    def tune(self):
        for d in self.drivers:
            d.tune()

    def diagnostics(self):
        return self.drivers[0].diagnostics()

    def destroy(self):
        for d in self.drivers:
            d.destroy()
        if self.comm:
            self.lib.b2_comm_destroy(self.comm)
            self.comm = None


class SlabDriver:
    """Host logic of one rank.  The substep is a fixed sequence of (compute, message) phases so that
    it can be driven by a torch.distributed communicator or stepped in lockstep in one process."""

    def __init__(self, backend, rank, world, cap_blocks=4096, cap_particles=65536):
        self.b, self.rank, self.world = backend, rank, world
        self.cap_blocks, self.cap_particles = cap_blocks, cap_particles
        nb_ = backend.new_buffer
        self.up, self.down = (rank + 1 if rank + 1 < world else None), (rank - 1 if rank > 0 else None)
        self.sendA, self.recvA = nb_(cap_blocks, BLOCK_REC), nb_(cap_blocks, BLOCK_REC)
        self.sendB, self.recvB = nb_(cap_blocks, BLOCK_REC), nb_(cap_blocks, BLOCK_REC)
        self.sendL, self.sendR = nb_(cap_particles, PART_REC), nb_(cap_particles, PART_REC)
        self.recvL, self.recvR = nb_(cap_particles, PART_REC), nb_(cap_particles, PART_REC)

    # each phase returns the list of messages (send_buf, dst, recv_buf, src) to run after it
    def phase_p2g(self):
        self.b.clear_grid()
        self.b.p2g()
        if self.up is not None:
            self.b.pack_blocks(self.b.hi, self.sendA, self.cap_blocks)
        return [(self.sendA, self.up, self.recvA, self.down)]

    def phase_gridop(self):
        if self.down is not None:
            self.b.unpack_blocks(self.recvA, self.cap_blocks, 0)
        self.b.grid_op()
        if self.down is not None:
            self.b.pack_blocks(self.b.lo, self.sendB, self.cap_blocks)
        return [(self.sendB, self.down, self.recvB, self.up)]

    def phase_g2p(self):
        if self.up is not None:
            self.b.unpack_blocks(self.recvB, self.cap_blocks, 1)
        self.b.g2p()
        msgs = []
        if self.down is not None:
            self.b.extract_migrants(0, self.sendL, self.cap_particles)
        if self.up is not None:
            self.b.extract_migrants(1, self.sendR, self.cap_particles)
        msgs.append((self.sendL, self.down, self.recvR, self.up))   # to the left;  from the right
        msgs.append((self.sendR, self.up, self.recvL, self.down))   # to the right; from the left
        return msgs

    def phase_migrate(self):
        if self.up is not None:
            self.b.insert_records(self.recvR, self.cap_particles)
        if self.down is not None:
            self.b.insert_records(self.recvL, self.cap_particles)
        return []

    PHASES = ("phase_p2g", "phase_gridop", "phase_g2p", "phase_migrate")


def exchange_distributed(msgs, group=None):
    """torch.distributed P2P for one phase's messages (NCCL: stream-ordered, no host sync)."""
    import torch.distributed as dist
    ops = []
    for send, dst, recv, src in msgs:
        if dst is not None:
            ops.append(dist.P2POp(dist.isend, send, dst, group))
        if src is not None:
            ops.append(dist.P2POp(dist.irecv, recv, src, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()


def substep_distributed(driver, nsteps=1, group=None):
    for _ in range(nsteps):
        for ph in SlabDriver.PHASES:
            exchange_distributed(getattr(driver, ph)(), group)


def diagnostics_distributed(backend, group=None):
    """Whole-domain diagnostics through a torch.distributed communicator (gloo on CPU, nccl on GPUs): the slab's
    local 8-vector, then all_reduce SUM of [0..5] and MAX of [6..7]."""
    import torch
    import torch.distributed as dist
    d = torch.from_numpy(np.asarray(backend.diagnostics(), np.float64).copy())
    if dist.get_backend(group) == "nccl":
        d = d.cuda()
    sums, maxs = d[:6].clone(), d[6:].clone()
    dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(maxs, op=dist.ReduceOp.MAX, group=group)
    return torch.cat([sums, maxs]).cpu().numpy()


def substep_local(drivers, nsteps=1):
    """All slabs in one process (several slabs emulated on one GPU / on the CPU): lockstep phases,
    messages handed over by copy."""
    for _ in range(nsteps):
        for ph in SlabDriver.PHASES:
            pending = [getattr(d, ph)() for d in drivers]
            for r, msgs in enumerate(pending):
                for k, (send, dst, recv, src) in enumerate(msgs):
                    if dst is not None:
                        # the matching receive of rank dst is its k-th message's recv buffer
                        drivers[dst_check(dst, len(drivers))]._recv_for(ph, k, r).copy_(send)


def dst_check(dst, n):
    assert 0 <= dst < n
    return dst


def _recv_for(self, ph, k, src):
    """Receive buffer that message k of phase `ph` coming from rank `src` lands in."""
    table = {"phase_p2g": [self.recvA], "phase_gridop": [self.recvB],
             "phase_g2p": [self.recvR, self.recvL]}
    buf = table[ph][k]
    return buf


SlabDriver._recv_for = _recv_for


def partition_particles(x, n_grid, world, bounds=None):
    """Rank that owns each particle: the slab of the block of its base cell (bounds: [(lo, hi)] per rank, default equal slabs)."""
    nb = n_grid // 4
    base = (x[:, 0].astype(np.float32) * np.float32(n_grid) - np.float32(0.5)).astype(np.int32)
    bx = base >> 2
    owner = np.zeros(len(x), np.int32)
    for r in range(world):
        lo, hi = bounds[r] if bounds is not None else slab_bounds(nb, world, r)
        owner[(bx >= lo) & (bx < hi)] = r
    return owner
