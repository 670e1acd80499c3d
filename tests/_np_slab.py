"""NumPy stand-in for one slab's compute (TEST INFRASTRUCTURE: lets the host logic of
taichi_b200/multigpu.py run under gloo on a CPU box).  Same interface and message formats as
multigpu.CudaSlab; arithmetic follows mpm3d.py:36-81 in fp32."""
import numpy as np
import torch

from taichi_b200.multigpu import BLOCK_REC, PART_REC, HEADER

BLOCK_DT = np.dtype([("key", "<i4"), ("pad", "<i4", 3), ("cells", "<f4", (64, 4))])
PART_DT = np.dtype([("p", "<f4", 16), ("id", "<u4"), ("pad", "<u4", 3)])
assert BLOCK_DT.itemsize == BLOCK_REC and PART_DT.itemsize == PART_REC


class NumpySlab:
    def __init__(self, n_grid, dt, lo, hi, E=400.0, p_rho=1.0, gravity=9.8, bound=3):
        self.n, self.nb, self.lo, self.hi = n_grid, n_grid // 4, lo, hi
        self.dt, self.E, self.gravity, self.bound = np.float32(dt), np.float32(E), np.float32(gravity), bound
        self.dx = np.float32(1.0 / n_grid)
        self.p_vol = (self.dx * np.float32(0.5)) ** 2
        self.p_mass = np.float32(self.p_vol * np.float32(p_rho))
        self.g = np.zeros((n_grid,) * 3 + (4,), np.float32)
        self.active = np.zeros((self.nb,) * 3, bool)
        self.x = np.zeros((0, 3), np.float32)
        self.v = np.zeros((0, 3), np.float32)
        self.C = np.zeros((0, 3, 3), np.float32)
        self.J = np.zeros(0, np.float32)
        self.ids = np.zeros(0, np.int32)

    def new_buffer(self, cap, rec):
        return torch.zeros(HEADER + cap * rec, dtype=torch.uint8)

    def set_particles(self, x, v, Cm, J, ids):
        self.x, self.v, self.C, self.J, self.ids = (np.array(a) for a in (x, v, Cm, J, ids))

    def export(self, cap):
        return self.x, self.v, self.C, self.J, self.ids

    def diagnostics(self):
        v = self.v.astype(np.float64)
        v2 = (v * v).sum(1)
        m = float(self.p_mass)
        return np.array([len(self.x), *(m * v.sum(0)), 0.5 * m * v2.sum(), self.J.astype(np.float64).sum(),
                         np.sqrt(v2).max() if len(v) else 0.0, np.abs(self.J - 1).max() if len(v) else 0.0], np.float64)

    def _stencil(self):
        X = self.x / self.dx
        base = (X - np.float32(0.5)).astype(np.int32)
        fx = X - base.astype(np.float32)
        w = np.stack([0.5 * (1.5 - fx) ** 2, 0.75 - (fx - 1.0) ** 2, 0.5 * (fx - 0.5) ** 2]).astype(np.float32)
        return base, fx.astype(np.float32), w

    def clear_grid(self):
        self.g[:] = 0

    def p2g(self):
        if not len(self.x):
            return
        base, fx, w = self._stencil()
        stress = -self.dt * 4 * self.E * self.p_vol * (self.J - 1) / self.dx ** 2
        A = self.p_mass * self.C + stress[:, None, None] * np.eye(3, dtype=np.float32)
        for i in range(3):
            for j in range(3):
                for k in range(3):
                    off = np.array([i, j, k], np.float32)
                    dpos = (off - fx) * self.dx
                    wt = w[i][:, 0] * w[j][:, 1] * w[k][:, 2]
                    mv = wt[:, None] * (self.p_mass * self.v + np.einsum("pab,pb->pa", A, dpos))
                    idx = base + np.array([i, j, k])
                    np.add.at(self.g, (idx[:, 0], idx[:, 1], idx[:, 2]), np.concatenate([mv, (wt * self.p_mass)[:, None]], 1).astype(np.float32))
                    b = idx >> 2
                    self.active[b[:, 0], b[:, 1], b[:, 2]] = True

    def grid_op(self):
        act = np.repeat(np.repeat(np.repeat(self.active, 4, 0), 4, 1), 4, 2)
        g = self.g
        m = g[..., 3]
        v = g[..., :3].copy()
        pos = m > 0
        v[pos] = v[pos] / m[pos][:, None]
        v[..., 1] -= self.dt * self.gravity
        I = np.stack(np.meshgrid(*[np.arange(self.n)] * 3, indexing="ij"), -1)
        cond = ((I < self.bound) & (v < 0)) | ((I > self.n - self.bound) & (v > 0))
        v[cond] = 0
        g[..., :3][act] = v[act]

    def g2p(self):
        if not len(self.x):
            return
        base, fx, w = self._stencil()
        nv = np.zeros_like(self.v)
        nC = np.zeros_like(self.C)
        for i in range(3):
            for j in range(3):
                for k in range(3):
                    off = np.array([i, j, k], np.float32)
                    dpos = (off - fx) * self.dx
                    wt = w[i][:, 0] * w[j][:, 1] * w[k][:, 2]
                    idx = base + np.array([i, j, k])
                    gv = self.g[idx[:, 0], idx[:, 1], idx[:, 2], :3]
                    nv += wt[:, None] * gv
                    nC += 4 * wt[:, None, None] * np.einsum("pa,pb->pab", gv, dpos) / self.dx ** 2
        self.v = nv.astype(np.float32)
        self.x = (self.x + self.dt * self.v).astype(np.float32)
        self.J = (self.J * (1 + self.dt * np.trace(nC, axis1=1, axis2=2))).astype(np.float32)
        self.C = nC.astype(np.float32)

    # ---- messages ----
    def _view(self, buf, dt, cap):
        a = buf.numpy()
        return a[:HEADER].view("<u4"), a[HEADER:HEADER + cap * dt.itemsize].view(dt)

    def pack_blocks(self, layer, buf, cap):
        cnt, recs = self._view(buf, BLOCK_DT, cap)
        blocks = np.argwhere(self.active[layer]) if 0 <= layer < self.nb else np.zeros((0, 2), int)
        assert len(blocks) <= cap
        cnt[0] = len(blocks)
        for r, (by, bz) in enumerate(blocks):
            recs[r]["key"] = (layer * self.nb + by) * self.nb + bz
            recs[r]["cells"] = self.g[layer * 4:layer * 4 + 4, by * 4:by * 4 + 4, bz * 4:bz * 4 + 4].reshape(64, 4)

    def unpack_blocks(self, buf, cap, mode):
        cnt, recs = self._view(buf, BLOCK_DT, cap)
        for r in range(int(cnt[0])):
            key = int(recs[r]["key"])
            bz, by, bx = key % self.nb, (key // self.nb) % self.nb, key // (self.nb * self.nb)
            sl = (slice(bx * 4, bx * 4 + 4), slice(by * 4, by * 4 + 4), slice(bz * 4, bz * 4 + 4))
            cells = recs[r]["cells"].reshape(4, 4, 4, 4)
            if mode == 0:
                self.g[sl] += cells
                self.active[bx, by, bz] = True
            elif self.active[bx, by, bz]:
                self.g[sl] = cells

    def extract_migrants(self, side, buf, cap):
        cnt, recs = self._view(buf, PART_DT, cap)
        base = (self.x[:, 0] / self.dx - np.float32(0.5)).astype(np.int32) >> 2 if len(self.x) else np.zeros(0, np.int32)
        sel = base < self.lo if side == 0 else base >= self.hi
        n = int(sel.sum())
        assert n <= cap
        cnt[0] = n
        if n:
            p = np.zeros((n, 16), np.float32)
            p[:, :3], p[:, 3] = self.x[sel], self.J[sel]
            p[:, 4:7] = self.v[sel]
            p[:, 7:16] = self.C[sel].reshape(n, 9)
            recs["p"][:n] = p
            recs["id"][:n] = self.ids[sel].astype(np.uint32)
        keep = ~sel
        self.x, self.v, self.C, self.J, self.ids = self.x[keep], self.v[keep], self.C[keep], self.J[keep], self.ids[keep]

    def insert_records(self, buf, cap):
        cnt, recs = self._view(buf, PART_DT, cap)
        n = int(cnt[0])
        if not n:
            return
        p = np.array(recs["p"][:n])
        self.x = np.concatenate([self.x, p[:, :3]])
        self.J = np.concatenate([self.J, p[:, 3]])
        self.v = np.concatenate([self.v, p[:, 4:7]])
        self.C = np.concatenate([self.C, p[:, 7:16].reshape(n, 3, 3)])
        self.ids = np.concatenate([self.ids, recs["id"][:n].astype(np.int32)])
