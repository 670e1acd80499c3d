import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
import taichi_b200 as tb
from test_gpu_mpm import cloud
rt = tb.Runtime(0)
for layout in ("dense", "pointer_bitmasked"):
    n_grid, N, dt = 64, 30000, 2e-4
    x0, v0, C0, J0 = cloud(11, N, vel=(0.5, -1.0, 0.25))
    s = tb.Mpm3D(rt, n_grid, N, dt, layout=layout, block_local=True)
    s.set_particles(x0, v0, C0, J0)
    def chk(tag):
        try:
            torch.cuda.synchronize()
            print(layout, tag, "ok", flush=True)
        except Exception as e:
            print(layout, tag, "FAILED", str(e)[:100], flush=True); sys.exit(1)
    chk("set")
    s.clear_grid(); chk("clear")
    s.p2g(); chk("p2g")
    s.grid_op(); chk("gridop")
    s.g2p(); chk("g2p")
    s.substep(3); chk("substep3")
    print(s.stats())
