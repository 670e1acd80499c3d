"""Substep time of the F / SVD materials model in the binned solver on the C3 cloud (512^3, 8M particles)."""
import json, os, sys, time
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch
import taichi_b200 as tb
import bench
n_grid, N = 512, 8 * 1024 * 1024
x, v, C, J = bench.make_cloud("c3", N, seed=3)
rng = np.random.default_rng(9)
mat = rng.integers(0, 3, N).astype(np.int32)
rt = tb.Runtime(0, device_memory_GB=4)
out = {}
for model in ("fluid", "materials"):
    s = tb.Mpm3D(rt, n_grid, N, 2e-5, layout="pointer_bitmasked", E=400.0 if model == "fluid" else 1000.0, model=model, deactivate_every=1,
                 p_vol=0.0 if model == "fluid" else (0.5 / n_grid) ** 3)
    if model == "fluid":
        s.set_particles(x, v, None, None)
    else:
        s.set_particles_model(x, v, None, None, None, mat)
    s.substep(25)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); s.substep(100); e1.record(); e1.synchronize()
    ms = e0.elapsed_time(e1) / 100
    s.set_profiling(True); s.substep(25); torch.cuda.synchronize(); ph = s.phase_ms(); s.set_profiling(False)
    st = s.stats()
    out[model] = {"ms_per_substep": ms, "updates_per_s": N / (ms * 1e-3), "particles": st["particles"], "active_blocks": st["active_blocks"], "phase_ms": ph}
    s.destroy()
print(json.dumps(out))
