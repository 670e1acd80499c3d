import numpy as np, sys
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import bench, taichi_b200 as tb
n_grid,N,dt,layout,_=bench.WORKLOADS['c3']
x,v,C,J=bench.make_cloud('c3',N,3)
rt=tb.Runtime(0,device_memory_GB=4)
s=tb.Mpm3D(rt,n_grid,N,dt,layout=layout)
s.set_particles(x,v,C,J)
s.substep(3); rt.wait()
