#!/bin/bash
mkdir -p gpurun_out
for ov in 1 0; do
  B2_SLAB_PROF=25 B2_SLAB_OVERLAP=$ov timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/bench_n2_prof$ov.json 2> gpurun_out/bench_n2_prof$ov.err; echo "ov=$ov rc=$?"
  grep "b2_slab rank 0" gpurun_out/bench_n2_prof$ov.err | sed -n "5,6p"
done
B2_SLAB_PROF=25 timeout 600 python bench.py --force-slabs --steps 4 --warmup 3 > gpurun_out/bench_n1_prof.json 2> gpurun_out/bench_n1_prof.err; grep "b2_slab rank" gpurun_out/bench_n1_prof.err | sed -n "4,6p"
