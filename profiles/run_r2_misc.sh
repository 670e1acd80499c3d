#!/bin/bash
# e2e against the number of timed steps; the slab driver on one GPU (weak-scaling workload and the 64M-particle C5 cloud)
mkdir -p gpurun_out
for k in 4 8 16; do
  timeout 300 python bench.py --steps $k --warmup 3 --no-cpu --no-extras > gpurun_out/e2e_k$k.json 2>/dev/null
  echo "K=$k $(tail -1 gpurun_out/e2e_k$k.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e9,2), round(d['e2e']['value']/1e9,2))")"
done
timeout 600 python bench.py --gpus 1 --force-slabs --steps 3 --warmup 3 > gpurun_out/bench_n1_slab.json 2> gpurun_out/bench_n1_slab.err; echo "n1 slab rc=$?"
timeout 900 python bench.py --gpus 1 --cloud c5 --strong --steps 3 --warmup 3 > gpurun_out/bench_c5strong_n1.json 2> gpurun_out/bench_c5strong_n1.err; echo "c5 n1 rc=$?"
for f in bench_n1_slab bench_c5strong_n1; do tail -1 gpurun_out/$f.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e9,2), round(d['e2e']['value']/1e9,2), d['ms_per_step'])"; done
