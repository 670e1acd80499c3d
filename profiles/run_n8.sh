#!/bin/bash
# N-GPU slab bench (weak scaling, 8M particles per GPU); N = $1
mkdir -p gpurun_out
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench n$N rc=$?"
tail -1 gpurun_out/bench_n$N.json | cut -c1-400
tail -3 gpurun_out/bench_n$N.err
