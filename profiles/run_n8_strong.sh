#!/bin/bash
# C5 cloud (six balls along x), 64M particles in total, strong scaling, static slabs cut by particle count; N = $1
mkdir -p gpurun_out
N=$1
SL=${2:-weighted}
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 3 --warmup 3 --cloud c5 --strong --slabs $SL > gpurun_out/bench_c5strong_n${N}_$SL.json 2> gpurun_out/bench_c5strong_n${N}_$SL.err; echo "bench c5 strong n$N $SL rc=$?"
tail -1 gpurun_out/bench_c5strong_n${N}_$SL.json | cut -c1-1200
tail -3 gpurun_out/bench_c5strong_n${N}_$SL.err
