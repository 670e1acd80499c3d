#!/bin/bash
# 2-GPU check: NCCL slab tests + the slab-decomposed bench (weak scaling, 8M particles per GPU)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multigpu.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_n2.log; cat gpurun_out/pytest_n2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"
tail -1 gpurun_out/bench_n2.json | cut -c1-900
