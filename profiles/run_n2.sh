#!/bin/bash
# 2-GPU call: NCCL slab tests, then the slab bench (balanced weak scaling), then a per-phase event profile of the slab substep
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multigpu.py -x -q > gpurun_out/pytest_n2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_n2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_n2.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e9,2), round(d['ms_per_step'],3), round(d['e2e']['value']/1e9,2), d['config']['particles_per_rank'])"
B2_SLAB_PROF=25 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/bench_n2_prof.json 2> gpurun_out/bench_n2_prof.err
grep "b2_slab rank" gpurun_out/bench_n2_prof.err | sed -n "9,12p"
