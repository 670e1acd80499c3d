#!/bin/bash
# one gpurun call: whole GPU suite, driver-style bench of both arms, listgen bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_c3.json | cut -c1-1500
timeout 900 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
tail -1 gpurun_out/bench_ref.json | cut -c1-900
timeout 600 python bench.py --workload c2 --steps 4 --warmup 3 --no-cpu > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"
timeout 900 python bench_listgen.py > gpurun_out/listgen.jsonl 2> gpurun_out/listgen.err; echo "listgen rc=$?"
