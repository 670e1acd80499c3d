#!/bin/bash
# r2 quick A/B: MPM GPU tests with the warp-autonomous P2G, then C3 bench with it and with r1's kernel
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mpm.py -x -q > gpurun_out/pytest_mpm.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_mpm.log
tail -15 gpurun_out/pytest_mpm.log
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/bench_c3_new.json 2> gpurun_out/bench_c3_new.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_c3_new.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['phase_ms'], d['roofline']['frac'])"


