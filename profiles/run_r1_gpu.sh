#!/bin/bash
# one gpurun call: GPU tests, bench, launch list, full ncu capture of P2G/G2P
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"
tail -2 gpurun_out/bench_c3.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_mpm_p2g|k_mpm_g2p' -s 20 -c 2 -o gpurun_out/mpm_full -f python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_full.log
