#!/bin/bash
# evidence of round 2: tests that the -x run skipped, smoke, ncu launch list of the bench command, ncu --set full of P2G / G2P
mkdir -p gpurun_out
timeout 600 python -m pytest "tests/test_gpu_snode.py::test_out_of_memory_never_writes_the_ambient_node" tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_rest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_rest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > gpurun_out/bench_short.json 2>&1; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_mpm_p2g|k_mpm_g2p" -s 30 -c 2 -o gpurun_out/p2g_g2p_final -f python bench.py --steps 1 --warmup 1 --no-cpu --no-extras > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
