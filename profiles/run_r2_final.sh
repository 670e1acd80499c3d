#!/bin/bash
# final check of round 2 on a fresh box: whole GPU suite, smoke, both bench arms with default flags, C2, listgen,
# ncu launch list of the bench command and ncu --set full of P2G / G2P / the life-cycle kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --workload c2 --steps 4 --warmup 3 --no-cpu --no-extras > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_mpm_p2g|k_mpm_g2p|k_deactivate_all_gc" -s 45 -c 3 -o gpurun_out/p2g_g2p_final -f python bench.py --steps 1 --warmup 1 --no-cpu --no-extras > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
