#!/bin/bash
# r2 G2P round 2: MPM GPU tests on an experiment library ($TEST_LIB), then bench variants, then optional ncu
mkdir -p gpurun_out
L=$PWD/taichi_b200/lib
if [ -n "$TEST_LIB" ]; then
  B2_LIB=$L/$TEST_LIB timeout ${TEST_TIMEOUT:-420} python -m pytest ${TESTS:-tests/test_gpu_mpm.py tests/test_gpu_multigpu.py} -x -q -m gpu > gpurun_out/pytest_$TEST_LIB.txt 2>&1; rc=$?; echo "pytest rc=$rc"; tail -5 gpurun_out/pytest_$TEST_LIB.txt
  if [ $rc -ne 0 ]; then echo "tests failed: benches skipped"; exit 1; fi
fi
bash profiles/run_r2_var.sh "$@"
if [ -n "$NCU_LIB" ]; then
  bash profiles/run_ncu.sh "${NCU_K:-k_mpm_g2p}" 30 1 g2p_$NCU_NAME B2_LIB=$L/$NCU_LIB
fi
