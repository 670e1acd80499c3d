#!/bin/bash
# r2 G2P round: no-alloc test alone, G2P variants (B2_LIB), ncu source capture of the G2P kernel
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/zz_alone.txt 2>&1; echo "zz alone rc=$?"; tail -3 gpurun_out/zz_alone.txt
L=$PWD/taichi_b200/lib
bash profiles/run_r2_var.sh "$@"
if [ -n "$NCU_LIB" ]; then
  bash profiles/run_ncu.sh k_mpm_g2p 30 1 g2p_$NCU_NAME B2_LIB=$L/$NCU_LIB
fi
