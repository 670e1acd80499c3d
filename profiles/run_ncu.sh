#!/bin/bash
# ncu --set full capture of kernels matching $1 (regex), $2 = launches to skip, $3 = count, $4 = output name; env in $5...
mkdir -p gpurun_out
K=$1; S=$2; C=$3; O=$4; shift 4
env "$@" timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$K" -s $S -c $C -o gpurun_out/$O -f python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_$O.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/ncu_$O.log | cut -c1-300
