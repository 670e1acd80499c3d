#!/bin/bash
# C2 (dense 128^3, 1M particles) and C3 with the r1 and r2 G2P bin prologue
mkdir -p gpurun_out
L=$PWD/taichi_b200/lib/libb2_sw.so
for w in c2 c3; do for v in r2 r1; do
  if [ $v = r1 ]; then E="B2_G2P_R1=1"; else E="B2_NONE=1"; fi
  env $E B2_LIB=$L timeout 150 python bench.py --workload $w --steps 4 --warmup 3 --no-cpu --no-extras > gpurun_out/sw_${w}_$v.json 2>/dev/null
  echo "$w $v $(tail -1 gpurun_out/sw_${w}_$v.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e9,2), {k:round(v,4) for k,v in d['phase_ms'].items()})")"
done; done
