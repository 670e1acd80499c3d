#!/bin/bash
# fused life cycle (deactivate_all + GC in one launch): tests, then C3 with the fused and the unfused form
mkdir -p gpurun_out
L=$PWD/taichi_b200/lib/libb2_lc.so
B2_LIB=$L timeout 400 python -m pytest tests/test_gpu_mpm.py tests/test_gpu_snode.py -x -q -m gpu 2>&1 | tail -2
for v in fused unfused; do
  if [ $v = unfused ]; then E="B2_LIFECYCLE_UNFUSED=1"; else E="B2_NONE=1"; fi
  env $E B2_LIB=$L timeout 150 python bench.py --steps 4 --warmup 3 --no-cpu --no-extras > gpurun_out/lc_$v.json 2>/dev/null
  echo "$v $(tail -1 gpurun_out/lc_$v.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e9,2), {k:round(v,4) for k,v in d['phase_ms'].items()}, d['gpu_launches'])")"
done
