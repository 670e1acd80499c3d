#!/bin/bash
# quick GPU loop: MPM parity tests + C3 bench under the env settings given as arguments ("A=1 B=2" ...)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mpm.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_mpm.log; cat gpurun_out/pytest_mpm.log
i=0
for cfg in "$@"; do
  i=$((i+1))
  echo "== $cfg"
  env $cfg timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_q$i.json 2> gpurun_out/bench_q$i.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_q$i.json").read().strip().splitlines()[-1])
    print("value %.3e e2e %.3e"%(d["value"], d["e2e"]["value"]), d["phase_ms"], d["roofline"]["frac"], d["roofline_substep"]["frac"])
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/bench_q$i.err").read()[-2000:])
PY
done
