#!/bin/bash
# r2 experiment runner: C3 bench under several environment variants ("NAME:ENV=val,ENV=val" arguments)
mkdir -p gpurun_out
for spec in "$@"; do
  name=${spec%%:*}; envs=${spec#*:}
  envs=$(echo "$envs" | tr ',' ' ')
  env $envs timeout ${VAR_TIMEOUT:-150} python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/var_$name.json 2> gpurun_out/var_$name.err
  echo "$name rc=$? $(tail -1 gpurun_out/var_$name.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e9,2), {k:round(v,4) for k,v in d['phase_ms'].items()}, round(d['roofline']['frac'],3), round(d['e2e']['value']/1e9,2))")"
done
