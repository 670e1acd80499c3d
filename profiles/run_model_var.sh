#!/bin/bash
for n in a b c; do echo "variant $n"; B2_LIB=$PWD/taichi_b200/lib/libb2_m$n.so timeout 200 python profiles/model_timing.py 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read())['materials']; print(d['ms_per_substep'], d['phase_ms'])"; done
