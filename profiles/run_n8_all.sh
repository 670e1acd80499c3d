#!/bin/bash
# 8-GPU lines of round 2: weak scaling (8M particles / GPU) and the C5 cloud, 64M particles, strong scaling, weighted slabs
bash profiles/run_n8.sh ${1:-8}
bash profiles/run_n8_strong.sh ${1:-8} weighted
