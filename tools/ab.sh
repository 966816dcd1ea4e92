#!/bin/bash
# A/B of two builds on ONE box: bash tools/ab.sh <libA.so> <libB.so> [rounds]   (run under gpurun)
A=$1; B=$2; N=${3:-3}
one() { MANIAC_GPU_LIB=$1 python bench.py --no-cpu-baseline --no-extras --steps 5 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$2', round(d['value']), r['phase_only_ms'])"; }
for i in $(seq $N); do one $A A; one $B B; done
