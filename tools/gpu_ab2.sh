#!/bin/bash
# A/B of builds under build_ab/ on one box: bash tools/gpu_ab2.sh <tag> <rounds> <a.so> <b.so> [...]
tag=$1; n=$2; shift 2
mkdir -p gpurun_out
one() { MANIAC_GPU_LIB=$PWD/build_ab/$1 python bench.py --no-cpu-baseline --no-extras --steps 8 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$1', round(d['value']), r['phase_only_ms'], d['clocks']['sm_mhz'])"; }
{ for i in $(seq $n); do for b in "$@"; do one $b; done; done; } 2>&1 | tee gpurun_out/${tag}_ab.log
