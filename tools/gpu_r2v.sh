#!/bin/bash
# round 2: tiles split into sub-items when there are few tiles per warp: bench A/B (must not change), few-chain rates, strict mode
tag=${1:-r2v}
mkdir -p gpurun_out
one() { MANIAC_GPU_LIB=$PWD/build_ab/$1 python bench.py --no-cpu-baseline --no-extras --steps 8 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$1', round(d['value']), r['phase_only_ms'])"; }
{ for i in 1 2 3; do one vC.so; one vD.so; done; } 2>&1 | tee gpurun_out/${tag}_ab.log
for b in vC vD; do echo "== $b"; MANIAC_GPU_LIB=$PWD/build_ab/$b.so python tools/few_chains.py; done 2>&1 | tee gpurun_out/${tag}_few_chains.log
for b in vC vD; do echo "== $b"; MANIAC_GPU_LIB=$PWD/build_ab/$b.so python tools/strict_rate.py; done 2>&1 | tee gpurun_out/${tag}_strict.log
timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -n 15 | tee gpurun_out/${tag}_pytest_gpu.log
