#!/bin/bash
# round 2: framework-energy cache, second build (uniform A(k) pointers, item dealing carried across the two sweeps)
tag=${1:-r2q}
mkdir -p gpurun_out
one() { MANIAC_GPU_LIB=$PWD/build_ab/$1 python bench.py --no-cpu-baseline --no-extras --steps 8 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$1', '$2', round(d['value']), r['phase_only_ms'])"; }
{
for i in 1 2 3; do one libhead.so; one new.so; MANIAC_FW_CACHE=0 one new.so cache_off; done
} 2>&1 | tee gpurun_out/${tag}_ab.log
for b in head_t new_t; do echo "== $b"; MANIAC_GPU_LIB=$PWD/build_ab/$b.so python tools/phase_timing.py; done 2>&1 | tee gpurun_out/${tag}_phase_timing.log
timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -n 15 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
