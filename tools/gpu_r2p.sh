#!/bin/bash
# round 2: why is the build with the framework-energy cache slower in EVERY phase?  builds side by side on one box
tag=${1:-r2p}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "framework_energy_cache" 2>&1 | tail -n 25 > gpurun_out/${tag}_cache_test.log; cat gpurun_out/${tag}_cache_test.log
one() { MANIAC_GPU_LIB=$PWD/build_ab/$1 python bench.py --no-cpu-baseline --no-extras --steps 8 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$1', '$2', round(d['value']), r['phase_only_ms'])"; }
{
for i in 1 2; do one libhead.so; one new.so; one new_allocend.so; MANIAC_FW_CACHE=0 one new_allocend.so cache_off; done
} 2>&1 | tee gpurun_out/${tag}_ab.log
for b in head_t new_t; do echo "== $b"; MANIAC_GPU_LIB=$PWD/build_ab/$b.so python tools/phase_timing.py; done 2>&1 | tee gpurun_out/${tag}_phase_timing.log
