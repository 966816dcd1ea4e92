#!/bin/bash
# round 2: single-precision tile cull: parity tests that exercise the cull, headline A/B, phase split
tag=${1:-r2x}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_long_parity.py -m gpu -q -x -k "culled or c4 or large_framework or bench_system or triclinic_cull or framework_energy_cache" 2>&1 | tail -n 12 | tee gpurun_out/${tag}_cull_tests.log
one() { MANIAC_GPU_LIB=$PWD/build_ab/$1 python bench.py --no-cpu-baseline --no-extras --steps 8 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$1', round(d['value']), r['phase_only_ms'])"; }
{ for i in 1 2 3; do one vE.so; one vF.so; done; } 2>&1 | tee gpurun_out/${tag}_ab.log
MANIAC_GPU_LIB=$PWD/build_ab/new_ts.so python tools/phase_timing.py 2>&1 | tee gpurun_out/${tag}_rs_split.log
MANIAC_GPU_LIB=$PWD/build_ab/vF.so python tools/strict_rate.py 2>&1 | tee gpurun_out/${tag}_strict.log
