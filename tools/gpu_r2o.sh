#!/bin/bash
# round 2: framework-energy cache: its own test first, the full GPU suite, then bench A/B: cache on / off (same build) / HEAD build
tag=${1:-r2o}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "framework_energy_cache" 2>&1 | tail -n 25 > gpurun_out/${tag}_cache_test.log; cat gpurun_out/${tag}_cache_test.log
timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -n 15 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
{
for rep in 1 2; do
  echo "== cache on";  python bench.py --no-cpu-baseline --no-extras --steps 10 --warmup 3
  echo "== cache off"; MANIAC_FW_CACHE=0 python bench.py --no-cpu-baseline --no-extras --steps 10 --warmup 3
  echo "== HEAD build (8f5bf66)"; MANIAC_GPU_LIB=$PWD/build_ab/libhead.so python bench.py --no-cpu-baseline --no-extras --steps 10 --warmup 3
done
} > gpurun_out/${tag}_ab_fw_cache.log 2>&1
grep -E "^==|\"value\"" gpurun_out/${tag}_ab_fw_cache.log | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if l.startswith('=='): print(l, end='  ')
    else:
        j = json.loads(l); r = j.get('roofline', {}); print(j['value'], j['ms_per_step'], r.get('frac'), r.get('phase_only_ms'))
"
python tools/phase_timing.py > gpurun_out/${tag}_phase_timing.log 2>&1; tail -n 12 gpurun_out/${tag}_phase_timing.log
