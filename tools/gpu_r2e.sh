#!/bin/bash
# round 2, fifth GPU session: Trial in shared memory + fused strict-mode kernel + tile prefetch: tests, phase timing, A/B, full bench line
tag=${1:-r2e}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -n 25 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
MANIAC_GPU_LIB=maniac_mc_b200/ab/lib_T_timing.so python tools/phase_timing.py > gpurun_out/${tag}_phase_timing.log 2>&1; cat gpurun_out/${tag}_phase_timing.log
one() { MANIAC_GPU_LIB=$1 python bench.py --no-cpu-baseline --no-extras --steps 5 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$2', round(d['value']), r['phase_only_ms'], 'frac', round(r['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
{ for i in 1 2; do for v in G_kd_pipe I_trialsmem H_trialsmem_tilepf; do one maniac_mc_b200/ab/lib_$v.so $v; done; done; } > gpurun_out/${tag}_ab.log 2>&1; cat gpurun_out/${tag}_ab.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -n 3 gpurun_out/${tag}_bench.err; python -c "
import json; d=json.load(open('gpurun_out/${tag}_bench.json')); print(d['value'], d['e2e']['value']); print(json.dumps(d.get('strict_mode'), indent=1)); print(json.dumps(d.get('c5_isotherm'), indent=1)); print(d.get('warnings'))"
