#!/bin/bash
# round 2, third GPU session: full GPU suite on the new default build, per-phase cycle breakdown of the fused kernel,
# A/B against the round-1 configuration (item loop + degree-9 table), start-up timing
tag=${1:-r2c}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -n 25 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
MANIAC_GPU_LIB=maniac_mc_b200/ab/lib_T_timing.so python tools/phase_timing.py > gpurun_out/${tag}_phase_timing.log 2>&1; cat gpurun_out/${tag}_phase_timing.log
one() { MANIAC_GPU_LIB=$1 python bench.py --no-cpu-baseline --no-extras --steps 5 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$2', round(d['value']), r['phase_only_ms'], 'frac', round(r['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'], 'startup_s', round(d.get('startup_system_energy_all_s',-1),2))"; }
{ for i in 1 2; do for v in A_items_tab93 E_dmma_tab84_cstride; do one maniac_mc_b200/ab/lib_$v.so $v; done; done; } > gpurun_out/${tag}_ab.log 2>&1; cat gpurun_out/${tag}_ab.log
