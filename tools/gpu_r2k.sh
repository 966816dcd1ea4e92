#!/bin/bash
# round 2: per-segment polynomial degree of the Coulomb table: GPU suite, A/B against the fixed degree
tag=${1:-r2k}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -n 15 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
one() { MANIAC_GPU_LIB=$1 python bench.py --no-cpu-baseline --no-extras --steps 5 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$2', round(d['value']), r['phase_only_ms'], 'frac', round(r['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
{ for i in 1 2 3; do for v in Q_fixeddeg P_vardeg; do one maniac_mc_b200/ab/lib_$v.so $v; done; done; } > gpurun_out/${tag}_ab.log 2>&1; cat gpurun_out/${tag}_ab.log
