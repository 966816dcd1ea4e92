#!/bin/bash
# round 2, sixth GPU session: parallel Philox draws, phase timing on both workloads, full bench line, ncu launch list + full capture
tag=${1:-r2f}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -n 15 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
MANIAC_GPU_LIB=maniac_mc_b200/ab/lib_T_timing.so python tools/phase_timing.py 592 c4_large_k > gpurun_out/${tag}_phase_timing.log 2>&1
MANIAC_GPU_LIB=maniac_mc_b200/ab/lib_T_timing.so python tools/phase_timing.py 512 zif8_h2o >> gpurun_out/${tag}_phase_timing.log 2>&1
MANIAC_GPU_LIB=maniac_mc_b200/ab/lib_T_timing.so python tools/phase_timing.py 64 zif8_h2o >> gpurun_out/${tag}_phase_timing.log 2>&1; cat gpurun_out/${tag}_phase_timing.log
one() { MANIAC_GPU_LIB=$1 python bench.py --no-cpu-baseline --no-extras --steps 5 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$2', round(d['value']), r['phase_only_ms'], 'frac', round(r['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
{ for i in 1 2; do for v in I_trialsmem J_philox16; do one maniac_mc_b200/ab/lib_$v.so $v; done; done; } > gpurun_out/${tag}_ab.log 2>&1; cat gpurun_out/${tag}_ab.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -n 3 gpurun_out/${tag}_bench.err; python -c "
import json; d=json.load(open('gpurun_out/${tag}_bench.json')); print(d['value'], d['e2e']['value']); print(json.dumps(d.get('strict_mode'), indent=1)); print(json.dumps(d.get('c5_isotherm'), indent=1)); print(d.get('warnings'))"
CMD="python bench.py --moves 32 --steps 1 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_chain_run|k_sys|k_advance' -c 300 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_chain_run' -s 3 -c 2 -f -o gpurun_out/${tag}_prof $CMD > gpurun_out/${tag}_ncu2.log 2>&1
tail -n 2 gpurun_out/${tag}_ncu2.log | cut -c1-300
