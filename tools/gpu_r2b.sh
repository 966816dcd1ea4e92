#!/bin/bash
# round 2, second GPU session: micro-benchmarks behind the DMMA reciprocal sweep, correctness of the new default build,
# A/B of {item loop, DMMA} x {degree-9 / 8 segments, degree-8 / 16 segments Coulomb table}, parity errors located
tag=${1:-r2b}
mkdir -p gpurun_out
tools/ubench/dmma_rate > gpurun_out/${tag}_ubench_dmma.log 2>&1; cat gpurun_out/${tag}_ubench_dmma.log
tools/ubench/lds_pattern > gpurun_out/${tag}_ubench_lds.log 2>&1; cat gpurun_out/${tag}_ubench_lds.log
python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_long_parity.py::test_strict_and_batched_trajectories_against_oracle 2>&1 | tail -n 25 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
one() { MANIAC_GPU_LIB=$1 python bench.py --no-cpu-baseline --steps 5 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$2', round(d['value']), r['phase_only_ms'], 'frac', round(r['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
{ for i in 1 2; do for v in A_items_tab93 B_items_tab84 C_dmma_tab84 D_dmma_tab93; do one maniac_mc_b200/ab/lib_$v.so $v; done; done; } > gpurun_out/${tag}_ab.log 2>&1; cat gpurun_out/${tag}_ab.log
timeout 900 python tools/long_parity.py --scale 0.1 --out gpurun_out/${tag}_long_parity.log > /dev/null 2> gpurun_out/${tag}_long_parity.err; echo "long_parity rc=$?"; grep -o '"run": "[^"]*"\|"worst": {[^}]*}\|flips": [0-9]*' gpurun_out/${tag}_long_parity.log | paste - - - ; tail -n 3 gpurun_out/${tag}_long_parity.err
