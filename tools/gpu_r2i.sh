#!/bin/bash
# round 2: padded, line-aligned row layout of A(k): GPU suite (default build), legacy item-loop build on a parity subset, A/B
tag=${1:-r2i}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -n 15 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
MANIAC_GPU_LIB=maniac_mc_b200/ab/lib_M_items_padded.so python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "batched_tape_replay or strict_mode_trajectory or triclinic_batched or k_sharded_steps" 2>&1 | tail -n 6 > gpurun_out/${tag}_pytest_items.log; cat gpurun_out/${tag}_pytest_items.log
one() { MANIAC_GPU_LIB=$1 python bench.py --no-cpu-baseline --no-extras --steps 5 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$2', round(d['value']), r['phase_only_ms'], 'frac', round(r['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
{ for i in 1 2 3; do for v in K_before_pad L_padded; do one maniac_mc_b200/ab/lib_$v.so $v; done; done; } > gpurun_out/${tag}_ab.log 2>&1; cat gpurun_out/${tag}_ab.log
