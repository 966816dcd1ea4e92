#!/bin/bash
# round 2: framework-energy cache in the cluster kernel too: tests, few-chain rates, headline A/B
tag=${1:-r2w}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "framework_energy_cache or cluster_per_chain" 2>&1 | tail -n 25 | tee gpurun_out/${tag}_cache_test.log
one() { MANIAC_GPU_LIB=$PWD/build_ab/$1 python bench.py --no-cpu-baseline --no-extras --steps 8 --warmup 3 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('$1', round(d['value']), r['phase_only_ms'])"; }
{ for i in 1 2; do one vC.so; one vE.so; done; } 2>&1 | tee gpurun_out/${tag}_ab.log
MANIAC_GPU_LIB=$PWD/build_ab/vE.so python tools/few_chains.py 2>&1 | tee gpurun_out/${tag}_few_chains.log
timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -n 15 | tee gpurun_out/${tag}_pytest_gpu.log
