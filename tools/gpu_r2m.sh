#!/bin/bash
# round 2: cluster-per-chain kernel: full GPU suite, few-chain rates per cluster size
tag=${1:-r2m}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -n 15 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
python tools/few_chains.py > gpurun_out/${tag}_few_chains.log 2>&1; cat gpurun_out/${tag}_few_chains.log
