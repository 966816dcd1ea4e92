#!/bin/bash
# round 2, first GPU session: full GPU test suite, trajectory-length parity log, Fortran-compiler probe, bench of both arms
tag=${1:-r2a}
mkdir -p gpurun_out
{ echo "# Fortran compiler probe on the GPU box ($(date -u +%FT%TZ))"; for c in gfortran flang flang-new nvfortran ifort ifx f95 f77 pgfortran; do printf "%s: " $c; command -v $c || echo "not found"; done;
  echo "# find / -name f951 -o -name 'flang*' -o -name nvfortran (first 20 hits)"; find / -xdev \( -name f951 -o -name 'flang*' -o -name nvfortran -o -name 'gfortran*' \) -not -path '/proc/*' 2>/dev/null | head -20; echo "# nproc: $(nproc)"; nvidia-smi --query-gpu=name,driver_version --format=csv,noheader; } > gpurun_out/${tag}_fortran_probe.log 2>&1
cat gpurun_out/${tag}_fortran_probe.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 25 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
timeout 1500 python tools/long_parity.py --out gpurun_out/${tag}_long_parity.log > /dev/null 2> gpurun_out/${tag}_long_parity.err; echo "long_parity rc=$?"; tail -n 30 gpurun_out/${tag}_long_parity.log; tail -n 5 gpurun_out/${tag}_long_parity.err
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -n 3 gpurun_out/${tag}_bench.err; cat gpurun_out/${tag}_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2>&1; cat gpurun_out/${tag}_bench_ref.json
