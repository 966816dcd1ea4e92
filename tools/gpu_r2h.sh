#!/bin/bash
# round 2: evidence session -- full-length trajectory parity (SURVEY 8d lengths), both bench arms, Fortran-compiler probe
tag=${1:-r2h}
mkdir -p gpurun_out
{ echo "# Fortran compiler probe on the GPU box ($(date -u +%FT%TZ))"; for c in gfortran flang flang-new nvfortran ifort ifx f95 f77 pgfortran; do printf "%s: " $c; command -v $c || echo "not found"; done;
  echo "# find / ( -name f951 -o -name 'flang*' -o -name nvfortran -o -name 'gfortran*' ) (first 20 hits)"; find / -xdev \( -name f951 -o -name 'flang*' -o -name nvfortran -o -name 'gfortran*' \) -not -path '/proc/*' 2>/dev/null | head -20; echo "# nproc: $(nproc)"; nvidia-smi --query-gpu=name,driver_version --format=csv,noheader; } > gpurun_out/${tag}_fortran_probe.log 2>&1
timeout 1800 python tools/long_parity.py --out gpurun_out/${tag}_long_parity.log > /dev/null 2> gpurun_out/${tag}_long_parity.err; echo "long_parity rc=$?"; tail -n 2 gpurun_out/${tag}_long_parity.log | cut -c1-300; tail -n 5 gpurun_out/${tag}_long_parity.err
python -m pytest tests/test_gpu_long_parity.py::test_h2o_gcmc_mean_n_three_way -m gpu -q -s 2>&1 | grep -E "<N>|passed|failed" > gpurun_out/${tag}_gcmc_mean_n.log; cat gpurun_out/${tag}_gcmc_mean_n.log
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -n 3 gpurun_out/${tag}_bench.err; cut -c1-400 gpurun_out/${tag}_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2>&1; cut -c1-200 gpurun_out/${tag}_bench_ref.json
