#!/bin/bash
# round 2, evidence session on the final build: GPU suite, full-length trajectory parity, both bench arms, ncu launch list + full capture
tag=${1:-r2z}
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -n 15 | tee gpurun_out/${tag}_pytest_gpu.log
timeout 1800 python tools/long_parity.py --out gpurun_out/${tag}_long_parity.log > /dev/null 2> gpurun_out/${tag}_long_parity.err; echo "long_parity rc=$?"; tail -n 1 gpurun_out/${tag}_long_parity.log | cut -c1-300; tail -n 3 gpurun_out/${tag}_long_parity.err
python -m pytest tests/test_gpu_long_parity.py::test_h2o_gcmc_mean_n_three_way -m gpu -q -s 2>&1 | grep -E "<N>|passed|failed" | tee gpurun_out/${tag}_gcmc_mean_n.log
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -n 3 gpurun_out/${tag}_bench.err; cut -c1-300 gpurun_out/${tag}_bench.json
python bench.py --impl reference > gpurun_out/${tag}_bench_ref.json 2>&1; cut -c1-200 gpurun_out/${tag}_bench_ref.json
CMD="python bench.py --moves 32 --steps 1 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_chain_run|k_sys|k_advance|k_rebuild' -c 300 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_chain_run' -s 3 -c 2 -f -o gpurun_out/${tag}_prof $CMD > gpurun_out/${tag}_ncu2.log 2>&1
tail -n 2 gpurun_out/${tag}_ncu2.log | cut -c1-300; ls -la gpurun_out/${tag}_prof.ncu-rep
