#!/bin/bash
# one GPU-box session: tests, bench (both arms), ncu launch list + full capture of the two sweeps.
# usage (under gpurun): bash tools/gpu_round.sh <tag> [notests]
tag=${1:-rX}
mkdir -p gpurun_out
if [ "$2" != "notests" ]; then
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
fi
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -3 gpurun_out/${tag}_bench.err; cat gpurun_out/${tag}_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2>&1; cat gpurun_out/${tag}_bench_ref.json
CMD="python bench.py --chains 64 --moves 6 --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_pair_sweep|k_recip_sweep|k_step_tail' -c 300 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
$CMD > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_pair_sweep|k_recip_sweep' -s 20 -c 4 -f -o gpurun_out/${tag}_prof $CMD > gpurun_out/${tag}_ncu2.log 2>&1
tail -3 gpurun_out/${tag}_ncu1.log gpurun_out/${tag}_ncu2.log
ls -la gpurun_out
