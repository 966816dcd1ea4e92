#!/bin/bash
# one GPU-box session: tests, bench (both arms), ncu launch list + full capture of the chain kernel.
# usage (under gpurun): bash tools/gpu_round.sh <tag> [notests|tests] [noref|ref] [noncu]
tag=${1:-rX}
mkdir -p gpurun_out
if [ "$2" != "notests" ]; then
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 15 > gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_pytest_gpu.log
fi
python bench.py --steps 5 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -n 3 gpurun_out/${tag}_bench.err; cat gpurun_out/${tag}_bench.json
if [ "$3" != "noref" ]; then
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2>&1; cat gpurun_out/${tag}_bench_ref.json
fi
if [ "$4" == "noncu" ]; then exit 0; fi
CMD="python bench.py --moves 32 --steps 1 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_chain_run|k_sweep|k_advance' -c 300 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
$CMD > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_chain_run' -s 3 -c 2 -f -o gpurun_out/${tag}_prof $CMD > gpurun_out/${tag}_ncu2.log 2>&1
tail -n 3 gpurun_out/${tag}_ncu1.log gpurun_out/${tag}_ncu2.log
ls -la gpurun_out
