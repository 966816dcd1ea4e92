#!/bin/bash
# round 2: cache kept across an identical re-upload (the end-to-end step): test, then the full bench line, cache on / off
tag=${1:-r2s}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "framework_energy_cache or cluster_per_chain or upload or roundtrip" 2>&1 | tail -n 8 | tee gpurun_out/${tag}_cache_test.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 3000 gpurun_out/${tag}_bench.json
MANIAC_FW_CACHE=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/${tag}_bench_nocache.json 2>> gpurun_out/${tag}_bench.err
python - <<'PY'
import json
for f in ("gpurun_out/TAG_bench.json", "gpurun_out/TAG_bench_nocache.json"):
    d = json.loads(open(f.replace("TAG", "r2s")).read().strip().splitlines()[-1])
    print(f, round(d["value"]), round(d["e2e"]["value"]), d["roofline"]["frac"], d["roofline"]["phase_only_ms"], d.get("c5_isotherm", {}).get("value"), d.get("strict_mode", {}))
PY
