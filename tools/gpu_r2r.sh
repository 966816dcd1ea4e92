#!/bin/bash
# round 2: where the real-space phase spends its cycles (cull / framework tiles / rest), cache on and off
tag=${1:-r2r}
mkdir -p gpurun_out
{
echo "== new_ts cache on";  MANIAC_GPU_LIB=$PWD/build_ab/new_ts.so python tools/phase_timing.py
echo "== new_ts cache off"; MANIAC_FW_CACHE=0 MANIAC_GPU_LIB=$PWD/build_ab/new_ts.so python tools/phase_timing.py
} 2>&1 | tee gpurun_out/${tag}_rs_split.log
