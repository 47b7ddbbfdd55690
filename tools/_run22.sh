#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python tools/prof_step.py 8192 16 5 > gpurun_out/r22_plain.log 2>&1 || exit 1
cat gpurun_out/r22_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 3 -c 1 -f -o gpurun_out/r22_tile python tools/prof_step.py 8192 16 3 > gpurun_out/r22_ncu.log 2>&1
tail -3 gpurun_out/r22_ncu.log
