#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/prof_step.py 256 16 3 > gpurun_out/r39_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --warp-sampling-interval 0 --clock-control none --import-source on -k regex:tile_kernel -s 3 -c 1 -f -o gpurun_out/r39_tile python tools/prof_step.py 256 16 3 > gpurun_out/r39_ncu.log 2>&1
tail -2 gpurun_out/r39_ncu.log
