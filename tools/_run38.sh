#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/quick_bench.py entry > gpurun_out/r38_qb.log 2>&1 || exit 1
cat gpurun_out/r38_qb.log
timeout 900 ncu --set full --sampling-interval 0 --clock-control none --import-source on -k regex:tile_kernel -s 3 -c 1 -f -o gpurun_out/r38_tile python tools/prof_step.py 256 16 3 > gpurun_out/r38_ncu.log 2>&1
tail -2 gpurun_out/r38_ncu.log
