#!/bin/bash
mkdir -p gpurun_out
{
for rep in 1 2; do
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_base.so timeout 300 python tools/quick_bench.py base
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_rtnt.so timeout 300 python tools/quick_bench.py mask
timeout 300 python tools/quick_bench.py mask+nt
done
timeout 900 python -m pytest tests/test_gpu_fused_step.py -x -q -m gpu 2>&1 | tail -3
} > gpurun_out/r27.log 2>&1
cat gpurun_out/r27.log
