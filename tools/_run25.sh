#!/bin/bash
mkdir -p gpurun_out
{
for rep in 1 2; do
timeout 300 python tools/quick_bench.py new
for v in lone2 ybr both; do
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_$v.so timeout 300 python tools/quick_bench.py $v
done
done
} > gpurun_out/r25.log 2>&1
cat gpurun_out/r25.log
