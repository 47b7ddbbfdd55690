#!/bin/bash
mkdir -p gpurun_out
{
for rep in 1 2; do
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_norow.so timeout 300 python tools/quick_bench.py norow+poll1
timeout 300 python tools/quick_bench.py rowscan+poll1
done
echo "=== clocks CTA 5"
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_clocks.so P2C_WARP_CLOCKS=1 P2C_TILE_CTA=5 P2C_TILE_CLOCKS=1 timeout 300 python tools/tile_clocks.py 256 16 2>&1 | tail -32
} > gpurun_out/r41.log 2>&1
cat gpurun_out/r41.log
