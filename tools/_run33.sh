#!/bin/bash
mkdir -p gpurun_out
export P2C_LIB=pedestrians_video_2_carla_b200/libp2c_clocks.so
{
echo "=== 256x16 CTA 5"
P2C_WARP_CLOCKS=1 P2C_TILE_CTA=5 P2C_TILE_CLOCKS=1 timeout 300 python tools/tile_clocks.py 256 16 2>&1 | tail -32
echo "=== 8192x16 CTA 5 tile 3"
P2C_WARP_CLOCKS=1 P2C_TILE_CTA=5 P2C_TILE_CLOCKS=3 timeout 300 python tools/tile_clocks.py 8192 16 2>&1 | tail -32
} > gpurun_out/r33.log 2>&1
cat gpurun_out/r33.log
