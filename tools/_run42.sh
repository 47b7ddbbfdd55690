#!/bin/bash
mkdir -p gpurun_out
{
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_clocks.so P2C_TILE_CTA=5 P2C_TILE_CLOCKS=1 timeout 300 python tools/tile_clocks.py 256 16 2>&1 | tail -12
} > gpurun_out/r42.log 2>&1
cat gpurun_out/r42.log
