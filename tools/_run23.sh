#!/bin/bash
mkdir -p gpurun_out
export P2C_LIB=pedestrians_video_2_carla_b200/libp2c_clocks.so
for cfg in "256 16" "8192 16"; do
  P2C_TILE_CLOCKS=2 timeout 300 python tools/tile_clocks.py $cfg 2>&1 | tail -22
done > gpurun_out/r23_clocks.log 2>&1
cat gpurun_out/r23_clocks.log
