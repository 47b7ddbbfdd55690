#!/bin/bash
mkdir -p gpurun_out
{
for rep in 1 2; do
timeout 300 python tools/quick_bench.py new
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_late.so timeout 300 python tools/quick_bench.py late
done
for v in clocks latec; do
  echo "=== $v CTA 5"
  P2C_LIB=pedestrians_video_2_carla_b200/libp2c_$v.so P2C_TILE_CTA=5 P2C_TILE_CLOCKS=1 timeout 300 python tools/tile_clocks.py 256 16 2>&1 | tail -19 | head -16
done
} > gpurun_out/r32.log 2>&1
cat gpurun_out/r32.log
