#!/bin/bash
mkdir -p gpurun_out
{
for rep in 1 2; do
timeout 300 python tools/quick_bench.py new
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_nf7.so timeout 300 python tools/quick_bench.py nf7
done
for cta in 125 119; do
  echo "=== nf7 CTA $cta"
  P2C_LIB=pedestrians_video_2_carla_b200/libp2c_nf7c.so P2C_TILE_CTA=$cta P2C_TILE_CLOCKS=1 timeout 300 python tools/tile_clocks.py 256 16 2>&1 | tail -18 | head -14
done
} > gpurun_out/r31.log 2>&1
cat gpurun_out/r31.log
