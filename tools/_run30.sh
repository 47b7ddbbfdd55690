#!/bin/bash
mkdir -p gpurun_out
export P2C_LIB=pedestrians_video_2_carla_b200/libp2c_clocks.so
{
for cta in 5 118 125 40; do
  echo "=== CTA $cta"
  P2C_CLOCKS_ALL=$([ $cta = 5 ] && echo 1) P2C_TILE_CTA=$cta P2C_TILE_CLOCKS=1 timeout 300 python tools/tile_clocks.py 256 16 2>&1 | tail -24
done
} > gpurun_out/r30_clocks.log 2>&1
cat gpurun_out/r30_clocks.log
