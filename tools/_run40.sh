#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_gpu_fused_step.py tests/test_gpu_exchange.py tests/test_gpu_module_api.py -x -q -m gpu 2>&1 | tail -5
for rep in 1 2; do
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_norow.so timeout 300 python tools/quick_bench.py norow
timeout 300 python tools/quick_bench.py rowscan
done
echo "=== clocks CTA 5"
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_clocks.so P2C_WARP_CLOCKS=1 P2C_TILE_CTA=5 P2C_TILE_CLOCKS=1 timeout 300 python tools/tile_clocks.py 256 16 2>&1 | tail -32
} > gpurun_out/r40.log 2>&1
cat gpurun_out/r40.log
