#!/bin/bash
mkdir -p gpurun_out
{
for i in 1 2 3; do
timeout 300 python tools/diag_loss.py 2 64 78 2>&1 | grep -v warn | tail -9
done
P2C_LIB=pedestrians_video_2_carla_b200/libp2c_nosvc.so timeout 300 python tools/diag_loss.py 2 64 78 2>&1 | grep -v warn | tail -9
} > gpurun_out/r36.log 2>&1
cat gpurun_out/r36.log
