#!/bin/bash
mkdir -p gpurun_out
{
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 300 python tools/quick_bench.py svc
} > gpurun_out/r37.log 2>&1
cat gpurun_out/r37.log
