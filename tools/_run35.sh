#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest "tests/test_gpu_fused_step.py::test_fused_step_shapes" -q -m gpu 2>&1 | grep -v "^  warn\|^$" | tail -60
} > gpurun_out/r35.log 2>&1
cat gpurun_out/r35.log
