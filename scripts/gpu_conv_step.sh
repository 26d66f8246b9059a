#!/bin/bash
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_conv_step_gpu.py -m gpu -q -s --timeout 600 > gpurun_out/conv_step.log 2>&1; echo rc=$? >> gpurun_out/conv_step.log
