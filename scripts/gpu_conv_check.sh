#!/bin/bash
# the convolutional path: kernel tests, then the ResNet-18 step tests
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_conv_kernels_gpu.py -m gpu -q -x --timeout 300 > gpurun_out/conv_kernels.log 2>&1; echo rc=$? >> gpurun_out/conv_kernels.log
timeout 1200 python -m pytest tests/test_conv_step_gpu.py -m gpu -q -s --timeout 600 > gpurun_out/conv_step.log 2>&1; echo rc=$? >> gpurun_out/conv_step.log
