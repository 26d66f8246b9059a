#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_conv_kernels_gpu.py tests/test_conv_step_gpu.py tests/test_task_resnet_gpu.py -m gpu -q --timeout 600 > gpurun_out/resnet_iter_tests.log 2>&1; echo rc=$? >> gpurun_out/resnet_iter_tests.log
