#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_conv_kernels_gpu.py tests/test_conv_step_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/resnet_iter_tests.log 2>&1; echo rc=$? >> gpurun_out/resnet_iter_tests.log
timeout 900 python bench.py --workload resnet18_bf16_b1024 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_resnet18_bf16_b1024.json 2> gpurun_out/bench_bf16.err; echo rc=$? >> gpurun_out/bench_bf16.err
bash scripts/gpu_ncu_resnet.sh
