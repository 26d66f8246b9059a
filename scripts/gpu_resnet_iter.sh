#!/bin/bash
# quick iteration on the ResNet path: conv kernel + step tests, then the bf16 and fp32 bench lines
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_conv_kernels_gpu.py tests/test_conv_step_gpu.py tests/test_task_resnet_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/resnet_iter_tests.log 2>&1; echo rc=$? >> gpurun_out/resnet_iter_tests.log
for wl in resnet18_bf16_b1024 resnet18_fp32_b1024; do
  timeout 900 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo rc=$? >> gpurun_out/bench_$wl.err
done
