#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_conv_kernels_gpu.py tests/test_conv_step_gpu.py tests/test_task_resnet_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/resnet_iter_tests.log 2>&1; echo rc=$? >> gpurun_out/resnet_iter_tests.log
for sp in 1 0; do
B200STEP_CONV_SPLIT=$sp timeout 900 python bench.py --workload resnet18_bf16_b1024 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_resnet18_bf16_b1024_sp$sp.json 2> gpurun_out/bench_sp$sp.err; echo rc=$? >> gpurun_out/bench_sp$sp.err
done
timeout 900 python bench.py --workload resnet18_tf32_b1024 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_resnet18_tf32_b1024.json 2> gpurun_out/bench_tf32.err; echo rc=$? >> gpurun_out/bench_tf32.err
