#!/bin/bash
cd "$(dirname "$0")/.."
timeout -k 10 900 python -m pytest tests/test_conv_kernels_gpu.py tests/test_kernels_gpu.py tests/test_conv_step_gpu.py -m gpu -q -x --timeout 300 > gpurun_out/narrow_tests.log 2>&1; echo rc=$? >> gpurun_out/narrow_tests.log
for v in 1 0; do
B200STEP_NARROW_MMA=$v timeout -k 10 600 python bench.py --workload resnet18_bf16_b1024 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_resnet18_bf16_b1024_nar$v.json 2> gpurun_out/bench_nar$v.err; echo rc=$? >> gpurun_out/bench_nar$v.err
done
