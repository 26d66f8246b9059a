#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_conv_step_gpu.py tests/test_task_resnet_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/resnet_iter_tests.log 2>&1; echo rc=$? >> gpurun_out/resnet_iter_tests.log
for dg in 1 0; do
B200STEP_IMPLICIT_DGRAD=$dg timeout 900 python bench.py --workload resnet18_bf16_b1024 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_resnet18_bf16_b1024_dg$dg.json 2> gpurun_out/bench_dg$dg.err; echo rc=$? >> gpurun_out/bench_dg$dg.err
done
timeout 900 python bench.py --workload resnet18_fp32_b1024 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_resnet18_fp32_b1024.json 2> gpurun_out/bench_fp32.err; echo rc=$? >> gpurun_out/bench_fp32.err
cp gpurun_out/bench_resnet18_bf16_b1024_dg1.json gpurun_out/bench_resnet18_bf16_b1024.json
