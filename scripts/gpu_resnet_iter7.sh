#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_conv_step_gpu.py tests/test_task_resnet_gpu.py -m gpu -q -s --timeout 600 > gpurun_out/resnet_iter_tests.log 2>&1; echo rc=$? >> gpurun_out/resnet_iter_tests.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/resnet_smoke.log 2>&1; echo rc=$? >> gpurun_out/resnet_smoke.log
timeout 900 python bench.py --workload resnet18_fp32_b1024 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_resnet18_fp32_b1024.json 2> gpurun_out/bench_fp32.err; echo rc=$? >> gpurun_out/bench_fp32.err
