#!/bin/bash
# ResNet-18 path: task.py test, smoke, and the three bench workloads (one GPU)
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_task_resnet_gpu.py tests/test_conv_step_gpu.py -m gpu -q -s --timeout 600 > gpurun_out/resnet_task.log 2>&1; echo rc=$? >> gpurun_out/resnet_task.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/resnet_smoke.log 2>&1; echo rc=$? >> gpurun_out/resnet_smoke.log
for wl in resnet18_bf16_b1024 resnet18_fp32_b1024 resnet18_tf32_b1024; do
  timeout 900 python bench.py --workload $wl --steps 20 --warmup 3 > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo rc=$? >> gpurun_out/bench_$wl.err
done
