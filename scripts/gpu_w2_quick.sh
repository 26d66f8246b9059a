#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_task_resnet_gpu.py tests/test_conv_step_gpu.py -m gpu -q --timeout 600 > gpurun_out/w2_tests.log 2>&1; echo rc=$? >> gpurun_out/w2_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --workload resnet18_bf16_b1024 --steps 20 --warmup 3 --sustain-seconds 0 > gpurun_out/w2_bench_resnet18_bf16_b1024.json 2> gpurun_out/w2_bench_resnet18_bf16_b1024.err; echo rc=$? >> gpurun_out/w2_bench_resnet18_bf16_b1024.err
