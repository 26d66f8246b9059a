#!/bin/bash
# two GPUs: task.py end-to-end tests (tabular + resnet18) and the bench lines with their parity legs
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_task_resnet_gpu.py tests/test_task_gpu.py -m gpu -q --timeout 900 > gpurun_out/w2_tests.log 2>&1; echo rc=$? >> gpurun_out/w2_tests.log
for wl in resnet18_bf16_b1024 t0_bf16_b65536; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --workload $wl --steps 20 --warmup 3 > gpurun_out/w2_bench_$wl.json 2> gpurun_out/w2_bench_$wl.err; echo rc=$? >> gpurun_out/w2_bench_$wl.err
done
