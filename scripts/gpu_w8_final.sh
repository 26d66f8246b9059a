#!/bin/bash
# 8 GPUs: ResNet-18 and the default workload with their parity legs
cd "$(dirname "$0")/.."
for wl in resnet18_bf16_b1024 t0_bf16_b65536; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 8 --workload $wl --steps 20 --warmup 3 --sustain-seconds 0.5 > gpurun_out/w8_bench_$wl.json 2> gpurun_out/w8_bench_$wl.err; echo rc=$? >> gpurun_out/w8_bench_$wl.err
done
