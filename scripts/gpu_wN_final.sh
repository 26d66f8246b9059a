#!/bin/bash
# usage: gpu_wN_final.sh N : ResNet-18 and the default workload on N GPUs with their parity legs
cd "$(dirname "$0")/.."
N=${1:-4}
for wl in resnet18_bf16_b1024 t0_bf16_b65536; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus $N --workload $wl --steps 20 --warmup 3 --sustain-seconds 0.5 > gpurun_out/w${N}_bench_$wl.json 2> gpurun_out/w${N}_bench_$wl.err; echo rc=$? >> gpurun_out/w${N}_bench_$wl.err
done
