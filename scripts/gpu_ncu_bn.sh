#!/bin/bash
# full ncu capture of a few BatchNorm statistics launches of the ResNet-18 step
cd "$(dirname "$0")/.."
ncu --set full --clock-control none --import-source on -k regex:"bn_stats_kernel|bn_bwd_reduce_kernel" --launch-skip 60 -c 6 \
  -o gpurun_out/ncu_bn -f python bench.py --workload resnet18_bf16_b1024 --steps 2 --warmup 3 --no-parity --no-cpu-baseline --sustain-seconds 0 \
  > gpurun_out/ncu_bn.log 2>&1; echo rc=$? >> gpurun_out/ncu_bn.log
ls -la gpurun_out/ncu_bn.ncu-rep >> gpurun_out/ncu_bn.log
