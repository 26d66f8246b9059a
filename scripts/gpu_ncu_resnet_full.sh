#!/bin/bash
# (1) DRAM traffic per kernel of the current ResNet-18 step; (2) ncu --set full of a few implicit-GEMM conv launches
cd "$(dirname "$0")/.."
bash scripts/gpu_ncu_resnet_traffic.sh resnet18_bf16_b1024
B200STEP_SIDE_WGRAD=0 ncu --set full --clock-control none --import-source on -k regex:"gemm_kernel" --launch-skip 470 -c 8 \
  -o gpurun_out/ncu_conv_gemm -f python bench.py --workload resnet18_bf16_b1024 --steps 1 --warmup 3 --no-parity --no-cpu-baseline --sustain-seconds 0 \
  > gpurun_out/ncu_conv_gemm.log 2>&1; echo rc=$? >> gpurun_out/ncu_conv_gemm.log
