#!/bin/bash
cd "$(dirname "$0")/.."
B200STEP_SIDE_WGRAD=0 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"gemm_kernel<.int.0, .bool.0, .bool.0, .int.0" --launch-skip 63 -c 6 \
  -o gpurun_out/ncu_conv_fwd -f python bench.py --workload resnet18_bf16_b1024 --steps 1 --warmup 3 --no-parity --no-cpu-baseline --sustain-seconds 0 \
  > gpurun_out/ncu_conv_fwd.log 2>&1; echo rc=$? >> gpurun_out/ncu_conv_fwd.log
