#!/bin/bash
cd "$(dirname "$0")/.."
timeout -k 10 300 python -m pytest tests/test_conv_kernels_gpu.py -m gpu -q -x --timeout 120 -k "fused_batchnorm" > gpurun_out/bn_fused_tests.log 2>&1; echo rc=$? >> gpurun_out/bn_fused_tests.log
grep -q "rc=0" gpurun_out/bn_fused_tests.log || exit 0
timeout -k 10 600 python -m pytest tests/test_conv_step_gpu.py tests/test_task_resnet_gpu.py -m gpu -q -x --timeout 300 > gpurun_out/resnet_iter_tests.log 2>&1; echo rc=$? >> gpurun_out/resnet_iter_tests.log
for v in 1 0; do
B200STEP_BN_FUSED=$v timeout -k 10 600 python bench.py --workload resnet18_bf16_b1024 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_resnet18_bf16_b1024_bnf$v.json 2> gpurun_out/bench_bnf$v.err; echo rc=$? >> gpurun_out/bench_bnf$v.err
done
