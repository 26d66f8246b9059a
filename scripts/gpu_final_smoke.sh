#!/bin/bash
cd "$(dirname "$0")/.."
timeout 600 python __graft_entry__.py smoke > gpurun_out/final_smoke.log 2>&1; echo rc=$? >> gpurun_out/final_smoke.log
timeout 900 python bench.py --workload resnet18_bf16_b1024 --steps 30 --warmup 5 > gpurun_out/bench_resnet18_bf16_b1024.json 2> gpurun_out/bench_bf16.err; echo rc=$? >> gpurun_out/bench_bf16.err
bash scripts/gpu_ncu_resnet.sh
