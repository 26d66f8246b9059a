#!/bin/bash
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/full_tests.log 2>&1; echo rc=$? >> gpurun_out/full_tests.log
bash scripts/gpu_ncu_resnet.sh
