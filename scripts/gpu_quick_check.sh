#!/bin/bash
# the whole GPU suite, then the default bench line (T0 bf16, one GPU)
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/quick_tests.log 2>&1; echo rc=$? >> gpurun_out/quick_tests.log
python bench.py --steps 30 --warmup 5 > gpurun_out/quick_bench.json 2> gpurun_out/quick_bench.err; echo rc=$? >> gpurun_out/quick_bench.err
