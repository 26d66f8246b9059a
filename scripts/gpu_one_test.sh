#!/bin/bash
# usage: gpu_one_test.sh <pytest -k expression> [file]: one slice of the GPU suite with its prints, under a hard timeout
cd "$(dirname "$0")/.."
timeout -k 10 ${3:-300} python -m pytest ${2:-tests} -m gpu -q -s -x --timeout 120 -k "$1" > gpurun_out/one_test.log 2>&1; echo rc=$? >> gpurun_out/one_test.log
