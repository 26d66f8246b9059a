#!/bin/bash
# usage: gpu_one_test.sh <pytest -k expression>: one slice of the GPU suite with its prints
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -q -s -k "$1" > gpurun_out/one_test.log 2>&1; echo rc=$? >> gpurun_out/one_test.log
