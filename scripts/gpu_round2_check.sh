#!/bin/bash
# Round-2 verification on one B200: 3xTF32 kernel tests first, the whole GPU suite, the config-2 bench lines
# (fp32 = 3xTF32, tf32 = single pass), then one step under ncu.
cd "$(dirname "$0")/.."
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "split_tf32 or three_pass" > gpurun_out/r2_t27a.log 2>&1; echo rc=$? >> gpurun_out/r2_t27a.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2_t27.log 2>&1; echo rc=$? >> gpurun_out/r2_t27.log
python bench.py --workload t0_fp32_b4096 --steps 50 --warmup 5 > gpurun_out/t0_fp32.json 2> gpurun_out/t0_fp32.err
python bench.py --workload t0_tf32_b4096 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/t0_tf32.json 2> gpurun_out/t0_tf32.err
bash scripts/gpu_ncu_step.sh
