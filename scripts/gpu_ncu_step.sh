#!/bin/bash
# One graph-replayed T0 step under ncu --set full (11 launches): skips the bf16 cast and the eager warm-up step.
cd "$(dirname "$0")/.."
timeout 500 ncu --set full --clock-control none -k regex:"^(preprocess|gemm_kernel|linear_ce|reduce_partials|emb_bwd|wgrad_grouped|sgd_kernel|cast_bf16)" --launch-skip 12 -c 11 -o gpurun_out/r02_step -f python bench.py --steps 2 --warmup 3 --no-parity --no-cpu-baseline --sustain-seconds 0 > gpurun_out/r02_ncu_full.log 2>&1; echo rc=$? >> gpurun_out/r02_ncu_full.log
