set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r1_smi.txt 2>&1
T="timeout 300 python -m pytest tests/test_kernels_gpu.py -q --tb=short -p no:cacheprovider -x --timeout=120 --timeout-method=thread"
$T -k "not linear and not dropout" > gpurun_out/r1_misc.log 2>&1; echo "misc rc=$?"
$T -k "linear_fwd and bf16" > gpurun_out/r1_fwd_bf16.log 2>&1; echo "fwd bf16 rc=$?"
$T -k "linear_fwd and f32" > gpurun_out/r1_fwd_f32.log 2>&1; echo "fwd f32 rc=$?"
$T -k "linear_dgrad and bf16" > gpurun_out/r1_dgrad_bf16.log 2>&1; echo "dgrad bf16 rc=$?"
$T -k "linear_dgrad and f32" > gpurun_out/r1_dgrad_f32.log 2>&1; echo "dgrad f32 rc=$?"
$T -k "linear_wgrad and bf16" > gpurun_out/r1_wgrad_bf16.log 2>&1; echo "wgrad bf16 rc=$?"
$T -k "linear_wgrad and f32" > gpurun_out/r1_wgrad_f32.log 2>&1; echo "wgrad f32 rc=$?"
$T -k "dropout or misaligned" > gpurun_out/r1_dropout.log 2>&1; echo "dropout rc=$?"
tail -n 5 gpurun_out/r1_*.log
timeout 900 python -m pytest tests/test_step_parity_gpu.py -q --tb=short -p no:cacheprovider --timeout=300 --timeout-method=thread > gpurun_out/r1_step.log 2>&1; echo "step rc=$?"
tail -n 40 gpurun_out/r1_step.log
