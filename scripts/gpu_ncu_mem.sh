set -x
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"ce_loss|preprocess|emb_bwd|sgd_kernel|reduce_partials" -s 10 -c 10 -o gpurun_out/prof_mem $CMD > gpurun_out/ncu_full2.log 2>&1
echo "rc=$?"
