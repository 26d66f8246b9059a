# ncu evidence for the round: launch list of a short bench run + one full capture of the heaviest GEMM.
# (run under gpurun; ncu only after the same command exited 0 without it)
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 12 -c 9 -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
ls -la gpurun_out | tail -8
