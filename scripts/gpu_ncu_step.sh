# One full training step under `ncu --set full` (T0, bf16, B=65536) + the launch list of the same command.
set -x
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_final.csv $CMD > gpurun_out/ncu_list3.log 2>&1
echo "list rc=$?"
ncu --set full --clock-control none --import-source on -s 60 -c 40 -o gpurun_out/prof_step $CMD > gpurun_out/ncu_full3.log 2>&1
echo "full rc=$?"
