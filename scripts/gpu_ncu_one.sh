set -x
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$1" -s 6 -c 2 -o gpurun_out/prof_one $CMD > gpurun_out/ncu_full4.log 2>&1
echo "rc=$?"
