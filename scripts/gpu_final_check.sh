set -x
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python -m pytest tests -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain5.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_final2.csv $CMD > gpurun_out/ncu_list5.log 2>&1
echo "list rc=$?"
ncu --set full --clock-control none --import-source on -s 60 -c 40 -o gpurun_out/prof_step2 $CMD > gpurun_out/ncu_full5.log 2>&1
echo "full rc=$?"
