mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_step_parity_gpu.py tests/test_task_gpu.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
run() { n=$1; tag=$2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/s2_n${n}_$tag.json 2> gpurun_out/s2_n${n}_$tag.err; echo "n=$n $tag rc=$?"; }
python bench.py --gpus 1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/s2_n1_a.json 2> gpurun_out/s2_n1_a.err
run 2 a; run 2 b; run 4 a; run 8 a; run 8 b
python - <<PY
import json,glob
base=None
for f in sorted(glob.glob("gpurun_out/s2_n*_*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1]); n=d["n_gpus"]
        if n==1: base=d["value"]
        print(f, "%.4f ms" % d["ms_per_step"], "%.1fM/s" % (d["value"]/1e6), "eff %.3f" % (d["value"]/(n*base)), "e2e %.1fM/s" % (d["e2e"]["value"]/1e6))
    except Exception as e: print(f, "ERR", e)
PY
