# 1 -> 8 GPU scaling of bench.py the way the driver launches it
mkdir -p gpurun_out
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then
    python bench.py --gpus 1 --steps 50 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 50 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  echo "n=$n rc=$?"
done
python bench.py --impl reference --gpus 1 --steps 5 --warmup 2 > gpurun_out/ref_n1.json 2> gpurun_out/ref_n1.err; echo "ref rc=$?"
python - <<PY
import json
base=None
for n in (1,2,4,8):
    try:
        d=json.loads([l for l in open(f"gpurun_out/scale_n{n}.json") if l.startswith("{")][-1])
        if n==1: base=d["value"]
        print(n, "%.4f ms" % d["ms_per_step"], "%.1fM/s" % (d["value"]/1e6), "eff %.3f" % (d["value"]/(n*base)), "e2e %.1fM/s" % (d["e2e"]["value"]/1e6), d["clocks"])
    except Exception as e: print(n, "ERR", e)
d=json.loads([l for l in open("gpurun_out/ref_n1.json") if l.startswith("{")][-1]); print("ref", d["value"], d["cpu_baseline"])
PY
