mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
run() { n=$1; wl=$2; tag=$3; if [ $n -eq 1 ]; then python bench.py --gpus 1 --steps 50 --warmup 5 --workload $wl > gpurun_out/fin_${tag}_n$n.json 2> gpurun_out/fin_${tag}_n$n.err; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800+n)) bench.py --gpus $n --steps 50 --warmup 5 --workload $wl > gpurun_out/fin_${tag}_n$n.json 2> gpurun_out/fin_${tag}_n$n.err; fi; echo "$tag n=$n rc=$?"; }
for n in 1 2 4 8; do run $n t0_bf16_b65536 t0; done
run 1 t1_bf16_b65536 t1; run 8 t1_bf16_b65536 t1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29850 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/fin_ref_n2.json 2> gpurun_out/fin_ref_n2.err; echo "ref n2 rc=$?"
python - <<PY
import json,glob
base={}
for f in sorted(glob.glob("gpurun_out/fin_t*_n*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1]); n=d["n_gpus"]; w=d["config"]["workload"]
        if n==1: base[w]=d["value"]
        print(f, "%.4f ms" % d["ms_per_step"], "%.1fM/s" % (d["value"]/1e6), "eff %.3f" % (d["value"]/(n*base[w])), "e2e %.1fM/s" % (d["e2e"]["value"]/1e6), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
    except Exception as e: print(f, "ERR", e)
print(open("gpurun_out/fin_ref_n2.json").read()[:300])
PY
