# 8 GPUs: default bench with the prefetched K-pre (multi-GPU default) and without it
mkdir -p gpurun_out
run() { tag=$1; pf=$2; B200STEP_PREFETCH_PRE=$pf timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29900+RANDOM%50)) bench.py --gpus 8 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/n8_$tag.json 2> gpurun_out/n8_$tag.err; echo "$tag rc=$?"; }
run a_on 1; run b_off 0; run c_on 1
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/n8_*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "%.4f ms" % d["ms_per_step"], "%.1fM/s" % (d["value"]/1e6), "e2e %.1fM/s" % (d["e2e"]["value"]/1e6), "prefetch", d["config"].get("prefetch_pre"))
        print("   ", " ".join("%s=%.4f" % (k["name"].split("[")[0][:12], k["ms"]) for k in d["kernels"]))
    except Exception as e: print(f, "ERR", e); print(open(f.replace(".json",".err")).read()[-1200:])
PY
