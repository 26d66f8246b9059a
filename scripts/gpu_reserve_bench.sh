# N GPUs: default bench under (B200STEP_SM_RESERVE, NCCL_MAX_CTAS) combinations
N=${1:-2}; WL=${2:-t0_bf16_b65536}
mkdir -p gpurun_out
i=0
for cfg in "0:" "32:" "16:16" "8:8" "24:24" "0:" "16:" "4:4"; do
  r=${cfg%%:*}; c=${cfg##*:}; i=$((i+1))
  if [ -n "$c" ]; then export NCCL_MAX_CTAS=$c; else unset NCCL_MAX_CTAS; fi
  B200STEP_SM_RESERVE=$r timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29840+i)) bench.py --gpus $N --steps 50 --warmup 5 --no-cpu-baseline --workload $WL > gpurun_out/rb_${WL}_n${N}_${i}_r${r}_c${c:-def}.json 2> gpurun_out/rb_${WL}_n${N}_${i}.err; echo "r=$r c=$c rc=$?"
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/rb_${WL}_n${N}_*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "%.4f ms" % d["ms_per_step"], "%.1fM/s" % (d["value"]/1e6), "e2e %.1fM/s" % (d["e2e"]["value"]/1e6))
    except Exception as e: print(f, "ERR", e)
PY
