# W GPUs: how many SMs must the GEMM grids leave free for the all-reduce to overlap?  (reserve, NCCL_MAX_CTAS) matrix
N=${1:-2}
mkdir -p gpurun_out
i=0
for cfg in "0:" "16:" "32:" "16:16" "32:32" "8:8" "24:24"; do
  r=${cfg%%:*}; c=${cfg##*:}; i=$((i+1))
  if [ -n "$c" ]; then export NCCL_MAX_CTAS=$c; else unset NCCL_MAX_CTAS; fi
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29820+i)) scripts/gpu_allreduce_sweep.py --min-mb 4 --max-mb 256 --sm-reserve $r > gpurun_out/mx_w${N}_r${r}_c${c:-def}.json 2> gpurun_out/mx_w${N}_r${r}_c${c:-def}.err; echo "r=$r c=$c rc=$?"
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/mx_w${N}_*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "pair %.4f" % d["gemm_pair_ms"])
        for r in d["rows"]:
            print("   S=%4d comm %.3f gemm %.3f both %.3f exposed %.3f busbw %.0f" % (r["S_MB"], r["comm_alone_ms"], r["gemm_alone_ms"], r["both_ms"], r["exposed_ms"], r["busbw_alone_gbs"]))
    except Exception as e: print(f, "ERR", e)
PY
