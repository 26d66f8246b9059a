# N GPUs: SM-budget + step parity tests, then bench: prefetched K-pre off / on with different SM reserves
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_step_parity_gpu.py tests/test_task_gpu.py -x -q -m gpu -p no:cacheprovider -k "sm_budget or step_parity or task or prefetch or preprocess" 2>&1 | tail -5
run() { tag=$1; pf=$2; rs=$3; if [ $N -gt 1 ]; then L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29860+RANDOM%100))"; else L=python; fi
  B200STEP_PREFETCH_PRE=$pf B200STEP_PREFETCH_RESERVE=$rs timeout 300 $L bench.py --gpus $N --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/pf_n${N}_$tag.json 2> gpurun_out/pf_n${N}_$tag.err; echo "$tag rc=$?"; }
run a_off 0 0; run b_on_r0 1 0; run c_on_r32 1 32; run d_on_r16 1 16; run e_on_r48 1 48; run f_off 0 0; run g_on_r32 1 32
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/pf_n${N}_*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "%.4f ms" % d["ms_per_step"], "%.1fM/s" % (d["value"]/1e6), "e2e %.1fM/s" % (d["e2e"]["value"]/1e6), "prefetch", d["config"].get("prefetch_pre"), "loss", d["final_loss"])
    except Exception as e: print(f, "ERR", e); print(open(f.replace(".json",".err")).read()[-1200:])
PY
