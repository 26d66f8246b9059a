# Round-2 first item (N >= 2 GPUs): verify K-opt-peer on real peers, then A/B it against the NCCL path.
#   gpurun --gpus 2 --timeout 600 -- 'bash scripts/gpu_peer_opt_verify.sh 2'
N=${1:-2}
mkdir -p gpurun_out
export B200STEP_PEER_TIMEOUT_S=20          # fail fast while this is unverified
B200STEP_TEST_PEER=1 CUDA_LAUNCH_BLOCKING=0 timeout 300 python -m pytest tests/test_task_gpu.py -x -q -m gpu -p no:cacheprovider -k peer 2>&1 | tail -30
run() { tag=$1; peer=$2; B200STEP_PEER_OPT=$peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29930+RANDOM%50)) bench.py --gpus $N --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/peer_n${N}_$tag.json 2> gpurun_out/peer_n${N}_$tag.err; echo "$tag rc=$?"; }
run a_nccl 0; run b_peer 1; run c_nccl 0; run d_peer 1
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/peer_n${N}_*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "%.4f ms" % d["ms_per_step"], "%.1fM/s" % (d["value"]/1e6), "e2e %.1fM/s" % (d["e2e"]["value"]/1e6), "peer", d["config"].get("peer_optimizer"), "loss", d["final_loss"])
    except Exception as e: print(f, "ERR", e); print(open(f.replace(".json",".err")).read()[-2000:])
PY
