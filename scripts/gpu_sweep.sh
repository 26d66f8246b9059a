# N GPUs: config-5 all-reduce sweep (+ stock PyTorch DDP context row at the same N)
N=${1:-2}
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29811 scripts/gpu_allreduce_sweep.py > gpurun_out/sweep_w$N.json 2> gpurun_out/sweep_w$N.err; echo "sweep rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29812 scripts/gpu_torch_eager_context.py --arch tabular_mlp --precision bf16 --batch 65536 > gpurun_out/ctx_eager_t0_bf16_n$N.json 2> gpurun_out/ctx_eager_t0_bf16_n$N.err; echo "eager ddp rc=$?"
tail -c 600 gpurun_out/sweep_w$N.err
python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/sweep_w$N.json") if l.startswith("{")][-1])
    print("pair_ms", d["gemm_pair_ms"])
    for r in d["rows"]:
        print("S=%4d MB bk=%2d comm %.3f gemm %.3f both %.3f exposed %.3f busbw %.0f GB/s (%.2f)" % (r["S_MB"], r["buckets"], r["comm_alone_ms"], r["gemm_alone_ms"], r["both_ms"], r["exposed_ms"], r["busbw_alone_gbs"], r["busbw_alone_frac_of_900"]))
except Exception as e: print("ERR", e)
try:
    d=json.loads([l for l in open("gpurun_out/ctx_eager_t0_bf16_n$N.json") if l.startswith("{")][-1]); print("eager ddp n=$N", d["ms_per_step"], d["value"]/1e6)
except Exception as e: print("ERR", e)
PY
