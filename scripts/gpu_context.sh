# 1 GPU: stock-PyTorch context rows, Zipf secondary dataset row, default bench on the same box
mkdir -p gpurun_out
python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/ctx_b200_t0.json 2> gpurun_out/ctx_b200_t0.err; echo "b200 t0 rc=$?"
python bench.py --steps 50 --warmup 5 --no-cpu-baseline --zipf-alpha 1.05 > gpurun_out/ctx_b200_t0_zipf.json 2> gpurun_out/ctx_b200_t0_zipf.err; echo "b200 t0 zipf rc=$?"
python bench.py --steps 30 --warmup 5 --no-cpu-baseline --workload t1_bf16_b65536 --zipf-alpha 1.05 > gpurun_out/ctx_b200_t1_zipf.json 2> gpurun_out/ctx_b200_t1_zipf.err; echo "b200 t1 zipf rc=$?"
timeout 300 python scripts/gpu_torch_eager_context.py --arch tabular_mlp --precision bf16 --batch 65536 > gpurun_out/ctx_eager_t0_bf16.json 2> gpurun_out/ctx_eager_t0_bf16.err; echo "eager t0 bf16 rc=$?"
timeout 300 python scripts/gpu_torch_eager_context.py --arch tabular_mlp --precision tf32 --batch 4096 > gpurun_out/ctx_eager_t0_tf32_b4096.json 2> gpurun_out/ctx_eager_t0_tf32.err; echo "eager t0 tf32 rc=$?"
timeout 300 python scripts/gpu_torch_eager_context.py --arch tabular_mlp_wide --precision bf16 --batch 65536 --steps 20 > gpurun_out/ctx_eager_t1_bf16.json 2> gpurun_out/ctx_eager_t1_bf16.err; echo "eager t1 bf16 rc=$?"
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/ctx_*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "%.4f ms" % d["ms_per_step"], "%.2fM/s" % (d["value"]/1e6), "loss", d.get("final_loss"))
    except Exception as e: print(f, "ERR", e); print(open(f.replace(".json",".err")).read()[-1500:])
PY
