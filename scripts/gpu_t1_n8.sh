timeout 600 python -m pytest tests/test_task_gpu.py tests/test_step_parity_gpu.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -2
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29910 bench.py --gpus 8 --steps 30 --warmup 5 --workload t1_bf16_b65536 --no-cpu-baseline > gpurun_out/t1b_n8.json 2> gpurun_out/t1b_n8.err; echo rc=$?; tail -2 gpurun_out/t1b_n8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29911 bench.py --gpus 8 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/t0b_n8.json 2> gpurun_out/t0b_n8.err; echo rc=$?
python - <<PY
import json
for f in ["t1b_n8","t0b_n8"]:
    d=json.loads([l for l in open("gpurun_out/%s.json"%f) if l.startswith("{")][-1]); print(f, d["ms_per_step"], d["value"]/1e6, d["e2e"]["value"]/1e6, d["final_loss"])
PY
