#!/bin/bash
cd "$(dirname "$0")/.."
bash scripts/gpu_window_ab.sh 8
B200STEP_NCCL_WINDOW=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29777 \
  bench.py --gpus 8 --steps 30 --warmup 5 --sustain-seconds 0 --workload t1_bf16_b65536 > gpurun_out/win1_t1_n8.json 2> gpurun_out/win1_t1_n8.err
python - <<'PY'
import json
try:
    d=json.loads([l for l in open("gpurun_out/win1_t1_n8.json") if l.startswith("{")][-1])
    print("T1 window=1 N=8 ms/step %.4f value %.1fM parity %s" % (d["ms_per_step"], d["value"]/1e6, d["parity"]["ok"]))
except Exception as e: print("T1 ERR", e); print(open("gpurun_out/win1_t1_n8.err").read()[-1500:])
PY
