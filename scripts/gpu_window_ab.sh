#!/bin/bash
# A/B at N GPUs: gradient buffer registered as an NCCL symmetric window (default) or not.
N=${1:-2}
cd "$(dirname "$0")/.."
for v in 0 1; do
  B200STEP_NCCL_WINDOW=$v timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29300 + RANDOM % 300)) \
    bench.py --gpus $N --steps 40 --warmup 5 --sustain-seconds 0 > gpurun_out/win${v}_n$N.json 2> gpurun_out/win${v}_n$N.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/win${v}_n$N.json") if l.startswith("{")][-1])
    print("window=$v N=$N ms/step %.4f value %.1fM parity %s" % (d["ms_per_step"], d["value"]/1e6, d["parity"]["ok"]))
except Exception as e: print("window=$v ERR", e); print(open("gpurun_out/win${v}_n$N.err").read()[-1500:])
PY
done
