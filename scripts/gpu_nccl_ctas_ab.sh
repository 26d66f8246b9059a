#!/bin/bash
# A/B of NCCL's CTA budget (NCCL_MAX_CTAS) at N GPUs on the T0 step: fewer NCCL CTAs leave the persistent GEMM grids alone.
N=${1:-2}
cd "$(dirname "$0")/.."
for c in default 2 4 8 16; do
  if [ "$c" = default ]; then unset NCCL_MAX_CTAS; else export NCCL_MAX_CTAS=$c; fi
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29300 + RANDOM % 300)) \
    bench.py --gpus $N --steps 40 --warmup 5 --sustain-seconds 0 --no-parity 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('NCCL_MAX_CTAS=$c N=$N ms/step %.4f value %.1fM' % (d['ms_per_step'], d['value']/1e6))"
done
