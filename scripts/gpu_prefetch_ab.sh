#!/bin/bash
# A/B of the overlapped K-pre (world = 1): off, and 1 / 2 / 3 / 4 CTA slots per SM.  Run under gpurun.
cd "$(dirname "$0")/.."
for v in off 1 2 3 4; do
  if [ "$v" = off ]; then export B200STEP_PREFETCH_PRE=0; unset B200STEP_PREFETCH_CTAS; else export B200STEP_PREFETCH_PRE=1 B200STEP_PREFETCH_CTAS=$v; fi
  python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-parity --sustain-seconds 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('prefetch=$v', 'ms/step %.4f' % d['ms_per_step'], 'value %.1fM' % (d['value']/1e6), 'e2e %.1fM' % (d['e2e']['value']/1e6), 'launches/step', d['gpu_launches']//d['steps'])
"
done
