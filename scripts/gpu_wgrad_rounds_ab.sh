#!/bin/bash
cd "$(dirname "$0")/.."
for r in 2 3 4 6 8; do
  B200STEP_WGRAD_ROUNDS=$r python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-parity --sustain-seconds 0 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
w=[k for k in d['kernels'] if 'wgrad' in k['name']][0]
print('rounds=$r ms/step %.4f wgrad %.1f us %.0f TF' % (d['ms_per_step'], w['ms']*1e3, w['tflops']))"
done
