#!/bin/bash
cd "$(dirname "$0")/.."
for v in 0 1; do
  B200STEP_L2_ORDER=$v python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-parity --sustain-seconds 0 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('l2_order=$v ms/step %.4f' % d['ms_per_step'], [(k['name'][:22], round(k['ms']*1e3,1)) for k in d['kernels']])"
done
