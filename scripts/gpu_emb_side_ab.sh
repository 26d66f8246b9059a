#!/bin/bash
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_step_parity_gpu.py tests/test_golden_gpu.py tests/test_full_size_gpu.py -m gpu -q -x --timeout 600 > gpurun_out/emb_side_tests.log 2>&1; echo rc=$? >> gpurun_out/emb_side_tests.log
for v in 1 0 1 0; do
B200STEP_EMB_SIDE=$v python bench.py --steps 40 --warmup 5 --no-cpu-baseline --sustain-seconds 0 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('emb_side=$v ms/step %.4f value %.1fM parity %s' % (d['ms_per_step'], d['value']/1e6, d['parity']['ok']))" >> gpurun_out/emb_side_ab.txt
done
