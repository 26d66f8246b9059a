#!/bin/bash
# 2-GPU verification: task.py tests (NCCL path), then bench A/B lines.
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_task_gpu.py -m gpu -q -x > gpurun_out/w2_tests.log 2>&1; echo rc=$? >> gpurun_out/w2_tests.log
run() { name=$1; shift; env "$@" python bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/w2_$name.json 2> gpurun_out/w2_$name.err; echo "$name rc=$?"; }
run default X=1
run grouped B200STEP_GROUPED_WGRAD=1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/w2_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'ms/step %.4f'%d['ms_per_step'], 'value %.1fM'%(d['value']/1e6), 'parity', d.get('parity',{}).get('ok'), d.get('parity',{}).get('loss_rel_max'), d.get('parity',{}).get('param_rel_l2_max'), d.get('parity',{}).get('ranks_bit_identical'))
    except Exception as e: print(f,'ERR',e)
PY
