#!/bin/bash
# 8-GPU box: bench.py A/B lines at N = 8 (T0 and T1), then the config-5 all-reduce sweep at W = 8 and W = 4.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { name=$1; shift; n=$1; shift; wl=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) \
    bench.py --gpus $n --steps 30 --warmup 5 --workload $wl --sustain-seconds 0 > gpurun_out/w8_$name.json 2> gpurun_out/w8_$name.err; echo "$name rc=$?"; }
run t0_default 8 t0_bf16_b65536 X=1
run t0_reg 8 t0_bf16_b65536 B200STEP_NCCL_REG=1
run t0_grouped 8 t0_bf16_b65536 B200STEP_GROUPED_WGRAD=1
run t0_grouped_reg 8 t0_bf16_b65536 B200STEP_GROUPED_WGRAD=1 B200STEP_NCCL_REG=1
run t1_default 8 t1_bf16_b65536 X=1
run t1_reg 8 t1_bf16_b65536 B200STEP_NCCL_REG=1
run t0_n4 4 t0_bf16_b65536 X=1
for w in 8 4; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $w --master-addr 127.0.0.1 --master-port $((29900 + w)) scripts/gpu_allreduce_sweep.py > gpurun_out/sweep_w$w.json 2> gpurun_out/sweep_w$w.err; echo "sweep w=$w rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/w8_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][-1]); p=d.get('parity',{})
        print(f, 'ms/step %.4f'%d['ms_per_step'], 'value %.1fM'%(d['value']/1e6), 'e2e %.1fM'%(d['e2e']['value']/1e6), 'parity', p.get('ok'), p.get('loss_rel_max'), p.get('param_rel_l2_max'), p.get('ranks_bit_identical'))
    except Exception as e: print(f,'ERR',e)
for w in (8,4):
    try:
        d=json.loads([l for l in open(f'gpurun_out/sweep_w{w}.json') if l.startswith('{')][-1])
        for r in d['rows']: print('W=%d S=%4d MB comm %.3f both %.3f exposed %.3f busbw %.0f GB/s (%.2f)'%(w,r['S_MB'],r['comm_alone_ms'],r['both_ms'],r['exposed_ms'],r['busbw_alone_gbs'],r['busbw_alone_frac_of_900']))
    except Exception as e: print('sweep',w,'ERR',e)
PY
