for cap in 1048576 26214400 4194304; do
 B200STEP_BUCKET_CAP_BYTES=$cap python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29900 bench.py --gpus 8 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bk_$cap.json 2> gpurun_out/bk_$cap.err
 python -c "
import json; d=json.loads([l for l in open('gpurun_out/bk_$cap.json') if l.startswith('{')][-1]); print($cap, d['ms_per_step'], d['value']/1e6, d['e2e']['value']/1e6)"
done
