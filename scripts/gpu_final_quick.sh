mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_golden_gpu.py tests/test_zz_no_alloc.py tests/test_step_parity_gpu.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
timeout 60 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/final_quick_n1.json 2> gpurun_out/final_quick_n1.err; echo "bench rc=$?"
python -c "
import json; d=json.loads([l for l in open('gpurun_out/final_quick_n1.json') if l.startswith('{')][-1]); print('%.4f ms %.1fM/s e2e %.1fM/s launches %d' % (d['ms_per_step'], d['value']/1e6, d['e2e']['value']/1e6, d['gpu_launches']))"
