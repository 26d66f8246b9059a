#!/bin/bash
# DRAM bytes + duration of every launch of one ResNet-18 step (ncu, two metrics passes) -> gpurun_out/ncu_resnet_traffic.json
cd "$(dirname "$0")/.."
WL=${1:-resnet18_bf16_b1024}
B200STEP_SIDE_WGRAD=0 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
  --log-file gpurun_out/ncu_resnet_traffic.csv \
  python bench.py --workload $WL --steps 1 --warmup 3 --no-parity --no-cpu-baseline --sustain-seconds 0 \
  > gpurun_out/ncu_resnet_traffic.log 2>&1; echo rc=$? >> gpurun_out/ncu_resnet_traffic.log
python - "$WL" <<'PY'
import csv, json, re, sys, collections
wl = sys.argv[1]
with open("gpurun_out/ncu_resnet_traffic.csv") as fh:
    lines = [l for l in fh if l.startswith('"')]
per = collections.OrderedDict()
for r in csv.DictReader(lines):
    k = (r["ID"], r["Kernel Name"])
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    if r["Metric Name"].startswith("gpu__time"):
        v = v / 1e3 if u.startswith("n") else (v if u.startswith("u") else v * 1e3)      # -> us
    else:
        v = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    per.setdefault(k, {})[r["Metric Name"]] = v
rows = [(k[1], d.get("gpu__time_duration.sum", 0), d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0)) for k, d in per.items()]
idx = [i for i, r in enumerate(rows) if "augment" in r[0]]
step = rows[idx[0]:idx[1]]
agg = collections.OrderedDict()
for n, us, by in step:
    k = re.sub(r"<.*", "", n.split("(")[0]).replace("void ", "").replace("b200::", "")
    a = agg.setdefault(k, [0, 0.0, 0.0]); a[0] += 1; a[1] += us; a[2] += by
out = {"workload": wl, "launches": len(step), "us": sum(r[1] for r in step), "dram_bytes": sum(r[2] for r in step),
       "kernels": {k: {"launches": a[0], "us": a[1], "dram_bytes": a[2]} for k, a in agg.items()}}
json.dump(out, open("gpurun_out/ncu_resnet_traffic.json", "w"), indent=1)
print(json.dumps({k: v for k, v in out.items() if k != "kernels"}))
PY
