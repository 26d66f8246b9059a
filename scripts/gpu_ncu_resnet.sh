#!/bin/bash
# per-launch device times of the ResNet-18 step under ncu (serialised, cold cache): which kernels carry the step
cd "$(dirname "$0")/.."
WL=${1:-resnet18_bf16_b1024}
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ncu_resnet_launches.csv \
  python bench.py --workload $WL --steps 2 --warmup 3 --no-parity --no-cpu-baseline --sustain-seconds 0 \
  > gpurun_out/ncu_resnet.log 2>&1; echo rc=$? >> gpurun_out/ncu_resnet.log
python - <<'PY'
import csv, collections, re
rows = []
with open("gpurun_out/ncu_resnet_launches.csv") as fh:
    lines = [l for l in fh if l.startswith('"')]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") == "gpu__time_duration.sum":
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
        rows.append((r["Kernel Name"], us))
# the first eager step = launches from the first augment kernel to the second one
idx = [i for i, (n, _) in enumerate(rows) if "augment" in n]
step = rows[idx[0]:idx[1]]
agg = collections.OrderedDict()
for n, us in step:
    k = re.sub(r"<.*", "", n.split("(")[0])
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += us
tot = sum(us for _, us in step)
with open("gpurun_out/ncu_resnet_summary.md", "w") as out:
    out.write(f"launches per step {len(step)}, summed device time {tot:.1f} us (serialised under ncu)\n\n| kernel | launches | us per step | share |\n|---|---|---|---|\n")
    for k, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.write(f"| {k} | {c} | {us:.1f} | {us / tot:.3f} |\n")
    out.write("\nlaunch list of one step (us):\n")
    for n, us in step:
        out.write(f"  {us:8.1f}  {n[:90]}\n")
PY
