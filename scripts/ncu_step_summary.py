"""Summarise an `ncu --set full` capture of bench.py: one training step (from one K-pre launch to the next), per launch
duration, DRAM bytes, pipe utilisation -> profiles/<tag>_ncu_step_kernels.md and profiles/ncu_traffic.json (DRAM bytes
per launch keyed like bench.py's kernel labels).
Usage: python scripts/ncu_step_summary.py gpurun_out/x.ncu-rep r02 t0_bf16_b65536"""
import csv, io, json, os, subprocess, sys

rep, tag, workload = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, rows = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}


def val(r, k):
    try:
        return float(r[ix[k]].replace(",", ""))
    except Exception:
        return float("nan")


def scaled(r, k, table):
    return val(r, k) * table.get(units[ix[k]].lower(), 1)


BYTES = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}
TIME = {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}
names = [r[ix["Kernel Name"]] for r in rows]
pre = [i for i, n in enumerate(names) if "preprocess" in n]
assert pre, "capture holds no K-pre launch"
if len(pre) >= 2:
    step = rows[pre[-2]:pre[-1]]       # the last complete step in the capture
else:                                  # one K-pre launch: the step runs to the optimizer launch that follows it
    end = next(i for i in range(pre[0], len(names)) if "sgd_kernel" in names[i] or "adam_kernel" in names[i])
    step = rows[pre[0]:end + 1]


def short(n):
    for key in ("preprocess_fast_kernel", "preprocess_kernel", "linear_ce_kernel", "wgrad_grouped_kernel",
                "reduce_partials_kernel", "emb_bwd_kernel", "sgd_kernel", "adam_kernel", "gemm_kernel", "advance_step"):
        if key in n:
            return key
    return n.split("(")[0][-40:]


seq = [short(r[ix["Kernel Name"]]) for r in step]
# bench.py labels in launch order (T0 / T1 default schedule at world 1)
labels, fwd, dgr = [], 0, 0
for i, s in enumerate(seq):
    labels.append(s)
out, traffic = [], {}
out.append(f"# One training step under `ncu --set full --clock-control none` ({workload}, 1 GPU, round {tag})\n")
out.append("Per-launch figures, cold-ish caches and serialised by the profiler: compare shares, not absolutes.\n")
out.append("| # | kernel | us | dram read MB | dram write MB | dram % of 6.55 TB/s | tensor pipe % | issue active % | regs | grid | cluster |")
out.append("|---|---|---|---|---|---|---|---|---|---|---|")
tot = 0.0
for i, r in enumerate(step):
    us = scaled(r, "gpu__time_duration.sum", TIME)
    rd, wr = scaled(r, "dram__bytes_read.sum", BYTES), scaled(r, "dram__bytes_write.sum", BYTES)
    tot += us
    traffic[f"{i}:{seq[i]}"] = rd + wr
    out.append("| %d | %s | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %d | %d | %d |" % (
        i, seq[i], us, rd / 1e6, wr / 1e6, 100.0 * (rd + wr) / (us * 1e-6) / 6550.7e9,   # of the measured copy bandwidth (MEASURED_PEAKS.json)
        val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
        val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"), val(r, "launch__registers_per_thread"),
        val(r, "launch__grid_size"), val(r, "launch__cluster_size") if "launch__cluster_size" in ix else 0))
out.append(f"\nSum of the {len(step)} launches: {tot:.1f} us; DRAM traffic of the step: "
           f"{sum(traffic.values()) / 1e9:.3f} GB.")
os.makedirs("profiles", exist_ok=True)
open(f"profiles/{tag}_ncu_step_kernels.md", "w").write("\n".join(out) + "\n")
tp = "profiles/ncu_traffic.json"
blob = json.load(open(tp)) if os.path.exists(tp) else {}
blob[workload + "@" + tag] = traffic
json.dump(blob, open(tp, "w"), indent=1)
print("\n".join(out))
