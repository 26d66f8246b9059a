"""Turn an `ncu --set full` capture of bench.py (T0 bf16) into per-kernel summaries:
   profiles/<tag>_ncu_step_kernels.md  and  profiles/ncu_traffic.json (dram bytes per launch, keyed like bench.py's
   kernel labels).  Usage: python scripts/ncu_to_traffic.py gpurun_out/prof_step.ncu-rep r01 t0_bf16_b65536"""
import csv, io, json, os, subprocess, sys

rep, tag, workload = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, rows = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
def val(r, k):
    try: return float(r[ix[k]].replace(",", ""))
    except Exception: return float("nan")
def to_bytes(r, k):
    v, u = val(r, k), units[ix[k]].lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
def to_us(r, k):
    v, u = val(r, k), units[ix[k]].lower()
    return v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1)
# label one step: start at the first preprocess kernel, stop at advance_step
names = [r[ix["Kernel Name"]] for r in rows]
start = next(i for i, n in enumerate(names) if "preprocess" in n)
gemm_labels = {"fwd": ["linear_fwd[65536x512x448]", "linear_fwd[65536x512x512]", "linear_fwd[65536x1000x512]"],
               "dgrad": ["linear_dgrad[65536x1000x512]+dbias", "linear_dgrad[65536x512x512]+dbias", "linear_dgrad[65536x512x448]"],
               "wgrad": ["linear_wgrad[65536x1000x512]", "linear_wgrad[65536x512x512]", "linear_wgrad[65536x512x448]"]}
cnt = {"fwd": 0, "dgrad": 0, "wgrad": 0}
out, traffic = [], {}
for r in rows[start:]:
    n = r[ix["Kernel Name"]]
    if "wgrad_grouped_kernel" in n:
        label = "linear_wgrad_grouped[65536x3 layers]"
    elif "gemm_kernel" in n:
        kind = "fwd" if ", 0, 2>" in n.replace("(int)", "").replace("(bool)", "") and False else None
        args = n.replace("(int)", "").replace("(bool)", "")
        epi = args.split("<")[1].split(">")[0].split(",")[3].strip()
        kind = {"0": "fwd", "1": "dgrad", "2": "wgrad"}[epi]
        label = gemm_labels[kind][cnt[kind] % 3]; cnt[kind] += 1
    elif "ce_loss" in n: label = "ce_loss_grad"
    elif "preprocess" in n: label = "preprocess"
    elif "emb_bwd" in n: label = "emb_bwd"
    elif "sgd_kernel" in n: label = "sgd_step"
    elif "reduce_partials" in n: label = "reduce_partials"
    elif "advance_step" in n: label = "advance_step"
    else: label = n[:30]
    rec = {"label": label, "us": to_us(r, "gpu__time_duration.sum"),
           "dram_read": to_bytes(r, "dram__bytes_read.sum"), "dram_write": to_bytes(r, "dram__bytes_write.sum"),
           "dram_pct": val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
           "tensor_pct": val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
           "sm_pct": val(r, "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
           "regs": val(r, "launch__registers_per_thread"), "grid": val(r, "launch__grid_size")}
    out.append(rec)
    if label not in traffic: traffic[label] = rec["dram_read"] + rec["dram_write"]
    if label == "advance_step": break
os.makedirs("profiles", exist_ok=True)
with open(f"profiles/{tag}_ncu_step_kernels.md", "w") as f:
    f.write(f"# One training step under `ncu --set full --clock-control none` ({workload}, 1 GPU)\n\n")
    f.write("Per-launch figures, cold-ish cache and serialised by the profiler: compare shares, not absolutes.\n\n")
    f.write("| kernel | us | dram read MB | dram write MB | dram % | tensor pipe % | sm % | regs | grid |\n|---|---|---|---|---|---|---|---|---|\n")
    for r in out:
        f.write("| %s | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %d | %d |\n" % (r["label"], r["us"], r["dram_read"] / 1e6,
                r["dram_write"] / 1e6, r["dram_pct"], r["tensor_pct"], r["sm_pct"], r["regs"], r["grid"]))
    f.write("\nTotal: %.1f us, dram traffic %.1f MB\n" % (sum(r["us"] for r in out), sum(r["dram_read"] + r["dram_write"] for r in out) / 1e6))
tp = "profiles/ncu_traffic.json"
allt = json.load(open(tp)) if os.path.exists(tp) else {}
allt[workload] = traffic
json.dump(allt, open(tp, "w"), indent=1)
print(open(f"profiles/{tag}_ncu_step_kernels.md").read())
