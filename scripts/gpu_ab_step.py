"""A/B timing of TrainStep variants in one process (T0 bf16 B=65536, graph replay, CUDA events)."""
import dataclasses, os, sys, time
ROOT = os.path.join(os.path.dirname(__file__), "..")
sys.path.insert(0, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")); sys.path.insert(0, ROOT)
import torch
from b200step.data import make_synthetic
from b200step.engine import StepConfig, TrainStep
from b200step.model import ARCHS, init_parameters
spec = ARCHS[os.environ.get("AB_ARCH", "tabular_mlp")]; B = int(os.environ.get("AB_BATCH", "65536"))
PREC = os.environ.get("AB_PRECISION", "bf16")
ds = make_synthetic(B * 8, spec)
import json, statistics
variants = json.loads(os.environ.get("AB_VARIANTS", '[["default", {}]]'))
steps = int(os.environ.get("AB_STEPS", "40"))
reps = int(os.environ.get("AB_REPS", "8"))
engs = []
for name, kw in variants:
    eng = TrainStep(spec, init_parameters(spec, 0), ds, torch.arange(B * 8), StepConfig(batch_size=B, precision=PREC, **kw))
    for _ in range(5): eng.step()
    eng.synchronize()
    engs.append((name, eng, []))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for rep in range(reps):            # interleaved, so clock / thermal drift hits every variant alike
    for name, eng, out in engs:
        e0.record(eng.main)
        for _ in range(steps): eng.step()
        e1.record(eng.main); eng.synchronize()
        out.append(e0.elapsed_time(e1) / steps)
for name, eng, out in engs:
    print("%-28s median %.4f  min %.4f  max %.4f ms" % (name, statistics.median(out), min(out), max(out)), flush=True)
