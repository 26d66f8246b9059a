"""A/B timing of TrainStep variants in one process (T0 bf16 B=65536, graph replay, CUDA events)."""
import dataclasses, os, sys, time
ROOT = os.path.join(os.path.dirname(__file__), "..")
sys.path.insert(0, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")); sys.path.insert(0, ROOT)
import torch
from b200step.data import make_synthetic
from b200step.engine import StepConfig, TrainStep
from b200step.model import ARCHS, init_parameters
spec = ARCHS["tabular_mlp"]; B = 65536
ds = make_synthetic(B * 8, spec)
def run(steps=200, **kw):
    eng = TrainStep(spec, init_parameters(spec, 0), ds, torch.arange(B * 8), StepConfig(batch_size=B, **kw))
    for _ in range(5): eng.step()
    eng.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    out = []
    for rep in range(3):
        e0.record(eng.main)
        for _ in range(steps): eng.step()
        e1.record(eng.main); eng.synchronize()
        out.append(e0.elapsed_time(e1) / steps)
    eng.close()
    return out
import json
variants = json.loads(os.environ.get("AB_VARIANTS", '[["default", {}]]'))
for name, kw in variants:
    print(name, ["%.4f" % x for x in run(steps=int(os.environ.get("AB_STEPS", "50")), **kw)], flush=True)
