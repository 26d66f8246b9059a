"""Fixtures made by running the REFERENCE'S OWN functions in the build container:

    python tests/golden/make_reference_pins.py        # writes tests/golden/reference_pins.pt

/root/reference/task.py is imported as it lies (only `google.cloud.storage`, which is not installed, is stubbed)
and two of its functions, which are plain CPU-runnable Python, are executed:

  task.set_random_seeds(seed)                  task.py:22-28   -> seeds torch / numpy / random
  task.evaluate(model, "cpu", test_loader)     task.py:30-46   -> top-1 accuracy over a shuffle=False loader of
                                                                  batch size 128 (task.py:267)

What is pinned with them:
  * `init`:  sha256 of every parameter tensor of the tabular model constructed right after
             task.set_random_seeds(0) (the order task.py:162 -> :171 builds its model in).  b200step's
             init_parameters() and an untrained TrainStep.state_dict() must reproduce them bit for bit.
  * `eval`:  the accuracy task.evaluate returns for that seeded model over the synthetic set, plus — captured with
             a forward hook during that very call — each row's argmax (torch.max(outputs.data, 1), task.py:40) and
             the gap between its two largest logits.  TrainStep.evaluate() must give the same prediction for every
             row whose gap is wider than the bf16 rounding of the logits (near-ties may flip), and an accuracy that
             differs by no more than the near-tie rows allow.
The model class (TabularMLP) and the synthetic data have no counterpart in the reference: the
reference contributes the seeding, the evaluation loop and, through them, torch's own initialisers and modules.
"""
import hashlib
import importlib.util
import os
import sys
import types

import torch
from torch.utils.data import DataLoader, TensorDataset

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def load_reference(path="/root/reference/task.py"):
    for name in ("google", "google.cloud", "google.cloud.storage"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["google"].cloud = sys.modules["google.cloud"]
    sys.modules["google.cloud"].storage = sys.modules["google.cloud.storage"]
    spec = importlib.util.spec_from_file_location("reference_task", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def tensor_sha(t: torch.Tensor) -> str:
    return hashlib.sha256(t.detach().contiguous().cpu().numpy().tobytes()).hexdigest()


class PackedInput(torch.nn.Module):
    """task.evaluate feeds the model ONE tensor per batch (`model(images)`, task.py:39): the standardised dense
    columns and the categorical ids (exact in fp32 below 2^24) travel side by side in it."""

    def __init__(self, model, num_dense):
        super().__init__()
        self.model, self.nd = model, num_dense

    def forward(self, packed):
        return self.model(packed[:, : self.nd], packed[:, self.nd:].to(torch.int64))


PIN_SPECS = {
    # name -> (ModelSpec overrides, dataset rows)
    "small": (dict(cardinality=50, hidden=(64, 32), num_classes=40), 1024),
    "t0": (dict(), 2048),
}


def make_pins(task):
    import dataclasses

    from b200step.data import make_synthetic
    from b200step.model import ARCHS
    from oracle.reference_step import TabularMLP, standardise

    out = {"torch": str(torch.__version__)}
    for name, (over, rows) in PIN_SPECS.items():
        spec = dataclasses.replace(ARCHS["tabular_mlp"], **over)
        sd = dataclasses.asdict(spec)
        sd.pop("name")
        task.set_random_seeds(0)                                    # task.py:162 -> :22-28
        model = TabularMLP(**sd)                                    # task.py:171 (model built under that seed)
        init = {k: tensor_sha(v) for k, v in model.state_dict().items()}
        ds = make_synthetic(rows, spec)
        packed = torch.cat([standardise(ds.dense, ds.mean, ds.std), ds.cat.to(torch.float32)], 1)
        loader = DataLoader(TensorDataset(packed, ds.label), batch_size=128, shuffle=False)   # task.py:267
        wrapped = PackedInput(model, spec.num_dense)
        seen = []
        hook = wrapped.register_forward_hook(lambda m, i, o: seen.append(o.detach().clone()))
        acc = task.evaluate(model=wrapped, device="cpu", test_loader=loader)                 # task.py:30-46
        hook.remove()
        logits = torch.cat(seen, 0)
        top2 = logits.topk(2, dim=1)
        out[name] = {"spec": sd, "rows": rows, "init_sha256": init, "accuracy": float(acc),
                     "argmax": top2.indices[:, 0].to(torch.int16), "gap": (top2.values[:, 0] - top2.values[:, 1]),
                     "logit_abs_max": logits.abs().max(dim=1).values}
        if name == "small":
            # the same call on a TRAINED parameter set (20 fp32 steps of torch's own autograd + SGD through the
            # oracle trainer), stored here so that the GPU engine can be loaded with exactly these parameters
            from oracle.reference_step import OracleTrainer

            tr = OracleTrainer(sd, ds, world=1, batch_size=256, precision="fp32")
            for _ in range(20):
                tr.step()
            wrapped = PackedInput(tr.model, spec.num_dense)
            seen = []
            hook = wrapped.register_forward_hook(lambda m, i, o: seen.append(o.detach().clone()))
            acc = task.evaluate(model=wrapped, device="cpu", test_loader=loader)
            hook.remove()
            logits = torch.cat(seen, 0)
            top2 = logits.topk(2, dim=1)
            out["small_trained"] = {"spec": sd, "rows": rows, "accuracy": float(acc),
                                    "params": {k: v.detach().clone() for k, v in tr.model.state_dict().items()},
                                    "argmax": top2.indices[:, 0].to(torch.int16),
                                    "gap": (top2.values[:, 0] - top2.values[:, 1]),
                                    "logit_abs_max": logits.abs().max(dim=1).values}
    return out


if __name__ == "__main__":
    pins = make_pins(load_reference())
    torch.save(pins, os.path.join(HERE, "reference_pins.pt"))
    for k, v in pins.items():
        if isinstance(v, dict):
            print(k, "accuracy", v["accuracy"], "rows", v["rows"], "min gap", float(v["gap"].min()),
                  "rows with gap < 2% of |logit|max:", int((v["gap"] < 0.02 * v["logit_abs_max"]).sum()))
