"""Generate the golden vectors under tests/golden/ from the CPU oracle (run here, in the build container):

    python tests/golden/make_golden.py

The reference ships no fixtures, so these are produced by the oracle restatement running the
installed torch 2.11.0 CPU kernels — the same library calls the reference makes.  Each file holds, for a small
seeded case: the sampler indices, the bit pattern of the preprocessed first batch, the per-step losses and a
checksum (sum, sum of squares) of every parameter tensor after the last step.  The GPU tests compare the CUDA
path with these; tests/test_oracle.py checks the oracle still reproduces them.
"""
import dataclasses
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "kubeflow-v2-distributed-pytorch_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from b200step.data import make_synthetic  # noqa: E402
from b200step.model import ARCHS  # noqa: E402
from oracle.reference_step import OracleTrainer, standardise  # noqa: E402

SMALL = dataclasses.replace(ARCHS["tabular_mlp"], cardinality=50, hidden=(64, 32), num_classes=40)

CASES = {
    "small_fp32_sgd": dict(spec=SMALL, rows=2048, batch=256, steps=16, precision="fp32", optimizer="sgd", lr=0.1),
    "small_bf16_sgd": dict(spec=SMALL, rows=2048, batch=256, steps=16, precision="bf16", optimizer="sgd", lr=0.1),
    "small_bf16_adam": dict(spec=SMALL, rows=2048, batch=256, steps=16, precision="bf16", optimizer="adam", lr=1e-3),
}


def spec_dict(spec):
    d = dataclasses.asdict(spec)
    d.pop("name")
    return d


def run_case(case):
    spec = case["spec"]
    ds = make_synthetic(case["rows"], spec)
    tr = OracleTrainer(spec_dict(spec), ds, world=1, batch_size=case["batch"], seed=0,
                       optimizer=case["optimizer"], precision=case["precision"], lr=case["lr"])
    rows0 = tr.batches[0][0]
    feats = tr.model.features(standardise(ds.dense[rows0], ds.mean, ds.std), ds.cat[rows0]).detach()
    if case["precision"] == "bf16":
        x0_bits = feats.to(torch.bfloat16).view(torch.int16).clone()
    else:
        x0_bits = feats.view(torch.int32).clone()
    losses = torch.tensor([tr.step()[0] for _ in range(case["steps"])], dtype=torch.float64)
    sd = tr.state_dict()
    chk = torch.tensor([[v.double().sum().item(), (v.double() ** 2).sum().item()] for v in sd.values()],
                       dtype=torch.float64)
    return {"row_idx": torch.cat(tr.batches[0]), "x0_bits": x0_bits, "losses": losses,
            "param_checksums": chk, "param_keys": list(sd.keys())}


if __name__ == "__main__":
    for name, case in CASES.items():
        out = run_case(case)
        torch.save(out, os.path.join(HERE, f"{name}.pt"))
        print(name, out["losses"][:3].tolist(), "...", out["losses"][-1].item())
