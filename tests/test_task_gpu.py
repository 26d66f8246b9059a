"""End-to-end drop-in checks on the GPU: `python task.py ...` launched the way the notebook launches it
(pytorch-pipeline.ipynb:159-163), single GPU and — when the box has >= 2 GPUs — two ranks over NCCL, compared
with the CPU oracle's DDP restatement (oracle/reference_step.py)."""
import dataclasses
import os
import re
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from b200step.data import make_synthetic  # noqa: E402
from b200step.model import ARCHS  # noqa: E402
from oracle.reference_step import OracleTrainer  # noqa: E402

HIDDEN, CARD, CLASSES = (64, 32), 50, 40
SPEC = dataclasses.replace(ARCHS["tabular_mlp"], cardinality=CARD, hidden=HIDDEN, num_classes=CLASSES)


def _sd(spec):
    d = dataclasses.asdict(spec)
    d.pop("name")
    return d


def _launch(ngpus, tmp, rows, batch, epochs, port, precision="bf16", extra=(), env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    env.update({"WORLD_SIZE": "1", "RANK": "0", "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port),
                "CUDA_VISIBLE_DEVICES": ",".join(str(i) for i in range(ngpus))})
    cmd = [sys.executable, os.path.join(ROOT, "task.py"), "--dist-url=env://", "--multiprocessing-distributed",
           f"--num_epochs={epochs}", "--local_training", "--arch", "tabular_mlp", "--hidden", "64,32",
           "--cardinality", str(CARD), "--num_classes", str(CLASSES), "--batch_size", str(batch),
           "--synthetic_rows", str(rows), "--model_dir", str(tmp), "--save_final", "--precision", precision,
           *extra]
    return subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600, cwd=ROOT)


def _compare(final, oracle, tol):
    so = oracle.state_dict()
    assert list(final.keys()) == list(so.keys())
    for k in so:
        assert final[k].shape == so[k].shape
        rel = ((final[k] - so[k]).norm() / so[k].norm().clamp_min(1e-12)).item()
        assert rel <= tol, (k, rel)


def test_task_py_single_gpu(tmp_path):
    rows, batch, epochs = 1400, 256, 2
    r = _launch(1, tmp_path, rows, batch, epochs, 29561)
    assert r.returncode == 0, r.stderr[-3000:]
    out = r.stdout
    assert "GPU x WORLD SIZE = 1" in out and "Process group initialized" in out
    assert out.count("training with gpu 0") == epochs * 6          # 5 full batches + the partial one per epoch
    assert re.search(r"Epoch: 0, Accuracy: 0\.\d+, Time: ", out) and "saving model to local folders" in out
    assert re.search(r"Epoch complete: \d{4}_", out) and re.search(r"Training complete: \d{4}_", out)
    ds = make_synthetic(rows, SPEC)
    oracle = OracleTrainer(_sd(SPEC), ds, world=1, batch_size=batch, precision="bf16")
    # the epoch-0 checkpoint is written BEFORE training (task.py:279): it must equal the seeded initial model exactly
    saved = torch.load(os.path.join(tmp_path, "resnet_distributed.pth"))
    for k, v in oracle.state_dict().items():
        assert torch.equal(saved[k], v), k
    for _ in range(epochs * 6):
        lo = oracle.step()[0]
    lg = float(re.search(r"final local loss \(rank 0\): ([0-9.eE+-]+)", out).group(1))
    assert abs(lo - lg) <= 2e-3 * max(1.0, abs(lo))
    _compare(torch.load(os.path.join(tmp_path, "resnet_distributed.pth.final")), oracle, 1e-2)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_task_py_two_gpus_matches_ddp_oracle(tmp_path):
    """DDP semantics (task.py:189): every rank ends with the same parameters = SGD on the mean gradient."""
    rows, batch, epochs = 2048, 256, 3
    r = _launch(2, tmp_path, rows, batch, epochs, 29562)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "GPU x WORLD SIZE = 2" in r.stdout
    ds = make_synthetic(rows, SPEC)
    oracle = OracleTrainer(_sd(SPEC), ds, world=2, batch_size=batch, precision="bf16")
    for _ in range(epochs * 4):
        l0 = oracle.step()[0]
    lg = float(re.search(r"final local loss \(rank 0\): ([0-9.eE+-]+)", r.stdout).group(1))
    assert abs(l0 - lg) <= 2e-3 * max(1.0, abs(l0))
    _compare(torch.load(os.path.join(tmp_path, "resnet_distributed.pth.final")), oracle, 1e-2)
