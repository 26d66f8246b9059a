"""`python task.py` with the reference's DEFAULT model (--arch resnet18, task.py:63) on the GPU, launched the way the
notebook launches it (pytorch-pipeline.ipynb:159-163): the checkpoint is a torchvision state_dict, and the training
loss follows the torchvision oracle (oracle/resnet_step.py).  Two ranks over NCCL when the box has two GPUs."""
import os
import re
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(ngpus, tmp, rows, batch, port, extra=()):
    env = dict(os.environ)
    env.update({"WORLD_SIZE": "1", "RANK": "0", "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port),
                "CUDA_VISIBLE_DEVICES": ",".join(str(i) for i in range(ngpus))})
    cmd = [sys.executable, os.path.join(ROOT, "task.py"), "--dist-url=env://", "--multiprocessing-distributed",
           "--num_epochs=1", "--local_training", "--batch_size", str(batch), "--synthetic_rows", str(rows),
           "--learning_rate", "0.01", "--model_dir", str(tmp), "--save_final", *extra]
    return subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=900, cwd=ROOT)


def _oracle(rows, batch, world):
    """Mirror the processes' use of torch's global generator: set_random_seeds + model construction, then (rank 0
    only) the evaluation pass's draws over the whole dataset (task.py:279-281), then this epoch's training draws."""
    from b200step.convnet import draw_augmentation, init_state, make_synthetic_images
    from b200step.sampler import rank_indices
    from oracle.resnet_step import ResNetOracle

    ds = make_synthetic_images(rows, seed=0)
    rank_rows = [rank_indices(rows, world, r) for r in range(world)]
    rank_draws = []
    for r in range(world):
        init_state("resnet18", 0)
        if r == 0:
            draw_augmentation(rows)
        rank_draws.append(draw_augmentation(len(rank_rows[r])))
    return ResNetOracle("resnet18", ds.images, ds.labels, rank_rows, batch_size=batch, lr=0.01, rank_draws=rank_draws)


def _check_checkpoints(tmp, oracle, steps, out):
    import torchvision

    saved = torch.load(os.path.join(tmp, "resnet_distributed.pth"))
    torch.manual_seed(0)
    ref = torchvision.models.resnet18()
    # the epoch-0 checkpoint is written BEFORE training (task.py:279-283): the seeded reference model, bit for bit,
    # and loadable by the reference's own class
    ref2 = torchvision.models.resnet18()
    ref2.load_state_dict({k[len("module."):]: v for k, v in saved.items()}, strict=True)
    for k, v in ref.state_dict().items():
        assert torch.equal(saved["module." + k], v), k
    for _ in range(steps):
        lo = oracle.step()[0]
    lg = float(re.search(r"final local loss \(rank 0\): ([0-9.eE+-]+)", out).group(1))
    print(f"final loss: oracle {lo} task.py {lg}")
    # step 2's loss: forward of the once-updated weights.  1e-3, not 1e-6: a batch may hold a pre-activation within
    # rounding of zero, and one flipped ReLU mask element is ~1 % of a layer4 gradient (profiles/
    # r02_resnet_parity_conditioning.md); measured 4e-5 .. 4e-4 on such batches, 1e-6 otherwise
    assert abs(lo - lg) <= 1e-3 * max(1.0, abs(lo))
    final = torch.load(os.path.join(tmp, "resnet_distributed.pth.final"))
    so = oracle.state_dict()
    assert list(final.keys()) == list(so.keys())
    keys = [k for k in so if so[k].dtype.is_floating_point and "running" not in k]
    uo = torch.cat([(so[k] - saved[k]).double().flatten() for k in keys])
    ug = torch.cat([(final[k] - saved[k]).double().flatten() for k in keys])
    cos = torch.nn.functional.cosine_similarity(uo, ug, dim=0).item()
    print(f"cosine between the two-step updates: {cos:.5f}")
    assert cos >= 0.97
    for k in so:
        if k.endswith("num_batches_tracked"):
            assert int(final[k]) == int(so[k]) == steps
        elif "running" in k:       # forward-side buffers of step 2 see step 1's update (cosine above): measured 3e-3 on
            rel = ((final[k] - so[k]).norm() / so[k].norm()).item()   # bn1.running_mean, whose entries are near zero
            assert rel <= 1e-2, (k, rel)


def test_task_py_default_arch_single_gpu(tmp_path):
    rows, batch = 128, 64
    r = _launch(1, tmp_path, rows, batch, 29571)
    assert r.returncode == 0, r.stderr[-3000:]
    out = r.stdout
    assert "=> creating model 'resnet18'" in out and "GPU x WORLD SIZE = 1" in out
    assert out.count("training with gpu 0") == 2
    assert re.search(r"Epoch: 0, Accuracy: 0\.\d+, Time: ", out) and "saving model to local folders" in out
    _check_checkpoints(tmp_path, _oracle(rows, batch, 1), 2, out)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_task_py_default_arch_two_gpus_matches_ddp_oracle(tmp_path):
    rows, batch = 256, 64
    r = _launch(2, tmp_path, rows, batch, 29572)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "GPU x WORLD SIZE = 2" in r.stdout
    _check_checkpoints(tmp_path, _oracle(rows, batch, 2), 2, r.stdout)
