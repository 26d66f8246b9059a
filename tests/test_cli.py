"""Host-side drop-in checks (CPU): the CLI of this repo's task.py against the reference's parser, and the
launcher / rank arithmetic / sampler partition over gloo with world_size 2 (no GPU, --plan_only)."""
import json
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "reference_cli.json")


def _mine():
    from b200step.cli import build_parser
    from tests.golden.make_cli_golden import describe

    env = {k: os.environ.pop(k, None) for k in ("AIP_MODEL_DIR", "WORLD_SIZE")}
    try:
        parser = build_parser()
    finally:
        for k, v in env.items():
            if v is not None:
                os.environ[k] = v
    return parser, describe


def test_reference_flags_kept_verbatim():
    """All 20 flags of task.py:56-94: same option strings, dest, type, default, action, in the same order."""
    ref = json.load(open(GOLDEN))
    parser, describe = _mine()
    mine = describe(parser, only={r["dest"] for r in ref})
    assert len(ref) == 20 and [m["dest"] for m in mine] == [r["dest"] for r in ref]
    for r, m in zip(ref, mine):
        assert r == m, (r, m)


@pytest.mark.skipif(not os.path.exists("/root/reference/task.py"), reason="reference tree only in the build container")
def test_golden_cli_matches_the_live_reference():
    from tests.golden.make_cli_golden import reference_actions

    assert reference_actions() == json.load(open(GOLDEN))


def test_additive_flags_default_to_reference_behaviour():
    parser, _ = _mine()
    a = parser.parse_args([])
    assert a.arch == "resnet18" and a.optimizer == "sgd" and a.dropout == 0.0 and a.synthetic_rows == 10000
    assert a.learning_rate == 0.1 and a.momentum == 0.9 and a.weight_decay == 1e-4 and a.batch_size == 1024
    assert a.precision == "fp32"            # the reference's nn.Linear arithmetic (task.py:309); bf16 / tf32 are opt-in
    a = parser.parse_args(["--dist-url=env://", "--multiprocessing-distributed", "--num_epochs=2"])  # notebook ARGS
    assert a.dist_url == "env://" and a.multiprocessing_distributed and a.num_epochs == 2
    assert parser.parse_args(["--precision", "tf32"]).precision == "tf32"
    a = parser.parse_args(["--arch", "tabular_mlp_wide", "--weight-decay", "0.01", "-a", "tabular_mlp"])
    assert a.arch == "tabular_mlp" and a.weight_decay == 0.01


def test_unbuilt_arch_fails_loudly():
    from b200step.cli import build_parser, spec_from_args

    a = build_parser().parse_args(["--arch", "vgg16"])         # a torchvision name without a B200 path
    with pytest.raises(NotImplementedError, match="only the BasicBlock ResNets"):
        spec_from_args(a)
    a = build_parser().parse_args(["--arch", "tabular_mlp", "--hidden", "64,32", "--cardinality", "50"])
    s = spec_from_args(a)
    assert s.hidden == (64, 32) and s.cardinality == 50 and s.num_classes == 1000


def _run_task(extra, env_extra, timeout=240):
    env = dict(os.environ)
    env.update({"MASTER_ADDR": "127.0.0.1", "CUDA_VISIBLE_DEVICES": ""})
    env.update(env_extra)
    cmd = [sys.executable, os.path.join(ROOT, "task.py")] + extra
    return subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=timeout, cwd=ROOT)


def test_launcher_plan_world2_gloo():
    """mp.spawn with 2 processes on one node over gloo: global rank = node_rank*ngpus + gpu (task.py:146), world =
    ngpus * nodes (task.py:120), each rank takes perm[r::W] (task.py:260) — the notebook's three-flag launch."""
    r = _run_task(["--dist-url", "env://", "--multiprocessing-distributed", "--num_epochs", "2", "--dist-backend",
                   "gloo", "--arch", "tabular_mlp", "--plan_only", "--synthetic_rows", "1000", "--batch_size", "128"],
                  {"WORLD_SIZE": "1", "RANK": "0", "MASTER_PORT": "29551", "B200STEP_PLAN_GPUS": "2"})
    assert r.returncode == 0, r.stderr[-2000:]
    out = r.stdout
    for line in ["os WORLD_SIZE=1", "os RANK=0", "Arg - distributed=True", "Arg - multiprocessing_distributed=True",
                 "Arg - dist_backend=gloo", "Arg - dist_url=env://", "ngpus_per_node=2", "GPU x WORLD SIZE = 2",
                 "Process group initialized", "random seeds", "=> creating model 'tabular_mlp'"]:
        assert line in out, line
    assert re.search(r"Starting training: \d{4}_\d\d_\d\d_\d\d_\d\d_\d\d", out)
    assert re.search(r"Training complete: \d{4}_", out)
    plans = sorted((json.loads(m) for m in re.findall(r"PLAN (\{.*\})", out)), key=lambda p: p["rank"])
    assert [p["rank"] for p in plans] == [0, 1] and all(p["world"] == 2 for p in plans)
    from b200step.sampler import rank_indices

    for p in plans:
        assert p["samples_this_rank"] == 500 and p["batches_per_epoch"] == 4 and p["last_batch_rows"] == 500 - 3 * 128
        assert p["first_indices"] == rank_indices(1000, 2, p["rank"])[:4].tolist()
        assert p["params"] == 1411816


def test_node_rank_arithmetic_second_node():
    """RANK env is the NODE rank: node 1 with 2 GPUs hosts global ranks 2 and 3 of world 4 — checked without
    a rendezvous by calling the same arithmetic main_worker uses."""
    r = _run_task(["--dist-url", "env://", "--num_epochs", "1", "--dist-backend", "gloo", "--arch", "tabular_mlp",
                   "--plan_only", "--synthetic_rows", "64", "--batch_size", "16"],
                  {"WORLD_SIZE": "1", "RANK": "0", "MASTER_PORT": "29552"})
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Arg - distributed=False" in r.stdout  # world_size 1, no --multiprocessing-distributed (task.py:100)
    plan = json.loads(re.search(r"PLAN (\{.*\})", r.stdout).group(1))
    assert plan["world"] == 1 and plan["samples_this_rank"] == 64 and plan["batches_per_epoch"] == 4


def test_training_without_gpu_exits_nonzero():
    """Fail-fast contract: the pipeline only reads the exit code."""
    r = _run_task(["--arch", "tabular_mlp", "--num_epochs", "1", "--local_training", "--synthetic_rows", "64",
                   "--batch_size", "32", "--gpu", "0"], {"MASTER_PORT": "29553"})
    assert r.returncode != 0
    assert "CUDA device" in r.stderr or "CUDA" in r.stderr


def test_default_arch_resnet18_plans_the_reference_model():
    """`python task.py` with no --arch is the reference's literal configuration: resnet18, 11 689 512 parameters, the
    10 000-image test split over the ranks, batches of 1024 (task.py:58,63,256,263)."""
    r = _run_task(["--plan_only", "--local_training"], {"WORLD_SIZE": "-1"})
    assert r.returncode == 0, r.stdout + r.stderr
    plan = json.loads([l for l in r.stdout.splitlines() if l.startswith("PLAN ")][0][5:])
    assert plan["arch"] == "resnet18" and plan["params"] == 11689512
    assert plan["samples_this_rank"] == 10000 and plan["batches_per_epoch"] == 10 and plan["last_batch_rows"] == 784


def test_launcher_plan_world2_gloo_default_arch():
    """The same two-process gloo launch with the reference's default model: resnet18 over the 10 000-image split, 5 000
    images per rank in DistributedSampler order (task.py:63,256,260)."""
    r = _run_task(["--dist-url", "env://", "--multiprocessing-distributed", "--dist-backend", "gloo", "--plan_only"],
                  {"WORLD_SIZE": "1", "RANK": "0", "MASTER_PORT": "29553", "B200STEP_PLAN_GPUS": "2"})
    assert r.returncode == 0, r.stderr[-2000:]
    assert "=> creating model 'resnet18'" in r.stdout and "GPU x WORLD SIZE = 2" in r.stdout
    plans = sorted((json.loads(m) for m in re.findall(r"PLAN (\{.*\})", r.stdout)), key=lambda p: p["rank"])
    assert [p["rank"] for p in plans] == [0, 1]
    from b200step.sampler import rank_indices

    for p in plans:
        assert p["arch"] == "resnet18" and p["params"] == 11689512 and p["world"] == 2
        assert p["samples_this_rank"] == 5000 and p["batches_per_epoch"] == 5 and p["last_batch_rows"] == 5000 - 4 * 1024
        assert p["first_indices"] == rank_indices(10000, 2, p["rank"])[:4].tolist()
